// brick.cuh -- brick index: the spatially coherent read view of the slot table.
//
// A brick is 4 x 4 x 4 voxels.  Every point that is FINDABLE through the reference's direct-mapped table
// (table[slot(cell(p))] == p, model/neural_points.py:216-247) sets one bit of its brick's 64-bit
// occupancy mask; the records of a brick are contiguous and ordered by bit.  The 81-cell stencil of a
// query touches exactly 2 x 2 x 2 bricks, so a neighbour search costs 8 probes of an open-addressing
// table (linear probing, 32-byte entries) instead of 81 probes of the slot view, and queries of the same
// brick read the same few cache lines.
//
// Equivalence with the slot table: lookup(cell) through the brick index returns the point p whose own
// cell is `cell` and which owns its slot.  The reference additionally sees far-away points that collide
// onto the probed slot -- and always rejects them by the distance test (|p - x|^2 <= max_valid_dist2)
// PROVIDED no two cells within the collision radius share a slot; the host verifies that property of
// (primes, buffer_size) before enabling this path and otherwise keeps the slot-probing kernels
// (model/neural_points.py of this package, _brick_safe()).
#pragma once
#include "query_common.cuh"

namespace pinq {

struct __align__(32) BrickEntry {
    unsigned long long mask;  // bit ((lx * 4 + ly) * 4 + lz) <-> voxel (lx, ly, lz) of the brick holds a findable point
    uint32_t base;            // rank of the brick's first record
    uint32_t pad0;
    unsigned long long key;   // packed brick coordinates, kBrickEmpty = unused entry
    unsigned long long pad1;
};
static_assert(sizeof(BrickEntry) == 32, "one sector per probe");

constexpr unsigned long long kBrickEmpty = ~0ull;

__device__ __forceinline__ unsigned long long brick_key(int bx, int by, int bz) {
    // 21 bits per axis, offset binary: |brick| < 2^20 <-> |cell| < 2^22
    return ((unsigned long long)(uint32_t)(bx + (1 << 20)) << 42) | ((unsigned long long)(uint32_t)(by + (1 << 20)) << 21) |
           (unsigned long long)(uint32_t)(bz + (1 << 20));
}
// 32-bit mix of the two key words (multiply-add, two xor-shift-multiply rounds of the murmur3 finaliser): 8 instructions;
// the 64-bit finaliser it replaces cost ~20 (two 64 x 64 multiplies) eight times per query in k_search_brick
__device__ __forceinline__ uint32_t brick_hash(unsigned long long k, uint32_t cap_mask) {
    uint32_t h = (uint32_t)k * 0x9E3779B1u + (uint32_t)(k >> 32) * 0x85EBCA77u;
    h ^= h >> 16;
    h *= 0x85EBCA6Bu;
    h ^= h >> 13;
    h *= 0xC2B2AE35u;
    h ^= h >> 16;
    return h & cap_mask;
}
__device__ __forceinline__ int brick_bit(int cx, int cy, int cz) { return (((cx & 3) * 4) + (cy & 3)) * 4 + (cz & 3); }

// one whole 32-byte entry with a single 256-bit load: words {mask lo, mask hi, base, pad, key lo, key hi, pad, pad}
__device__ __forceinline__ void ldg_entry(const BrickEntry* e, uint32_t (&w)[8]) {
    asm("ld.global.nc.v8.u32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
        : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "l"(e));
}

// read-only lookup: entry index of `key`, or -1
__device__ __forceinline__ int brick_find(const BrickEntry* tab, uint32_t cap_mask, unsigned long long key) {
    uint32_t h = brick_hash(key, cap_mask);
    for (;;) {
        const unsigned long long k = __ldg(&tab[h].key);
        if (k == key) return (int)h;
        if (k == kBrickEmpty) return -1;
        h = (h + 1) & cap_mask;
    }
}

struct BrickDev {
    const BrickEntry* tab;
    uint32_t cap_mask;
    const float4* rec;  // brick-ordered candidate records {x, y, z, bits(index)}
    int radius;         // stencil radius in voxels (<= 2)
    bool exact_keys;    // PINQ_F_EXACT_KEYS
};

}  // namespace pinq

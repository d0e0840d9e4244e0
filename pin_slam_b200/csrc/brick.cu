// brick.cu -- builds the brick index (csrc/brick.cuh) from the rank-ordered records of the occupancy index.
// Three passes over the occupied slots (not over the table):
//   k_brick_insert  claim the brick's entry (atomicCAS on the key), set the voxel's bit (atomicOr)
//   k_brick_base    the record owning the lowest set bit of a brick reserves the brick's record range
//   k_brick_pack    rec_brick[base + popc(mask below bit)] = record
// The order of bricks in the record array is whatever the atomics produce; nothing observable depends on it.
#include "brick.cuh"

namespace pinq {

__global__ void k_brick_clear(BrickEntry* tab, uint32_t cap, int32_t* counters) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < cap) {
        BrickEntry e;
        e.mask = 0ull; e.base = 0u; e.pad0 = 0u; e.key = kBrickEmpty; e.pad1 = 0ull;
        tab[i] = e;
    }
    if (i < 2) counters[i] = 0;
}

__device__ __forceinline__ void record_cell(const float4 r, float res, float inv_res, bool recip, int& cx, int& cy, int& cz) {
    cx = cell_of(r.x, res, inv_res, recip);
    cy = cell_of(r.y, res, inv_res, recip);
    cz = cell_of(r.z, res, inv_res, recip);
}

__global__ void k_brick_insert(const float4* __restrict__ rec, const int32_t* __restrict__ d_count, float res, float inv_res,
                               bool recip, BrickEntry* tab, uint32_t cap_mask, int32_t* counters) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= *d_count) return;
    int cx, cy, cz;
    record_cell(rec[r], res, inv_res, recip, cx, cy, cz);
    const unsigned long long key = brick_key(cx >> 2, cy >> 2, cz >> 2);
    const unsigned long long bit = 1ull << brick_bit(cx, cy, cz);
    uint32_t h = brick_hash(key, cap_mask);
    for (;;) {
        const unsigned long long prev = atomicCAS(&tab[h].key, kBrickEmpty, key);
        if (prev == kBrickEmpty || prev == key) {
            const unsigned long long old = atomicOr(&tab[h].mask, bit);
            if (old & bit) atomicAdd(&counters[1], 1);  // two findable points in one voxel: the table invariant is broken
            return;
        }
        h = (h + 1) & cap_mask;
    }
}

__global__ void k_brick_base(const float4* __restrict__ rec, const int32_t* __restrict__ d_count, float res, float inv_res,
                             bool recip, BrickEntry* tab, uint32_t cap_mask, int32_t* counters) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= *d_count) return;
    int cx, cy, cz;
    record_cell(rec[r], res, inv_res, recip, cx, cy, cz);
    const int e = brick_find(tab, cap_mask, brick_key(cx >> 2, cy >> 2, cz >> 2));
    if (e < 0) return;
    const unsigned long long mask = tab[e].mask;
    if (__ffsll((long long)mask) - 1 == brick_bit(cx, cy, cz)) tab[e].base = (uint32_t)atomicAdd(&counters[0], __popcll(mask));
}

__global__ void k_brick_pack(const float4* __restrict__ rec_g, const float4* __restrict__ rec_l,
                             const int32_t* __restrict__ d_count, float res, float inv_res, bool recip,
                             const BrickEntry* __restrict__ tab, uint32_t cap_mask, float4* __restrict__ out_g,
                             float4* __restrict__ out_l) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= *d_count) return;
    const float4 v = rec_g[r];
    int cx, cy, cz;
    record_cell(v, res, inv_res, recip, cx, cy, cz);
    const int e = brick_find(tab, cap_mask, brick_key(cx >> 2, cy >> 2, cz >> 2));
    if (e < 0) return;
    const unsigned long long mask = tab[e].mask;
    const uint32_t rank = tab[e].base + (uint32_t)__popcll(mask & ((1ull << brick_bit(cx, cy, cz)) - 1ull));
    if (out_g) out_g[rank] = v;
    if (out_l) out_l[rank] = rec_l[r];
}

}  // namespace pinq

using namespace pinq;

extern "C" {

int64_t pinq_brick_capacity(int64_t m) {
    int64_t cap = 1024;
    while (cap < 2 * m) cap <<= 1;  // bricks <= occupied slots <= m: load factor <= 0.5
    return cap;
}

int pinq_brick_build(const float* rec_global4, const float* rec_local4, const int32_t* d_count, int64_t m, float res,
                     uint32_t flags, void* tab, uint32_t cap, int32_t* d_counters2, float* brick_rec_global4,
                     float* brick_rec_local4, void* stream) {
    PINQ_REQUIRE(rec_global4 && d_count && tab && d_counters2, "null pointer");
    PINQ_REQUIRE(cap >= 1024 && (cap & (cap - 1)) == 0 && (int64_t)cap >= 2 * m, "cap must be a power of two >= 2 m");
    PINQ_REQUIRE(!brick_rec_local4 || rec_local4, "local brick records need the local slot records");
    cudaStream_t st = as_stream(stream);
    BrickEntry* t = reinterpret_cast<BrickEntry*>(tab);
    const bool recip = (flags & PINQ_F_RECIP_DIV) != 0;
    const float inv = 1.0f / res;
    k_brick_clear<<<grid_for(cap, 256), 256, 0, st>>>(t, cap, d_counters2);
    if (m > 0) {
        const float4* rg = reinterpret_cast<const float4*>(rec_global4);
        const unsigned g = grid_for(m, 256);
        k_brick_insert<<<g, 256, 0, st>>>(rg, d_count, res, inv, recip, t, cap - 1, d_counters2);
        k_brick_base<<<g, 256, 0, st>>>(rg, d_count, res, inv, recip, t, cap - 1, d_counters2);
        k_brick_pack<<<g, 256, 0, st>>>(rg, reinterpret_cast<const float4*>(rec_local4), d_count, res, inv, recip, t, cap - 1,
                                        reinterpret_cast<float4*>(brick_rec_global4),
                                        reinterpret_cast<float4*>(brick_rec_local4));
    }
    return check_launch("pinq_brick_build");
}

}  // extern "C"

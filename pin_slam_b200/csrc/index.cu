// index.cu -- L2-resident read view of the slot table ("occupancy index").
//
// The reference's direct-mapped table (model/neural_points.py:73) is 5e7 slots of which a few per cent
// are occupied; probing it directly costs one DRAM sector per probe (ncu, profiles/r1a: 8.1 KB of DRAM
// reads per query, L2 hit rate 19 %).  The query kernels therefore read a succinct, bit-identical view:
//   occ[w] = { bits : occupancy of slots 32w..32w+31,  prefix : number of occupied slots below 32w }
//   rec[r] = packed candidate of the r-th occupied slot (xyz + filtered local index | global index)
// lookup(slot) == table[slot] by construction: empty iff the bit is clear, else the record at rank
// prefix + popc(bits below).  12.5 MB + 16 B per neural point instead of 200 MB -> stays in the 126 MB L2.
// The dense int32 table remains the master copy for insertion (update / recreate_hash); the view is
// rebuilt from it with two streaming passes (~2 x 31 us at 6.5 TB/s) whenever the map changes.
#include "common.cuh"

namespace pinq {

constexpr int kScanTileW = 1024;

// one warp per 1024 consecutive slots: 32 coalesced loads per lane, ballot -> 32 occupancy words
__global__ void k_occ_bits(const int32_t* __restrict__ table, uint32_t B, uint32_t nwords, uint2* __restrict__ occ) {
    const int lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t nchunks = (nwords + 31) / 32;
    for (uint32_t chunk = warp; chunk < nchunks; chunk += nwarps) {
        const uint64_t base = (uint64_t)chunk * 1024u;
        int32_t v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            const uint64_t s = base + (uint64_t)i * 32u + lane;
            v[i] = s < B ? __ldcs(&table[s]) : -1;
        }
        uint32_t mine = 0;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            const uint32_t b = __ballot_sync(0xffffffffu, v[i] >= 0);
            if (lane == i) mine = b;
        }
        const uint32_t w = chunk * 32u + lane;
        if (w < nwords) occ[w].x = mine;
    }
}

__global__ void k_occ_tile_count(const uint2* __restrict__ occ, uint32_t nwords, int32_t* __restrict__ tile_sums) {
    __shared__ int wsum[32];
    const uint32_t i = blockIdx.x * kScanTileW + threadIdx.x;
    int c = i < nwords ? __popc(occ[i].x) : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x < 32) {
        int v = wsum[threadIdx.x];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (threadIdx.x == 0) tile_sums[blockIdx.x] = v;
    }
}

__global__ void k_occ_prefix(uint2* __restrict__ occ, uint32_t nwords, const int32_t* __restrict__ tile_offs) {
    __shared__ int wsum[32];
    const uint32_t i = blockIdx.x * kScanTileW + threadIdx.x;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int c = i < nwords ? __popc(occ[i].x) : 0;
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) wsum[w] = incl;
    __syncthreads();
    if (threadIdx.x < 32) {
        const int v = wsum[threadIdx.x];
        int wi = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, wi, o);
            if (threadIdx.x >= o) wi += t;
        }
        wsum[threadIdx.x] = wi - v;
    }
    __syncthreads();
    if (i < nwords) occ[i].y = (uint32_t)(tile_offs[blockIdx.x] + wsum[w] + incl - c);
}

// second streaming pass: rec[rank(slot)] = { xyz[v], bits(index) } for every occupied slot (v = table[slot]).
// index = v (global view) or, for the local view, global2local[v] after the travel-distance filter of
// model/neural_points.py:480-488 (-1 = not a candidate for local queries).
__global__ void k_occ_pack(const int32_t* __restrict__ table, uint32_t B, uint32_t nwords, const uint2* __restrict__ occ,
                           const float* __restrict__ xyz, const int64_t* __restrict__ ts_c, const int64_t* __restrict__ g2l,
                           const uint8_t* __restrict__ lmask, const float* __restrict__ td, int64_t cur_ts, float diff_td, float4* __restrict__ rec_global,
                           float4* __restrict__ rec_local) {
    const int lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t nchunks = (nwords + 31) / 32;
    for (uint32_t chunk = warp; chunk < nchunks; chunk += nwarps) {
        const uint32_t w = chunk * 32u + lane;
        const uint2 mine = w < nwords ? occ[w] : make_uint2(0u, 0u);
        if (__ballot_sync(0xffffffffu, mine.x != 0u) == 0u) continue;  // 1024 empty slots: nothing to emit
        const uint64_t base = (uint64_t)chunk * 1024u;
#pragma unroll 4
        for (int i = 0; i < 32; ++i) {
            const uint32_t bits = __shfl_sync(0xffffffffu, mine.x, i);
            const uint32_t pre = __shfl_sync(0xffffffffu, mine.y, i);
            if (!((bits >> lane) & 1u)) continue;
            const uint64_t s = base + (uint64_t)i * 32u + lane;
            const int32_t v = table[s];
            const uint32_t r = pre + __popc(bits & ((1u << lane) - 1u));
            const float x = xyz[3 * (int64_t)v], y = xyz[3 * (int64_t)v + 1], z = xyz[3 * (int64_t)v + 2];
            if (rec_global) rec_global[r] = make_float4(x, y, z, __int_as_float(v));
            if (rec_local) {
                int32_t li = lmask[v] ? (int32_t)g2l[v] : -2;  // -2: outside the local map -> index 1 (hazard H13)
                const float dt = fabsf(__fsub_rn(td[cur_ts], td[ts_c[v]]));
                if (!(dt < diff_td)) li = -1;
                rec_local[r] = make_float4(x, y, z, __int_as_float(li));
            }
        }
    }
}

}  // namespace pinq

using namespace pinq;

extern "C" {

int64_t pinq_index_words(uint32_t buffer_size) { return ((int64_t)buffer_size + 31) / 32; }

int64_t pinq_index_scratch_bytes(uint32_t buffer_size) {
    const int64_t nwords = pinq_index_words(buffer_size);
    return ((nwords + kScanTileW - 1) / kScanTileW + 1) * (int64_t)sizeof(int32_t);
}

int pinq_index_build(const int32_t* table, uint32_t buffer_size, uint32_t* occ2, int32_t* d_count, void* scratch,
                     void* stream) {
    PINQ_REQUIRE(table && occ2 && d_count && scratch && buffer_size > 0, "null pointer");
    cudaStream_t st = as_stream(stream);
    const uint32_t nwords = (uint32_t)pinq_index_words(buffer_size);
    uint2* occ = reinterpret_cast<uint2*>(occ2);
    k_occ_bits<<<148 * 8, 256, 0, st>>>(table, buffer_size, nwords, occ);
    const int64_t ntiles = (nwords + kScanTileW - 1) / kScanTileW;
    int32_t* tiles = reinterpret_cast<int32_t*>(scratch);
    k_occ_tile_count<<<(unsigned)ntiles, kScanTileW, 0, st>>>(occ, nwords, tiles);
    launch_tile_scan(tiles, ntiles, d_count, st);
    k_occ_prefix<<<(unsigned)ntiles, kScanTileW, 0, st>>>(occ, nwords, tiles);
    return check_launch("pinq_index_build");
}

int pinq_index_pack(const int32_t* table, uint32_t buffer_size, const uint32_t* occ2, const float* map_xyz,
                    const int64_t* ts_create, const int64_t* global2local, const uint8_t* local_mask,
                    const float* travel_dist, int64_t cur_ts, float diff_travel_dist_local, float* rec_global4, float* rec_local4, void* stream) {
    PINQ_REQUIRE(table && occ2 && map_xyz, "null pointer");
    PINQ_REQUIRE(!rec_local4 || (ts_create && global2local && travel_dist && local_mask), "local records need ts_create, global2local, travel_dist");
    if (!rec_global4 && !rec_local4) return PINQ_OK;
    const uint32_t nwords = (uint32_t)pinq_index_words(buffer_size);
    k_occ_pack<<<148 * 8, 256, 0, as_stream(stream)>>>(table, buffer_size, nwords, reinterpret_cast<const uint2*>(occ2),
                                                       map_xyz, ts_create, global2local, local_mask, travel_dist, cur_ts,
                                                       diff_travel_dist_local, reinterpret_cast<float4*>(rec_global4),
                                                       reinterpret_cast<float4*>(rec_local4));
    return check_launch("pinq_index_pack");
}

}  // extern "C"

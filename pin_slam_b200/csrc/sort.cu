// sort.cu -- spatial binning of a query batch (one-pass counting sort on a 16-bit brick key).
//
// ncu on the unsorted batch (profiles/r1d_*): the decode and backward kernels wait on gathers that miss
// L1 and half of L2 (1.5-2.2 GB of DRAM reads per 4 Mi queries for 53 MB of features), and the search is
// bound by L1TEX wavefronts of fully divergent probes.  Processing the queries brick by brick (a brick =
// 4 x 4 x 4 voxels) turns those gathers into L1 hits.  Only the PROCESSING ORDER changes: every kernel
// reads its inputs and writes its per-query outputs through perm[], so results are those of the unsorted
// batch.  The order inside a bin is whatever the atomics produce -- no output depends on it.
//   k_bin_count   key[i] = brick key of query i, hist[key]++
//   k_bin_scan    exclusive scan of the 65536 counters (one CTA)
//   k_bin_scatter perm[cursor[key[i]]++] = i
#include "query_common.cuh"

namespace pinq {

constexpr int kBinBits = 16;
constexpr int kBins = 1 << kBinBits;

__device__ __forceinline__ uint32_t brick_key16(int gx, int gy, int gz) {
    // 6 + 6 + 4 bits of the brick coordinates (cell >> 2): bins alias with a period of 102 m / 102 m / 26 m,
    // which only interleaves far-apart bricks in the processing order
    return (uint32_t)((gx >> 2) & 63) | ((uint32_t)((gy >> 2) & 63) << 6) | ((uint32_t)((gz >> 2) & 15) << 12);
}

__global__ void k_bin_count(const float* __restrict__ xyz, int64_t n, float res, float inv_res, bool recip,
                            uint16_t* __restrict__ keys, uint32_t* __restrict__ hist) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t key = brick_key16(cell_of(xyz[3 * i], res, inv_res, recip), cell_of(xyz[3 * i + 1], res, inv_res, recip),
                                     cell_of(xyz[3 * i + 2], res, inv_res, recip));
    keys[i] = (uint16_t)key;
    atomicAdd(&hist[key], 1u);
}

__global__ void __launch_bounds__(1024) k_bin_scan(uint32_t* __restrict__ hist) {
    __shared__ uint32_t wsum[32];
    constexpr int per = kBins / 1024;
    uint32_t v[per], s = 0;
#pragma unroll
    for (int j = 0; j < per; ++j) {
        v[j] = hist[threadIdx.x * per + j];
        s += v[j];
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t incl = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) wsum[w] = incl;
    __syncthreads();
    if (threadIdx.x < 32) {
        const uint32_t x = wsum[threadIdx.x];
        uint32_t wi = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
            if (threadIdx.x >= o) wi += t;
        }
        wsum[threadIdx.x] = wi - x;
    }
    __syncthreads();
    uint32_t run = wsum[w] + incl - s;
#pragma unroll
    for (int j = 0; j < per; ++j) {
        hist[threadIdx.x * per + j] = run;
        run += v[j];
    }
}

__global__ void k_bin_scatter(const uint16_t* __restrict__ keys, int64_t n, uint32_t* __restrict__ cursor,
                              int32_t* __restrict__ perm) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    perm[atomicAdd(&cursor[keys[i]], 1u)] = (int32_t)i;
}

int64_t bin_scratch_bytes(int64_t n) { return (int64_t)kBins * 4 + ((n * 2 + 15) / 16) * 16; }

int launch_spatial_bin(const float* xyz, int64_t n, float res, bool recip, void* scratch, int32_t* perm, cudaStream_t st) {
    uint32_t* hist = reinterpret_cast<uint32_t*>(scratch);
    uint16_t* keys = reinterpret_cast<uint16_t*>(hist + kBins);
    cudaMemsetAsync(hist, 0, (size_t)kBins * 4, st);
    k_bin_count<<<grid_for(n, 256), 256, 0, st>>>(xyz, n, res, 1.0f / res, recip, keys, hist);
    k_bin_scan<<<1, 1024, 0, st>>>(hist);
    k_bin_scatter<<<grid_for(n, 256), 256, 0, st>>>(keys, n, hist, perm);
    return check_launch("spatial binning");
}

}  // namespace pinq

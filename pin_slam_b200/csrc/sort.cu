// sort.cu -- spatial binning of a query batch (one-pass counting sort on a 21-bit brick key).
//
// ncu on the unsorted batch (profiles/r1d_*): the decode and backward kernels wait on gathers that miss
// L1 and half of L2 (1.5-2.2 GB of DRAM reads per 4 Mi queries for 53 MB of features), and the search is
// bound by L1TEX wavefronts of fully divergent probes.  Processing the queries brick by brick (a brick =
// 4 x 4 x 4 voxels) turns those gathers into L1 hits.  Only the PROCESSING ORDER changes: every kernel
// reads its inputs and writes its per-query outputs through perm[], so results are those of the unsorted
// batch.  The order inside a bin is whatever the atomics produce -- no output depends on it.
//   k_bin_count   key[i] = brick key of query i, arrival[i] = hist[key]++  (the atomic's return value)
//   k_bin_tile_*  exclusive scan of the 2^21 counters (tile sums, scan of the sums, per-tile prefix)
//   k_bin_scatter perm[base[key[i]] + arrival[i]] = i   (no second round of atomics)
#include "query_common.cuh"

namespace pinq {

constexpr int kBinBits = 21;
constexpr int kBins = 1 << kBinBits;
constexpr int kScanTile = 1024;
constexpr int kScanTiles = kBins / kScanTile;  // 2048

__device__ __forceinline__ uint32_t brick_key21(int gx, int gy, int gz) {
    // 8 + 8 + 5 bits of the brick coordinates (cell >> 2), x fastest: one bin = one brick up to a period of
    // 410 m / 410 m / 51 m (aliases only interleave far-apart bricks in the processing order)
    return (uint32_t)((gx >> 2) & 255) | ((uint32_t)((gy >> 2) & 255) << 8) | ((uint32_t)((gz >> 2) & 31) << 16);
}

__global__ void k_bin_count(const float* __restrict__ xyz, int64_t n, float res, float inv_res, bool recip,
                            uint32_t* __restrict__ keys, uint32_t* __restrict__ arrival, uint32_t* __restrict__ hist) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t key = brick_key21(cell_of(xyz[3 * i], res, inv_res, recip), cell_of(xyz[3 * i + 1], res, inv_res, recip),
                                     cell_of(xyz[3 * i + 2], res, inv_res, recip));
    keys[i] = key;
    arrival[i] = atomicAdd(&hist[key], 1u);
}

// block-wide exclusive scan of one value per thread (1024 threads); returns the exclusive prefix, total in *tot
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* wsum, uint32_t* tot) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) wsum[w] = incl;
    __syncthreads();
    if (w == 0) {
        const uint32_t x = wsum[lane];
        uint32_t wi = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += t;
        }
        wsum[lane] = wi - x;
        if (lane == 31 && tot) *tot = wi;
    }
    __syncthreads();
    return wsum[w] + incl - v;
}

// The scan kernels work on tiles of 1024 counters with 256 threads x 4 counters (uint4): 8 resident CTAs per SM cover the
// 2048 tiles in ONE wave and hide each other's load latency (1024-thread CTAs needed 7 waves of 2 per SM: 12 us per pass).
constexpr int kScanThreads = kScanTile / 4;

// exclusive scan of one value per thread over a 256-thread CTA; returns the exclusive prefix, the total through *tot
__device__ __forceinline__ uint32_t block_excl_scan256(uint32_t v, uint32_t* wsum, uint32_t* tot) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) wsum[w] = incl;
    __syncthreads();
    uint32_t before = 0u, all = 0u;
#pragma unroll
    for (int i = 0; i < kScanThreads / 32; ++i) {
        const uint32_t x = wsum[i];
        before += i < w ? x : 0u;
        all += x;
    }
    if (tot) *tot = all;
    return before + incl - v;
}

__global__ void __launch_bounds__(kScanThreads) k_bin_tile_sums(const uint4* __restrict__ hist4, uint32_t* __restrict__ tiles) {
    __shared__ uint32_t wsum[kScanThreads / 32];
    const uint4 v = hist4[blockIdx.x * kScanThreads + threadIdx.x];
    uint32_t tot;
    block_excl_scan256(v.x + v.y + v.z + v.w, wsum, &tot);
    if (threadIdx.x == 0) tiles[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(kScanTile) k_bin_tile_scan(uint32_t* __restrict__ tiles) {  // one CTA, 2048 tile sums
    __shared__ uint32_t wsum[32];
    __shared__ uint32_t tot;
    const uint32_t a = tiles[2 * threadIdx.x], b = tiles[2 * threadIdx.x + 1];
    const uint32_t ex = block_excl_scan(a + b, wsum, &tot);
    tiles[2 * threadIdx.x] = ex;
    tiles[2 * threadIdx.x + 1] = ex + a;
}

__global__ void __launch_bounds__(kScanThreads) k_bin_tile_apply(uint4* __restrict__ hist4, const uint32_t* __restrict__ tiles) {
    __shared__ uint32_t wsum[kScanThreads / 32];
    const uint32_t idx = blockIdx.x * kScanThreads + threadIdx.x;
    const uint4 v = hist4[idx];
    const uint32_t ex = tiles[blockIdx.x] + block_excl_scan256(v.x + v.y + v.z + v.w, wsum, nullptr);
    hist4[idx] = make_uint4(ex, ex + v.x, ex + v.x + v.y, ex + v.x + v.y + v.z);
}

__global__ void k_bin_scatter(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ arrival, int64_t n,
                              const uint32_t* __restrict__ base, int32_t* __restrict__ perm) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    perm[__ldg(&base[keys[i]]) + arrival[i]] = (int32_t)i;
}

int64_t bin_scratch_bytes(int64_t n) { return (int64_t)(kBins + kScanTiles) * 4 + 2 * (((n * 4 + 15) / 16) * 16); }

int launch_spatial_bin(const float* xyz, int64_t n, float res, bool recip, void* scratch, int32_t* perm, cudaStream_t st) {
    uint32_t* hist = reinterpret_cast<uint32_t*>(scratch);
    uint32_t* tiles = hist + kBins;
    uint32_t* keys = tiles + kScanTiles;
    uint32_t* arrival = keys + ((n + 3) / 4) * 4;
    cudaMemsetAsync(hist, 0, (size_t)kBins * 4, st);
    k_bin_count<<<grid_for(n, 256), 256, 0, st>>>(xyz, n, res, 1.0f / res, recip, keys, arrival, hist);
    k_bin_tile_sums<<<kScanTiles, kScanThreads, 0, st>>>(reinterpret_cast<const uint4*>(hist), tiles);
    k_bin_tile_scan<<<1, kScanTile, 0, st>>>(tiles);
    k_bin_tile_apply<<<kScanTiles, kScanThreads, 0, st>>>(reinterpret_cast<uint4*>(hist), tiles);
    k_bin_scatter<<<grid_for(n, 256), 256, 0, st>>>(keys, arrival, n, hist, perm);
    return check_launch("spatial binning");
}

}  // namespace pinq

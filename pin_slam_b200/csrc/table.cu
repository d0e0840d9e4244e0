// table.cu -- voxel-hash neural-point index: insertion, lookup, local-map window, compaction.
// Replaces the ATen index/index_put_/unique/scatter_reduce chains of
// model/neural_points.py:205-313, 372-428 and utils/tools.py:409-477 (reference paths).
#include <stdarg.h>
#include <string.h>

#include <mutex>

#include "common.cuh"

namespace pinq {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: CUDA error %d (%s)", what, (int)e, cudaGetErrorString(e));
        return PINQ_ERR_CUDA;
    }
    return PINQ_OK;
}

// Optional per-kernel timing of the query path (bench.py roofline): CUDA events on the launching stream between
// the kernels of one pinq_query_fused_fwd / _bwd call.  Off by default; marks cost nothing when off.
static bool g_ktime_on = false;
static cudaEvent_t g_ktime_ev[PINQ_KT_MARKS];
static bool g_ktime_set[PINQ_KT_MARKS];

void ktime_mark(int mark, cudaStream_t st) {
    if (!g_ktime_on || mark < 0 || mark >= PINQ_KT_MARKS) return;
    if (!g_ktime_ev[mark]) cudaEventCreate(&g_ktime_ev[mark]);
    cudaEventRecord(g_ktime_ev[mark], st);
    g_ktime_set[mark] = true;
}

// ---------------------------------------------------------------------------------------------
__global__ void k_fill_i32(int32_t* p, int64_t n, int32_t v) {
    // 16-byte stores: the table is 200 MB at buffer_size = 5e7, so this is a pure HBM-write kernel.
    int64_t i4 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t n4 = n >> 2;
    int4 vv = make_int4(v, v, v, v);
    for (int64_t i = i4; i < n4; i += stride) reinterpret_cast<int4*>(p)[i] = vv;
    for (int64_t i = (n4 << 2) + i4; i < n; i += stride) p[i] = v;
}

__global__ void k_hash_points(const float* __restrict__ xyz, int64_t n, float res, float inv_res, bool recip,
                              uint32_t B, int32_t* __restrict__ slots) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int gx = cell_of(xyz[3 * i + 0], res, inv_res, recip);
    int gy = cell_of(xyz[3 * i + 1], res, inv_res, recip);
    int gz = cell_of(xyz[3 * i + 2], res, inv_res, recip);
    slots[i] = (int32_t)slot_of_cell(gx, gy, gz, B);
}

__global__ void k_gather_i32(const int32_t* __restrict__ table, const int32_t* __restrict__ slots, int64_t n,
                             int32_t* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = table[slots[i]];
}

// Deterministic "sequential index_put_" (hazard H4): phase 1 tags each slot with the highest batch
// position that targets it (encoded as -(2+i), below every legal value), phase 2 lets that position
// store its value.
__global__ void k_scatter_tag(int32_t* table, const int32_t* __restrict__ slots, int64_t n) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) atomicMin(&table[slots[i]], -(int32_t)(2 + i));
}
__global__ void k_scatter_commit(int32_t* table, const int32_t* __restrict__ slots,
                                 const int32_t* __restrict__ values, int64_t n) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int32_t s = slots[i];
    if (table[s] == -(int32_t)(2 + i)) table[s] = values[i];
}

__global__ void k_update_mask(const float* __restrict__ sp, const int32_t* __restrict__ hash_idx, int64_t n,
                              const float* __restrict__ map_xyz, const int64_t* __restrict__ ts_update, int64_t m,
                              const float* __restrict__ travel_dist, int64_t cur_ts, float thr_d2, float diff_td,
                              uint8_t* __restrict__ mask) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (m == 0) {
        mask[i] = 1;
        return;
    }
    int32_t h = hash_idx[i];
    int64_t j = h < 0 ? m - 1 : h;  // -1 is a live (wrapping) index in the reference (hazard H8)
    float d2 = dist2_rn(map_xyz[3 * j], map_xyz[3 * j + 1], map_xyz[3 * j + 2], sp[3 * i], sp[3 * i + 1],
                        sp[3 * i + 2]);
    float dt = __fsub_rn(travel_dist[cur_ts], travel_dist[ts_update[j]]);
    mask[i] = (h == -1) | (d2 > thr_d2) | (dt > diff_td);
}

// ---- order-preserving compaction: 3-kernel scan over 1024-element tiles -----------------------
constexpr int kScanTile = 1024;

__global__ void k_tile_count(const uint8_t* __restrict__ flags, int64_t n, int32_t* __restrict__ tile_sums) {
    __shared__ int wsum[32];
    int64_t i = (int64_t)blockIdx.x * kScanTile + threadIdx.x;
    int f = (i < n) ? (flags[i] != 0) : 0;
    unsigned b = __ballot_sync(0xffffffffu, f);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = __popc(b);
    __syncthreads();
    if (threadIdx.x < 32) {
        int v = wsum[threadIdx.x];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (threadIdx.x == 0) tile_sums[blockIdx.x] = v;
    }
}

__global__ void k_tile_scan(int32_t* tile_sums, int64_t ntiles, int32_t* d_count) {
    // single block: exclusive scan of the tile sums in place
    __shared__ int wsum[32];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int64_t base = 0; base < ntiles; base += blockDim.x) {
        int64_t i = base + threadIdx.x;
        int v = (i < ntiles) ? tile_sums[i] : 0;
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, incl, o);
            if ((threadIdx.x & 31) >= o) incl += t;
        }
        if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = incl;
        __syncthreads();
        if (threadIdx.x < 32) {
            int w = wsum[threadIdx.x];
            int wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int t = __shfl_up_sync(0xffffffffu, wi, o);
                if (threadIdx.x >= o) wi += t;
            }
            wsum[threadIdx.x] = wi - w;  // exclusive warp offsets
        }
        __syncthreads();
        int excl = carry + wsum[threadIdx.x >> 5] + incl - v;
        if (i < ntiles) tile_sums[i] = excl;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) *d_count = carry;
}

// host launcher shared with index.cu: in-place exclusive scan of `ntiles` int32 sums, total -> *d_count
void launch_tile_scan(int32_t* tile_sums, int64_t ntiles, int32_t* d_count, cudaStream_t st) {
    k_tile_scan<<<1, 1024, 0, st>>>(tile_sums, ntiles, d_count);
}

__global__ void k_tile_emit(const uint8_t* __restrict__ flags, int64_t n, const int32_t* __restrict__ tile_offs,
                            int32_t* __restrict__ out_index) {
    __shared__ int wsum[32];
    int64_t i = (int64_t)blockIdx.x * kScanTile + threadIdx.x;
    int f = (i < n) ? (flags[i] != 0) : 0;
    unsigned b = __ballot_sync(0xffffffffu, f);
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) wsum[w] = __popc(b);
    __syncthreads();
    if (threadIdx.x < 32) {
        int v = wsum[threadIdx.x], incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (threadIdx.x >= o) incl += t;
        }
        wsum[threadIdx.x] = incl - v;
    }
    __syncthreads();
    if (f) out_index[tile_offs[blockIdx.x] + wsum[w] + __popc(b & ((1u << lane) - 1u))] = (int32_t)i;
}

// ---- voxel down-sampling (utils/tools.py:409-477) --------------------------------------------
// scratch layout (uint32, order-preserving float encoding): [0..2] min xyz, [3..5] max xyz, [6] max score
__global__ void k_vds_init(uint32_t* s) {
    if (threadIdx.x < 3) s[threadIdx.x] = 0xffffffffu;
    else if (threadIdx.x < 8) s[threadIdx.x] = 0u;
}

__device__ __forceinline__ float vds_center_dist(float x, float y, float z, float v) {
    // dist = sqrt(sum((p - (floor(p/v)+0.5)*v)^2)), tools.py:423-426, rounded op by op
    float gx = floorf(__fdiv_rn(x, v)), gy = floorf(__fdiv_rn(y, v)), gz = floorf(__fdiv_rn(z, v));
    float cx = __fmul_rn(__fadd_rn(gx, 0.5f), v), cy = __fmul_rn(__fadd_rn(gy, 0.5f), v),
          cz = __fmul_rn(__fadd_rn(gz, 0.5f), v);
    return __fsqrt_rn(dist2_rn(x, y, z, cx, cy, cz));
}

__global__ void k_vds_reduce(const float* __restrict__ xyz, int64_t n, float v, const float* __restrict__ value,
                             int mode, uint32_t* s) {
    float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY}, ms = -INFINITY;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
        mn[0] = fminf(mn[0], x); mn[1] = fminf(mn[1], y); mn[2] = fminf(mn[2], z);
        mx[0] = fmaxf(mx[0], x); mx[1] = fmaxf(mx[1], y); mx[2] = fmaxf(mx[2], z);
        float sc = mode == 0 ? vds_center_dist(x, y, z, v) : value[i];
        ms = fmaxf(ms, sc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
            mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
        }
        ms = fmaxf(ms, __shfl_xor_sync(0xffffffffu, ms, o));
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            atomicMin(&s[a], f2ord(mn[a]));
            atomicMax(&s[3 + a], f2ord(mx[a]));
        }
        atomicMax(&s[6], f2ord(ms));
    }
}

__global__ void k_vds_keys(const float* __restrict__ xyz, int64_t n, float v, const float* __restrict__ value,
                           int mode, const uint32_t* __restrict__ s, int64_t* __restrict__ keys,
                           int32_t* __restrict__ qscore) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    long long off[3], V = 0;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        off[a] = (long long)floorf(__fdiv_rn(ord2f(s[a]), v));              // tools.py:421
        long long top = (long long)floorf(__fdiv_rn(ord2f(s[3 + a]), v)) - off[a];
        V = top > V ? top : V;                                               // grid.max(): one scalar (H5)
    }
    float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
    long long g0 = (long long)floorf(__fdiv_rn(x, v)) - off[0];
    long long g1 = (long long)floorf(__fdiv_rn(y, v)) - off[1];
    long long g2 = (long long)floorf(__fdiv_rn(z, v)) - off[2];
    keys[i] = g0 + g1 * V + g2 * V * V;                                      // tools.py:431
    float sc = mode == 0 ? vds_center_dist(x, y, z, v) : value[i];
    float q = __fmul_rn(__fdiv_rn(sc, ord2f(s[6])), 999.0f);                 // tools.py:427 / :462
    qscore[i] = (int32_t)q;                                                  // .long(): truncation
}

constexpr uint64_t kEmptyKey = ~0ull;

__device__ __forceinline__ uint64_t mix64(uint64_t k) {
    k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
    return k;
}

__global__ void k_fill_u64(uint64_t* p, uint64_t n, uint64_t v) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        p[i] = v;
}

__global__ void k_vds_select(const int64_t* __restrict__ keys, const int32_t* __restrict__ qscore, int64_t n,
                             unsigned long long* tab_keys, unsigned long long* tab_best, uint64_t cap) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned long long key = (unsigned long long)keys[i];
    // same ordering as the reference's decimal packing i + q*10^digits: smaller q first, then smaller i
    unsigned long long packed = ((unsigned long long)(uint32_t)qscore[i] << 32) | (unsigned long long)(uint32_t)i;
    uint64_t h = mix64(key) & (cap - 1);
    while (true) {
        unsigned long long prev = atomicCAS(&tab_keys[h], (unsigned long long)kEmptyKey, key);
        if (prev == kEmptyKey || prev == key) {
            atomicMin(&tab_best[h], packed);
            return;
        }
        h = (h + 1) & (cap - 1);
    }
}

__global__ void k_vds_emit(const unsigned long long* __restrict__ tab_keys,
                           const unsigned long long* __restrict__ tab_best, uint64_t cap, int64_t* out_keys,
                           int32_t* out_index, int32_t* d_count) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool occ = i < cap && tab_keys[i] != kEmptyKey;
    unsigned b = __ballot_sync(0xffffffffu, occ);
    int lane = threadIdx.x & 31;
    int base = 0;
    if (lane == 0 && b) base = atomicAdd(d_count, __popc(b));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (occ) {
        int o = base + __popc(b & ((1u << lane) - 1u));
        out_keys[o] = (int64_t)tab_keys[i];
        out_index[o] = (int32_t)(tab_best[i] & 0xffffffffull);
    }
}

// ---- local-map window -----------------------------------------------------
__global__ void k_local_mask(const float* __restrict__ xyz, const int64_t* __restrict__ ts_c,
                             const int64_t* __restrict__ ts_u, int64_t m, const float* __restrict__ sensor3,
                             const float* __restrict__ td, int64_t cur_ts, float r2, float diff_td, int use_mid,
                             uint8_t* __restrict__ flags) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    // the sensor position lives on the device (cur_pose[:3,3]); read it through the pointer
    const float sx = sensor3[0], sy = sensor3[1], sz = sensor3[2];
    float d2 = dist2_rn(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], sx, sy, sz);
    int64_t t = ts_c[i];
    if (use_mid) t = (int64_t)((float)(ts_c[i] + ts_u[i]) / 2.0f);  // ((a+b)/2).long(): fp32 true-div then trunc
    float dt = fabsf(__fsub_rn(td[cur_ts], td[t]));
    flags[i] = (d2 < r2) & (dt < diff_td);
}

}  // namespace pinq

// built-in scratch of pinq_compact_flags_static: one int32 tile sum per 1024 flags, + 1 (file scope, see the entry point)
static __device__ __align__(256) int32_t g_compact_static_tiles[PINQ_COMPACT_STATIC_MAX_N / 1024 + 1];

// device address of g_compact_static_tiles on the CURRENT device (resolved once per device; NULL + error text on failure)
static int32_t* compact_static_scratch() {
    static int32_t* addr[64] = {nullptr};
    static std::mutex mu;
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) {
        pinq::set_error("pinq_compact_flags_static: cudaGetDevice failed or device index %d out of range: %s", dev,
                        cudaGetErrorString(e));
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_compact_static_tiles);
        if (e != cudaSuccess || !p) {
            pinq::set_error("pinq_compact_flags_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
            cudaGetLastError();
            return nullptr;
        }
        addr[dev] = reinterpret_cast<int32_t*>(p);
    }
    return addr[dev];
}

using namespace pinq;

extern "C" {

const char* pinq_last_error(void) { return g_err; }
int pinq_version(void) { return 100; }

void pinq_kernel_timing(int enable) {
    g_ktime_on = enable != 0;
    for (int i = 0; i < PINQ_KT_MARKS; ++i) g_ktime_set[i] = false;
}

int pinq_kernel_times(float* ms_out, int cap) {
    // ms_out[i] = time between mark i and the next recorded mark (phase i of pinq.h's PINQ_KT_* list); -1 if absent
    int nout = 0;
    for (int i = 0; i + 1 < PINQ_KT_MARKS && nout < cap; ++i, ++nout) {
        ms_out[i] = -1.f;
        if (!g_ktime_set[i] || i == PINQ_KT_FWD_END) continue;
        int j = i + 1;
        while (j < PINQ_KT_MARKS && !g_ktime_set[j]) ++j;
        if (j >= PINQ_KT_MARKS || (i < PINQ_KT_FWD_END && j > PINQ_KT_FWD_END)) continue;
        if (cudaEventSynchronize(g_ktime_ev[j]) != cudaSuccess) continue;
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, g_ktime_ev[i], g_ktime_ev[j]) == cudaSuccess) ms_out[i] = ms;
    }
    return nout;
}

int pinq_table_clear(int32_t* table, uint32_t buffer_size, void* stream) {
    PINQ_REQUIRE(table != nullptr && buffer_size > 0, "null table");
    k_fill_i32<<<148 * 8, 256, 0, as_stream(stream)>>>(table, (int64_t)buffer_size, -1);
    return check_launch("pinq_table_clear");
}

int pinq_hash_points(const float* xyz, int64_t n, float res, uint32_t buffer_size, uint32_t flags, int32_t* slots,
                     void* stream) {
    PINQ_REQUIRE(n >= 0 && res > 0.f && buffer_size > 0 && buffer_size < 0x7fffffffu, "bad size");
    if (n == 0) return PINQ_OK;
    k_hash_points<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(xyz, n, res, 1.0f / res,
                                                                   (flags & PINQ_F_RECIP_DIV) != 0, buffer_size, slots);
    return check_launch("pinq_hash_points");
}

int pinq_table_gather(const int32_t* table, const int32_t* slots, int64_t n, int32_t* out, void* stream) {
    if (n == 0) return PINQ_OK;
    k_gather_i32<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(table, slots, n, out);
    return check_launch("pinq_table_gather");
}

int pinq_table_scatter_last_wins(int32_t* table, const int32_t* slots, const int32_t* values, int64_t n,
                                 void* stream) {
    PINQ_REQUIRE(n < 0x7ffffff0LL, "batch too large");
    if (n == 0) return PINQ_OK;
    k_scatter_tag<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(table, slots, n);
    k_scatter_commit<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(table, slots, values, n);
    return check_launch("pinq_table_scatter_last_wins");
}

int pinq_update_mask(const float* sample_xyz, const int32_t* hash_idx, int64_t n, const float* map_xyz,
                     const int64_t* map_ts_update, int64_t m, const float* travel_dist, int64_t cur_ts, float res,
                     float diff_travel_dist_local, uint8_t* update_mask, void* stream) {
    if (n == 0) return PINQ_OK;
    // 3*res**2 is evaluated in double by Python and compared against an fp32 tensor: the scalar is
    // converted to fp32 for the comparison kernel (neural_points.py:231).
    float thr = (float)(3.0 * (double)res * (double)res);
    k_update_mask<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(sample_xyz, hash_idx, n, map_xyz, map_ts_update, m,
                                                                   travel_dist, cur_ts, thr, diff_travel_dist_local,
                                                                   update_mask);
    return check_launch("pinq_update_mask");
}

int64_t pinq_compact_scratch_bytes(int64_t n) { return ((n + kScanTile - 1) / kScanTile + 1) * (int64_t)sizeof(int32_t); }

int pinq_compact_flags(const uint8_t* flags, int64_t n, int32_t* out_index, int32_t* d_count, void* scratch,
                       void* stream) {
    PINQ_REQUIRE(n >= 0 && n < 0x7fffffffLL, "bad n");
    cudaStream_t st = as_stream(stream);
    if (n == 0) {
        cudaMemsetAsync(d_count, 0, sizeof(int32_t), st);
        return check_launch("pinq_compact_flags");
    }
    int64_t ntiles = (n + kScanTile - 1) / kScanTile;
    int32_t* tiles = reinterpret_cast<int32_t*>(scratch);
    k_tile_count<<<(unsigned)ntiles, kScanTile, 0, st>>>(flags, n, tiles);
    k_tile_scan<<<1, 1024, 0, st>>>(tiles, ntiles, d_count);
    k_tile_emit<<<(unsigned)ntiles, kScanTile, 0, st>>>(flags, n, tiles, out_index);
    return check_launch("pinq_compact_flags");
}

// ---- allocation-free variant of pinq_compact_flags (no counterpart in the reference; added
// for callers that cannot allocate: CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
// The tile sums live in a file-scope device array instead of caller scratch.  In HOST code the array's name is only a
// stand-in, not a device address: compact_static_scratch() asks the runtime for the address, once per device with that
// device current, and the kernels are given that pointer.  Every GPU that loads the library holds the 64 KB until exit.
// Nothing has to start at zero: k_tile_count overwrites every element a call reads.  All calls share the array, so
// calls that overlap on the GPU (other streams, other host threads) race; calls on one stream are ordered.
int pinq_compact_flags_static(const uint8_t* flags, int64_t n, int32_t* out_index, int32_t* d_count, void* stream) {
    PINQ_REQUIRE(n >= 0, "bad n");
    if (n > PINQ_COMPACT_STATIC_MAX_N) {
        set_error("pinq_compact_flags_static: n = %lld exceeds the built-in scratch (PINQ_COMPACT_STATIC_MAX_N = %lld flags); "
                  "use pinq_compact_flags with caller scratch", (long long)n, (long long)PINQ_COMPACT_STATIC_MAX_N);
        return PINQ_ERR_BAD_ARG;
    }
    PINQ_REQUIRE(d_count && (n == 0 || (flags && out_index)), "null pointer");
    cudaStream_t st = as_stream(stream);
    if (n == 0) {
        cudaMemsetAsync(d_count, 0, sizeof(int32_t), st);
        return check_launch("pinq_compact_flags_static");
    }
    int32_t* tiles = compact_static_scratch();
    if (!tiles) return PINQ_ERR_CUDA;  // (error text set by compact_static_scratch)
    const int64_t ntiles = (n + kScanTile - 1) / kScanTile;
    k_tile_count<<<(unsigned)ntiles, kScanTile, 0, st>>>(flags, n, tiles);
    k_tile_scan<<<1, 1024, 0, st>>>(tiles, ntiles, d_count);
    k_tile_emit<<<(unsigned)ntiles, kScanTile, 0, st>>>(flags, n, tiles, out_index);
    return check_launch("pinq_compact_flags_static");
}

int pinq_vds_prepare(const float* xyz, int64_t n, float voxel, const float* value, int mode, float* d_scratch8,
                     int64_t* keys, int32_t* qscore, void* stream) {
    PINQ_REQUIRE(n > 0 && voxel > 0.f, "bad args");
    PINQ_REQUIRE(mode == 0 || value != nullptr, "mode 1 needs value");
    cudaStream_t st = as_stream(stream);
    uint32_t* s = reinterpret_cast<uint32_t*>(d_scratch8);
    k_vds_init<<<1, 32, 0, st>>>(s);
    unsigned g = grid_for(n, 256);
    if (g > 148 * 8) g = 148 * 8;
    k_vds_reduce<<<g, 256, 0, st>>>(xyz, n, voxel, value, mode, s);
    k_vds_keys<<<grid_for(n, 256), 256, 0, st>>>(xyz, n, voxel, value, mode, s, keys, qscore);
    return check_launch("pinq_vds_prepare");
}

int pinq_vds_select(const int64_t* keys, const int32_t* qscore, int64_t n, uint64_t* tab_keys, uint64_t* tab_best,
                    uint64_t cap, void* stream) {
    PINQ_REQUIRE(cap >= 2 * (uint64_t)n && (cap & (cap - 1)) == 0, "cap must be a power of two >= 2n");
    cudaStream_t st = as_stream(stream);
    k_fill_u64<<<148 * 4, 256, 0, st>>>(tab_keys, cap, kEmptyKey);
    k_fill_u64<<<148 * 4, 256, 0, st>>>(tab_best, cap, ~0ull);
    k_vds_select<<<grid_for(n, 256), 256, 0, st>>>(keys, qscore, n, (unsigned long long*)tab_keys,
                                                   (unsigned long long*)tab_best, cap);
    return check_launch("pinq_vds_select");
}

int pinq_vds_emit(const uint64_t* tab_keys, const uint64_t* tab_best, uint64_t cap, int64_t n, int64_t* out_keys,
                  int32_t* out_index, int32_t* d_count, void* stream) {
    (void)n;
    cudaStream_t st = as_stream(stream);
    cudaMemsetAsync(d_count, 0, sizeof(int32_t), st);
    k_vds_emit<<<grid_for((int64_t)cap, 256), 256, 0, st>>>((const unsigned long long*)tab_keys,
                                                            (const unsigned long long*)tab_best, cap, out_keys,
                                                            out_index, d_count);
    return check_launch("pinq_vds_emit");
}

int pinq_local_mask(const float* map_xyz, const int64_t* ts_create, const int64_t* ts_update, int64_t m,
                    const float* sensor_xyz3, const float* travel_dist, int64_t cur_ts, float radius2,
                    float diff_travel_dist_local, int use_mid_ts, uint8_t* flags, void* stream) {
    if (m == 0) return PINQ_OK;
    k_local_mask<<<grid_for(m, 256), 256, 0, as_stream(stream)>>>(map_xyz, ts_create, ts_update, m, sensor_xyz3,
                                                                      travel_dist, cur_ts, radius2,
                                                                      diff_travel_dist_local, use_mid_ts, flags);
    return check_launch("pinq_local_mask");
}

}  // extern "C"

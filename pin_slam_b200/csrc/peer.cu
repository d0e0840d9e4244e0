// peer.cu -- the exchange step of data-parallel mapping over NVLink peer memory.
//
// The reference trains on one device: `cur_loss.backward(); opt.step()` (utils/mapper.py:572-573, utils/tools.py:89-116).
// With the batch sharded over R ranks the gradients of the replicated feature table and decoder have to be summed
// between those two lines.  Instead of NCCL all-reduce + a separate optimizer kernel, ONE kernel per rank does
//   barrier-in  ->  owner-sliced reduction with P2P loads from every rank's gradient block  ->  [Adam on the slice]
//   ->  P2P stores of the result (summed gradients, or the updated parameters) into every rank's block  ->  barrier-out
// (two-shot all-reduce: each rank moves (R-1)/R of the block in each direction over NVSwitch, every element crosses a
// link once per direction; barriers are flag words in peer memory, release/acquire at system scope).
//
// Every rank owns one "peer block" (cudaMalloc'ed so that it can be exported with cudaIpc), laid out identically:
//   [ signal pad: PINQ_PEER_PAD_BYTES ][ payload ... ]
// pad: uint32 flag[PINQ_PEER_MAX_GRID][8] (flag[b][r] = last phase value rank r's CTA b signalled), then an error word.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace pinq {

constexpr int kPeerThreads = 512;
// default of PINQ_PEER_CARVEOUT; -1 = leave the driver's choice
#ifndef PINQ_PEER_CARVEOUT_PCT
#define PINQ_PEER_CARVEOUT_PCT -1
#endif
constexpr int kPeerCarveoutPct = PINQ_PEER_CARVEOUT_PCT;
constexpr int kPadErrWord = PINQ_PEER_MAX_GRID * 8;  // uint32 index of the error word in the pad

struct PeerDev {
    float* block[PINQ_PEER_MAX];
    float* mc;                 // NVLS multicast address of the same block (all ranks' copies at once), or nullptr
    int rank, world;
    int64_t sum_off4, nsum4;   // fp32 sum region, in float4 units from the block base
    int64_t max_off2, nmax2;   // int64 max region, in longlong2 units
    int64_t par_off4, nadam4;  // parameter region (float4 units); first nadam4 float4 of the sum region are its gradients
    float* exp_avg;
    float* exp_avg_sq;
    float beta1, beta2, eps, step_size, bc2_sqrt;
    uint32_t phase;            // barrier-in value; barrier-out = phase + 1
    long long timeout_cycles;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// CTA b of every rank meets CTA b of every other rank.  Thread t < world signals rank t and waits for rank t.
__device__ __forceinline__ void peer_barrier(const PeerDev& a, uint32_t value) {
    __threadfence_system();
    __syncthreads();
    if ((int)threadIdx.x < a.world) {
        uint32_t* theirs = reinterpret_cast<uint32_t*>(a.block[threadIdx.x]) + blockIdx.x * 8 + a.rank;
        st_release_sys(theirs, value);
        uint32_t* mine = reinterpret_cast<uint32_t*>(a.block[a.rank]) + blockIdx.x * 8 + threadIdx.x;
        const long long t0 = clock64();
        while ((int32_t)(ld_acquire_sys(mine) - value) < 0) {
            if (clock64() - t0 > a.timeout_cycles) {  // a peer never arrived: flag it, do not hang the device
                atomicExch(reinterpret_cast<uint32_t*>(a.block[a.rank]) + kPadErrWord, value | 0x80000000u);
                break;
            }
            __nanosleep(64);
        }
    }
    __syncthreads();
}

// NVLS (NVSwitch multicast): ONE load returns the element summed over every rank's copy, reduced inside the switch, and ONE
// store lands in every rank's copy -- 2 memory instructions per 16 bytes of the owned slice instead of 2 x world.
__device__ __forceinline__ float4 multimem_ld_reduce_add(const float4* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st(float4* mc, const float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 :: "l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// torch.optim.Adam single-tensor arithmetic (k_adam, gn.cu) on float4 i of the owned slice; state lives on the owner only
__device__ __forceinline__ float4 adam_update4(const PeerDev& a, int64_t i, const float4 s) {
    float4 m = reinterpret_cast<float4*>(a.exp_avg)[i], q = reinterpret_cast<float4*>(a.exp_avg_sq)[i];
    float4 w = __ldcg(reinterpret_cast<const float4*>(a.block[a.rank]) + a.par_off4 + i);
    const float g[4] = {s.x, s.y, s.z, s.w};
    float* mm = &m.x; float* qq = &q.x; float* ww = &w.x;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        mm[c] = mm[c] + (g[c] - mm[c]) * (1.f - a.beta1);
        qq[c] = qq[c] * a.beta2 + (1.f - a.beta2) * g[c] * g[c];
        const float denom = sqrtf(qq[c]) / a.bc2_sqrt + a.eps;
        ww[c] = ww[c] - a.step_size * (mm[c] / denom);
    }
    reinterpret_cast<float4*>(a.exp_avg)[i] = m;
    reinterpret_cast<float4*>(a.exp_avg_sq)[i] = q;
    return w;
}

// T = threads per CTA.  512 (one CTA fills most of an SM's registers) is the stand-alone form.  128 with at most 56 registers
// is the form that runs UNDER another kernel: one of its CTAs fits into exactly the registers ONE retiring CTA of the
// neighbour search (128 threads x 56 registers, 9 per SM) leaves behind, so the exchange gets onto the SMs as soon as
// the first search CTAs retire instead of waiting for an SM that has drained completely.
template <int W, bool ADAM, int T>
__global__ void __launch_bounds__(T, T == 128 ? 9 : 1) k_peer_exchange(PeerDev a) {
    peer_barrier(a, a.phase);  // every rank's backward has finished writing its gradient block

    const int64_t tid = (int64_t)blockIdx.x * T + threadIdx.x;
    const int64_t nthr = (int64_t)gridDim.x * T;
    {   // fp32 sum region: this rank owns float4 [lo, hi).  U elements per thread and trip, all W x U loads issued before the
        // first add: the link wants megabytes in flight (latency ~2 us x 0.8 TB/s), not one 16-byte load per thread.
        // (T = 128: four 16-byte loads in flight per thread -- 56 registers without spills; the grid is 4 x larger instead)
        constexpr int U = T == 128 ? (W >= 4 ? 1 : 4 / W) : (W >= 8 ? 2 : (W == 4 ? 4 : 8));
        constexpr int PC = (T == 128 && W > 4) ? 4 : W;  // ranks loaded per round
        const int64_t lo = a.nsum4 * a.rank / W, hi = a.nsum4 * (a.rank + 1) / W;
        if (a.mc) {  // NVLS form: same slices, same element -> thread mapping (the zero_grad below relies on it)
            constexpr int UM = 4;
            float4* mc = reinterpret_cast<float4*>(a.mc);
            for (int64_t i0 = lo + tid; i0 < hi; i0 += nthr * UM) {
                float4 v[UM];
#pragma unroll
                for (int u = 0; u < UM; ++u) {
                    const int64_t i = i0 + u * nthr;
                    if (i < hi) v[u] = multimem_ld_reduce_add(mc + a.sum_off4 + i);
                }
#pragma unroll
                for (int u = 0; u < UM; ++u) {
                    const int64_t i = i0 + u * nthr;
                    if (i >= hi) break;
                    if (ADAM && i < a.nadam4) multimem_st(mc + a.par_off4 + i, adam_update4(a, i, v[u]));
                    else multimem_st(mc + a.sum_off4 + i, v[u]);
                }
            }
        } else
        for (int64_t i0 = lo + tid; i0 < hi; i0 += nthr * U) {
            float4 acc[U];
#pragma unroll
            for (int p0 = 0; p0 < W; p0 += PC) {
                float4 v[U][PC];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int64_t i = i0 + u * nthr;
                    if (i < hi) {
#pragma unroll
                        for (int p = 0; p < PC; ++p)
                            v[u][p] = __ldcg(reinterpret_cast<const float4*>(a.block[p0 + p]) + a.sum_off4 + i);
                    }
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {  // fixed rank order 0, 1, ..., W - 1: the sum does not depend on which rank owns the slice
                    float4 s = p0 == 0 ? v[u][0] : acc[u];
#pragma unroll
                    for (int p = (p0 == 0 ? 1 : 0); p < PC; ++p) { s.x += v[u][p].x; s.y += v[u][p].y; s.z += v[u][p].z; s.w += v[u][p].w; }
                    acc[u] = s;
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t i = i0 + u * nthr;
                if (i >= hi) break;
                const float4 s = acc[u];
                if (ADAM && i < a.nadam4) {
                    const float4 w = adam_update4(a, i, s);
#pragma unroll
                    for (int p = 0; p < W; ++p)
                        __stcg(reinterpret_cast<float4*>(a.block[p]) + a.par_off4 + i, w);  // parameter broadcast
                } else {
#pragma unroll
                    for (int p = 0; p < W; ++p) __stcg(reinterpret_cast<float4*>(a.block[p]) + a.sum_off4 + i, s);
                }
            }
        }
    }
    {   // int64 max region (local_point_ts_update: utils/../neural_points.py:643-645 amax semantics across ranks)
        const int64_t lo = a.nmax2 * a.rank / W, hi = a.nmax2 * (a.rank + 1) / W;
        for (int64_t i = lo + tid; i < hi; i += nthr) {
            longlong2 s = __ldcg(reinterpret_cast<const longlong2*>(a.block[0]) + a.max_off2 + i);
#pragma unroll
            for (int p = 1; p < W; ++p) {
                const longlong2 v = __ldcg(reinterpret_cast<const longlong2*>(a.block[p]) + a.max_off2 + i);
                s.x = v.x > s.x ? v.x : s.x;
                s.y = v.y > s.y ? v.y : s.y;
            }
#pragma unroll
            for (int p = 0; p < W; ++p) __stcg(reinterpret_cast<longlong2*>(a.block[p]) + a.max_off2 + i, s);
        }
    }
    peer_barrier(a, a.phase + 1);  // every slice has landed in this rank's block
    if (ADAM) {
        // zero_grad: every rank clears its OWN gradient block once all owners have read it (local stores at HBM speed
        // instead of W - 1 remote zero streams over the links).  peer_barrier only orders CTA b of this rank after CTA b
        // of the owners, so thread (b, t) may clear exactly the elements that thread (b, t) of owner p loaded:
        // lo_p + tid + j * nthr.  (Clearing with i = tid + j * nthr raced with slower CTAs of the owner -- ADVICE r1.)
        float4* g = reinterpret_cast<float4*>(a.block[a.rank]) + a.sum_off4;
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int p = 0; p < W; ++p) {
            const int64_t lo = a.nsum4 * p / W, hi0 = a.nsum4 * (p + 1) / W;
            const int64_t hi = hi0 < a.nadam4 ? hi0 : a.nadam4;
            for (int64_t i = lo + tid; i < hi; i += nthr) __stcg(g + i, z);
        }
    }
}

// Cooperative launch: the flag barrier needs every CTA of the grid resident at the same time on every rank.  A plain
// launch only has that by luck when another kernel shares the device (e.g. the next step's search on another stream);
// cudaLaunchCooperativeKernel makes it a launch-time guarantee (or a launch error) instead of a barrier timeout.
template <int W, int T>
static cudaError_t launch_peer_t(PeerDev& d, bool adam, bool cooperative, int grid, cudaStream_t st) {
    static bool configured = false;
    if (!configured) {  // PINQ_PEER_CARVEOUT: shared-memory carve-out (percent) the exchange asks for; -1 = the driver's choice
        configured = true;
        const char* ce = getenv("PINQ_PEER_CARVEOUT");
        const int carve = ce ? atoi(ce) : kPeerCarveoutPct;
        if (carve >= 0) {
            cudaFuncSetAttribute(k_peer_exchange<W, true, T>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            cudaFuncSetAttribute(k_peer_exchange<W, false, T>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
        }
    }
    if (!cooperative) {  // several simulated ranks in one process: their kernels wait for each other, the caller vouches
        if (adam) k_peer_exchange<W, true, T><<<grid, T, 0, st>>>(d);
        else k_peer_exchange<W, false, T><<<grid, T, 0, st>>>(d);
        return cudaGetLastError();
    }
    void* args[] = {&d};
    const void* fn = adam ? reinterpret_cast<const void*>(&k_peer_exchange<W, true, T>)
                          : reinterpret_cast<const void*>(&k_peer_exchange<W, false, T>);
    return cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(T), args, 0, st);
}

// Cooperative launch: the flag barrier needs every CTA of the grid resident at the same time on every rank.  A plain
// launch only has that by luck when another kernel shares the device (e.g. the next step's search on another stream);
// cudaLaunchCooperativeKernel makes it a launch-time guarantee (or a launch error) instead of a barrier timeout.
template <int W>
static cudaError_t launch_peer(PeerDev& d, bool adam, bool cooperative, int grid, int threads, cudaStream_t st) {
    return threads == 128 ? launch_peer_t<W, 128>(d, adam, cooperative, grid, st)
                          : launch_peer_t<W, kPeerThreads>(d, adam, cooperative, grid, st);
}

}  // namespace pinq

using namespace pinq;

extern "C" {

int pinq_peer_alloc(int64_t bytes, void** d_out) {
    PINQ_REQUIRE(d_out && bytes >= PINQ_PEER_PAD_BYTES, "need at least the signal pad");
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, (size_t)bytes);
    if (e == cudaSuccess) e = cudaMemset(p, 0, (size_t)bytes);
    if (e != cudaSuccess) {
        set_error("pinq_peer_alloc: %s", cudaGetErrorString(e));
        if (p) cudaFree(p);
        return PINQ_ERR_CUDA;
    }
    *d_out = p;
    return PINQ_OK;
}

int pinq_peer_free(void* d_block) {
    if (!d_block) return PINQ_OK;
    cudaError_t e = cudaFree(d_block);
    if (e != cudaSuccess) { set_error("pinq_peer_free: %s", cudaGetErrorString(e)); return PINQ_ERR_CUDA; }
    return PINQ_OK;
}

int pinq_peer_export(void* d_block, void* h_handle) {
    PINQ_REQUIRE(d_block && h_handle, "null pointer");
    static_assert(sizeof(cudaIpcMemHandle_t) == PINQ_PEER_HANDLE_BYTES, "handle size");
    cudaError_t e = cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(h_handle), d_block);
    if (e != cudaSuccess) { set_error("pinq_peer_export: %s", cudaGetErrorString(e)); return PINQ_ERR_CUDA; }
    return PINQ_OK;
}

int pinq_peer_open(const void* h_handle, void** d_out) {
    PINQ_REQUIRE(h_handle && d_out, "null pointer");
    cudaIpcMemHandle_t h;
    memcpy(&h, h_handle, sizeof(h));
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { set_error("pinq_peer_open: %s", cudaGetErrorString(e)); return PINQ_ERR_CUDA; }
    *d_out = p;
    return PINQ_OK;
}

int pinq_peer_close(void* d_mapped) {
    if (!d_mapped) return PINQ_OK;
    cudaError_t e = cudaIpcCloseMemHandle(d_mapped);
    if (e != cudaSuccess) { set_error("pinq_peer_close: %s", cudaGetErrorString(e)); return PINQ_ERR_CUDA; }
    return PINQ_OK;
}

int64_t pinq_peer_error_offset(void) { return (int64_t)kPadErrWord * 4; }

int pinq_peer_exchange(const pinq_peer_args* a, void* stream) {
    PINQ_REQUIRE(a, "null args");
    PINQ_REQUIRE(a->world >= 1 && a->world <= PINQ_PEER_MAX && a->rank >= 0 && a->rank < a->world, "bad rank/world");
    PINQ_REQUIRE(a->world == 1 || a->world == 2 || a->world == 4 || a->world == 8, "world must be 1, 2, 4 or 8");
    for (int p = 0; p < a->world; ++p) PINQ_REQUIRE(a->block[p], "null peer block");
    PINQ_REQUIRE(a->n_sum >= 0 && a->n_sum % 4 == 0 && a->sum_off % 16 == 0 && a->sum_off >= PINQ_PEER_PAD_BYTES,
                 "sum region: offset 16-byte aligned behind the pad, count a multiple of 4");
    PINQ_REQUIRE(a->n_max >= 0 && a->n_max % 2 == 0 && (a->n_max == 0 || (a->max_off % 16 == 0 && a->max_off >= PINQ_PEER_PAD_BYTES)),
                 "max region: offset 16-byte aligned behind the pad, count a multiple of 2");
    PINQ_REQUIRE(a->epoch >= 1 && a->epoch < 0x3fffffffu, "epoch counts calls from 1");
    const bool adam = a->n_adam > 0;
    if (adam) {
        PINQ_REQUIRE(a->n_adam % 4 == 0 && a->n_adam <= a->n_sum && a->param_off % 16 == 0 && a->param_off >= PINQ_PEER_PAD_BYTES,
                     "adam region: a multiple of 4 leading elements of the sum region, parameters 16-byte aligned");
        PINQ_REQUIRE(a->exp_avg && a->exp_avg_sq && a->adam_step >= 1, "adam state");
    }
    int grid = a->grid > 0 ? a->grid : 148;  // one CTA per SM, launched cooperatively: co-residency is guaranteed
    PINQ_REQUIRE(grid <= PINQ_PEER_MAX_GRID, "grid too large for the signal pad");
    PINQ_REQUIRE(a->threads == 0 || a->threads == 128 || a->threads == kPeerThreads, "threads: 0 (= 512), 128 or 512");
    PeerDev d{};
    for (int p = 0; p < a->world; ++p) d.block[p] = reinterpret_cast<float*>(a->block[p]);
    d.mc = a->world > 1 ? reinterpret_cast<float*>(a->mc_block) : nullptr;
    d.rank = a->rank;
    d.world = a->world;
    d.sum_off4 = a->sum_off / 16;
    d.nsum4 = a->n_sum / 4;
    d.max_off2 = a->max_off / 16;
    d.nmax2 = a->n_max / 2;
    d.par_off4 = a->param_off / 16;
    d.nadam4 = a->n_adam / 4;
    d.exp_avg = a->exp_avg;
    d.exp_avg_sq = a->exp_avg_sq;
    if (adam) {
        const double bc1 = 1.0 - pow((double)a->beta1, (double)a->adam_step);
        const double bc2 = 1.0 - pow((double)a->beta2, (double)a->adam_step);
        d.beta1 = a->beta1;
        d.beta2 = a->beta2;
        d.eps = a->adam_eps;
        d.step_size = (float)((double)a->lr / bc1);
        d.bc2_sqrt = (float)sqrt(bc2);
    }
    d.phase = a->epoch * 2u;
    const double ms = a->timeout_ms > 0 ? a->timeout_ms : 5000.0;
    d.timeout_cycles = (long long)(ms * 1.9e6);  // SM clock ~1.9 GHz; a bound, not a measurement
    cudaStream_t st = as_stream(stream);
    cudaError_t le;
    switch (a->world) {
        case 1: le = launch_peer<1>(d, adam, a->cooperative != 0, grid, a->threads, st); break;
        case 2: le = launch_peer<2>(d, adam, a->cooperative != 0, grid, a->threads, st); break;
        case 4: le = launch_peer<4>(d, adam, a->cooperative != 0, grid, a->threads, st); break;
        default: le = launch_peer<8>(d, adam, a->cooperative != 0, grid, a->threads, st); break;
    }
    if (le != cudaSuccess) {
        set_error("pinq_peer_exchange: %slaunch of %d CTAs failed: %s", a->cooperative ? "cooperative " : "", grid, cudaGetErrorString(le));
        cudaGetLastError();
        return PINQ_ERR_CUDA;
    }
    return check_launch("pinq_peer_exchange");
}

}  // extern "C"

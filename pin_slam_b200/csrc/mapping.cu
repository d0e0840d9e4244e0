// mapping.cu -- the mapper's training iterations as one device-side loop (reference: utils/mapper.py:425-593).
//
// Mapper.mapping() in the reference is ~110 small launches per iteration (index_select of the replay pool,
// 7 query passes, loss arithmetic, autograd bookkeeping, multi-kernel Adam).  Here an iteration is
//   k_map_gather -> fused query forward -> k_map_loss_grad -> fused query backward -> k_adam_multi
// and `pinq_mapping_run` enqueues all iterations of a frame back to back: the host is not on the path.
#include <math.h>

#include <stdlib.h>

#include <mutex>

#include "common.cuh"

namespace pinq {

struct GatherParams {
    const float* pool_xyz;
    const float* pool_label;
    const float* pool_weight;
    const int64_t* pool_ts;
    const int64_t* draws_hist;  // this iteration's row
    const int64_t* draws_new;
    const int64_t* new_idx;
    int bs, bs_hist, ng, decim, eik_mode, eik_start;
    float eps, surface_range;
    float* xyz;
    int64_t* ts;
    float* label;
    float* weight;
    int32_t* eik_count;  // this iteration's counter
};

// pool row of batch sample i: history draws first, then the draws among the new samples (mapper.py:336-343)
__device__ __forceinline__ int64_t pool_row(const GatherParams& p, int i) {
    return i < p.bs_hist ? p.draws_hist[i] : p.new_idx[p.draws_new[i - p.bs_hist]];
}

// One thread per query of the iteration's batch: [0, bs) the training samples, then six groups of ng points
// x[::decim] +/- eps along x, y, z in the order mapper.py:697 concatenates them (data-parallel form: the members of the
// GLOBAL batch's x[::decim] that fall into this rank's slice, i.e. local rows eik_start, eik_start + decim, ...).
__global__ void k_map_gather(GatherParams p) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = p.bs + 6 * p.ng;
    if (i >= n) return;
    if (i < p.bs) {
        const int64_t r = pool_row(p, i);
        p.xyz[3 * i + 0] = p.pool_xyz[3 * r + 0];
        p.xyz[3 * i + 1] = p.pool_xyz[3 * r + 1];
        p.xyz[3 * i + 2] = p.pool_xyz[3 * r + 2];
        p.ts[i] = p.pool_ts[r];
        const float lab = p.pool_label[r];
        p.label[i] = lab;
        p.weight[i] = fabsf(p.pool_weight[r]);
        if (p.eik_mode != PINQ_EIK_ALL && p.decim > 0 && i % p.decim == 0) {  // (single device only: eik_start == 0)
            const bool surface = fabsf(lab) < p.surface_range;
            if (surface == (p.eik_mode == PINQ_EIK_SURFACE)) atomicAdd(p.eik_count, 1);
        }
        return;
    }
    const int e = i - p.bs;
    const int group = e / p.ng, j = e - group * p.ng;
    const int64_t r = pool_row(p, p.eik_start + j * p.decim);
    float v[3] = {p.pool_xyz[3 * r + 0], p.pool_xyz[3 * r + 1], p.pool_xyz[3 * r + 2]};
    const int axis = group >> 1;
    v[axis] = (group & 1) ? __fsub_rn(v[axis], p.eps) : __fadd_rn(v[axis], p.eps);
    p.xyz[3 * i + 0] = v[0];
    p.xyz[3 * i + 1] = v[1];
    p.xyz[3 * i + 2] = v[2];
    p.ts[i] = -1;  // Mapper.sdf() passes no timestamps (H9): amax with -1 never changes a frame id
}

struct LossParams {
    const float* sdf;     // [bs + 6 ng]
    const float* label;   // [bs]
    const float* weight;  // [bs]
    int bs, ng, decim, eik_mode, weighted, eik_start;
    int norm_bs, norm_ng;  // the GLOBAL batch's sample / eikonal counts (data-parallel form; = bs, ng on one device)
    float sigma, weight_e, two_eps, surface_range;
    const int32_t* eik_count;
    float* grad_sdf;
    float* loss;  // [2] or NULL
};

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

// dL/dsdf of  mean_i w_i bce(sdf_i / sigma, sigmoid(label_i / sigma))  +  weight_e * mean_j (|g_j| - 1)^2,
// g_j the two-sided difference quotient (loss.py:40-47, mapper.py:526-547, 697-712).
__global__ void k_map_loss_grad(LossParams p) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float l_bce = 0.f, l_eik = 0.f;
    if (i < p.bs) {
        const float z = p.sdf[i] / p.sigma;
        const float t = sigmoidf_(p.label[i] / p.sigma);
        const float w = p.weighted ? p.weight[i] : 1.f;
        p.grad_sdf[i] = ((sigmoidf_(z) - t) * w) * (1.f / (float)p.norm_bs) / p.sigma;
        const float mz = fmaxf(-z, 0.f);
        l_bce = w * ((1.f - t) * z + mz + logf(expf(-mz) + expf(-z - mz)));
    } else if (i < p.bs + p.ng) {
        const int j = i - p.bs;
        bool selected = true;
        int count = p.norm_ng;
        if (p.eik_mode != PINQ_EIK_ALL) {
            const bool surface = fabsf(p.label[p.eik_start + j * p.decim]) < p.surface_range;
            selected = surface == (p.eik_mode == PINQ_EIK_SURFACE);
            count = *p.eik_count;
        }
        const float* s = p.sdf + p.bs + j;
        float* gs = p.grad_sdf + p.bs + j;
        float g[3], d[3] = {0.f, 0.f, 0.f};
#pragma unroll
        for (int a = 0; a < 3; ++a) g[a] = (s[(2 * a) * p.ng] - s[(2 * a + 1) * p.ng]) / p.two_eps;
        if (selected && count > 0) {
            const float nrm = sqrtf(g[0] * g[0] + g[1] * g[1] + g[2] * g[2]);
            const float c = p.weight_e * (2.f * (nrm - 1.f) / (float)count);
            if (nrm > 0.f) {
#pragma unroll
                for (int a = 0; a < 3; ++a) d[a] = c * (g[a] / nrm) / p.two_eps;
            }
            l_eik = (nrm - 1.f) * (nrm - 1.f);
        }
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            gs[(2 * a) * p.ng] = d[a];
            gs[(2 * a + 1) * p.ng] = -d[a];
        }
    }
    if (p.loss) {  // logging only
        l_bce = warp_sum(l_bce);
        l_eik = warp_sum(l_eik);
        if ((threadIdx.x & 31) == 0) {
            if (l_bce != 0.f) atomicAdd(&p.loss[0], l_bce / (float)p.norm_bs);
            if (l_eik != 0.f) {
                const int count = p.eik_mode == PINQ_EIK_ALL ? p.norm_ng : *p.eik_count;
                atomicAdd(&p.loss[1], l_eik / (float)count);
            }
        }
    }
}

struct AdamPack {
    pinq_adam_tensor t[8];
    int64_t end[8];  // exclusive prefix end in groups of 4 elements
    int count;
};

// torch.optim.Adam single-tensor arithmetic (see k_adam in gn.cu) over up to 8 tensors in one launch; reads
// and clears the gradient (zero_grad), treats the moments as zero at the first step of a fresh optimizer.
template <bool VEC>
__global__ void k_adam_multi(AdamPack pk, int64_t groups, float beta1, float beta2, float eps, float step_size,
                             float bc2_sqrt, float weight_decay, int first) {
    const int64_t gi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gi >= groups) return;
    int t = 0;
    while (t < pk.count - 1 && gi >= pk.end[t]) ++t;
    const int64_t j0 = (gi - (t ? pk.end[t - 1] : 0)) * 4;
    const pinq_adam_tensor T = pk.t[t];
    const int cnt = (int)min((int64_t)4, T.n - j0);
    float p[4], g[4], m[4] = {0.f, 0.f, 0.f, 0.f}, v[4] = {0.f, 0.f, 0.f, 0.f};
    const bool vec = VEC && cnt == 4;
    if (vec) {
        *reinterpret_cast<float4*>(p) = *reinterpret_cast<const float4*>(T.param + j0);
        *reinterpret_cast<float4*>(g) = *reinterpret_cast<const float4*>(T.grad + j0);
        if (!first) {
            *reinterpret_cast<float4*>(m) = *reinterpret_cast<const float4*>(T.exp_avg + j0);
            *reinterpret_cast<float4*>(v) = *reinterpret_cast<const float4*>(T.exp_avg_sq + j0);
        }
    } else {
        for (int e = 0; e < cnt; ++e) {
            p[e] = T.param[j0 + e];
            g[e] = T.grad[j0 + e];
            if (!first) {
                m[e] = T.exp_avg[j0 + e];
                v[e] = T.exp_avg_sq[j0 + e];
            }
        }
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        if (e >= cnt) break;
        float ge = g[e];
        if (weight_decay != 0.f) ge = ge + weight_decay * p[e];
        m[e] = m[e] + (ge - m[e]) * (1.f - beta1);
        v[e] = v[e] * beta2 + (1.f - beta2) * ge * ge;
        const float denom = sqrtf(v[e]) / bc2_sqrt + eps;
        p[e] = p[e] - step_size * (m[e] / denom);
        g[e] = 0.f;
    }
    if (vec) {
        *reinterpret_cast<float4*>(T.param + j0) = *reinterpret_cast<float4*>(p);
        *reinterpret_cast<float4*>(T.grad + j0) = *reinterpret_cast<float4*>(g);
        *reinterpret_cast<float4*>(T.exp_avg + j0) = *reinterpret_cast<float4*>(m);
        *reinterpret_cast<float4*>(T.exp_avg_sq + j0) = *reinterpret_cast<float4*>(v);
    } else {
        for (int e = 0; e < cnt; ++e) {
            T.param[j0 + e] = p[e];
            T.grad[j0 + e] = 0.f;
            T.exp_avg[j0 + e] = m[e];
            T.exp_avg_sq[j0 + e] = v[e];
        }
    }
}

// Data-parallel form: the exchange + optimizer kernel of iteration i runs on a side stream UNDER the batch gather, the binning
// and the neighbour search of iteration i + 1 (they read sample positions and map geometry only); the decode of i + 1 -- the
// first kernel that reads the parameters the optimizer updates -- waits for it (pinq_query_args.decode_wait_event).  One
// side stream + two events per calling stream, created on first use and kept (the only objects this file creates).
struct MapSide {
    cudaStream_t main, side;
    cudaEvent_t bwd_done, exchanged;
};
static MapSide* map_side_for(cudaStream_t st) {
    static MapSide pool[32];
    static int used = 0;
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    for (int i = 0; i < used; ++i)
        if (pool[i].main == st) return &pool[i];
    if (used == 32) return nullptr;
    MapSide& m = pool[used];
    m.main = st;
    // highest priority: the exchange's few CTAs take the SM slots the next iteration's search frees first
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    if (cudaStreamCreateWithPriority(&m.side, cudaStreamNonBlocking, prio_hi) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&m.bwd_done, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&m.exchanged, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    ++used;
    return &m;
}

}  // namespace pinq

using namespace pinq;

extern "C" int pinq_mapping_run(const pinq_mapping_args* a, int32_t iters, void* stream) {
    PINQ_REQUIRE(a, "null args");
    PINQ_REQUIRE(iters >= 0, "negative iteration count");
    if (iters == 0) return PINQ_OK;
    PINQ_REQUIRE(a->bs > 0 && a->bs_new >= 0 && a->bs_new <= a->bs, "bad batch sizes");
    PINQ_REQUIRE(a->pool_xyz && a->pool_label && a->pool_weight && a->pool_ts, "null replay pool");
    PINQ_REQUIRE(a->draws_hist || a->bs_new == a->bs, "null history draws");
    PINQ_REQUIRE(a->bs_new == 0 || (a->draws_new && a->new_idx), "null new-sample draws");
    PINQ_REQUIRE(a->batch_xyz && a->batch_ts && a->batch_label && a->batch_weight && a->grad_sdf, "null batch buffers");
    PINQ_REQUIRE(a->tensor_count >= 1 && a->tensor_count <= 8, "1..8 optimizer tensors");
    PINQ_REQUIRE(a->eik_mode >= PINQ_EIK_ALL && a->eik_mode <= PINQ_EIK_FREESPACE, "bad eik_mode");
    const int decim = a->decimation;
    // data-parallel form: this rank holds samples [shard_first, shard_first + bs) of a global batch of global_bs samples
    const int global_bs = a->global_bs > 0 ? a->global_bs : a->bs;
    const bool sharded = a->global_bs > 0 || a->peer;
    PINQ_REQUIRE(a->shard_first >= 0 && a->shard_first + a->bs <= global_bs, "shard outside the global batch");
    if (sharded && (a->eik_mode != PINQ_EIK_ALL || a->weight_decay != 0.f)) {
        set_error("pinq_mapping_run: the data-parallel form supports ekional_add_to = all and no weight decay");
        return PINQ_ERR_UNSUPPORTED;
    }
    const int eik_start = decim > 0 ? (decim - a->shard_first % decim) % decim : 0;
    const int ng = decim > 0 && eik_start < a->bs ? (a->bs - eik_start + decim - 1) / decim : 0;
    const int global_ng = decim > 0 ? (global_bs + decim - 1) / decim : 0;
    const int64_t n = (int64_t)a->bs + 6 * (int64_t)ng;
    PINQ_REQUIRE(a->query.n == n && a->bwd.n == n, "query / bwd n must be bs + 6 * (eikonal samples of this shard)");
    PINQ_REQUIRE(a->query.xyz == a->batch_xyz && a->query.sdf && a->bwd.grad_sdf == a->grad_sdf, "batch buffers not wired");
    PINQ_REQUIRE(a->eik_mode == PINQ_EIK_ALL || a->eik_count, "eik_count needed for surface / freespace eikonal");
    cudaStream_t st = as_stream(stream);

    AdamPack pk;
    int64_t groups = 0;
    bool aligned = true;
    pk.count = a->tensor_count;
    for (int t = 0; t < 8; ++t) {
        pk.t[t] = t < pk.count ? a->tensors[t] : pinq_adam_tensor{nullptr, nullptr, nullptr, nullptr, 0};
        if (t < pk.count) {
            const pinq_adam_tensor& T = a->tensors[t];
            PINQ_REQUIRE(T.param && T.grad && T.exp_avg && T.exp_avg_sq && T.n > 0, "bad optimizer tensor");
            groups += (T.n + 3) / 4;
            aligned = aligned && (((uintptr_t)T.param | (uintptr_t)T.grad | (uintptr_t)T.exp_avg | (uintptr_t)T.exp_avg_sq) & 15) == 0;
        }
        pk.end[t] = groups;
    }

    GatherParams gp;
    gp.pool_xyz = a->pool_xyz; gp.pool_label = a->pool_label; gp.pool_weight = a->pool_weight; gp.pool_ts = a->pool_ts;
    gp.new_idx = a->new_idx;
    gp.bs = a->bs; gp.bs_hist = a->bs - a->bs_new; gp.ng = ng; gp.decim = decim; gp.eik_mode = a->eik_mode; gp.eik_start = eik_start;
    gp.eps = a->eps_step; gp.surface_range = a->surface_range;
    gp.xyz = a->batch_xyz; gp.ts = a->batch_ts; gp.label = a->batch_label; gp.weight = a->batch_weight;

    LossParams lp;
    lp.sdf = a->query.sdf; lp.label = a->batch_label; lp.weight = a->batch_weight;
    lp.bs = a->bs; lp.ng = ng; lp.decim = decim; lp.eik_mode = a->eik_mode; lp.weighted = a->weighted; lp.eik_start = eik_start;
    lp.norm_bs = global_bs; lp.norm_ng = global_ng;
    lp.sigma = a->sigma; lp.weight_e = ng > 0 ? a->weight_e : 0.f;
    lp.two_eps = (float)(2.0 * (double)a->eps_step);
    lp.surface_range = a->surface_range;
    lp.grad_sdf = a->grad_sdf;

    MapSide* ms = nullptr;
    if (a->peer && a->peer->world > 1) {
        static int overlap = -1;
        if (overlap < 0) {
            const char* e = getenv("PINQ_MAPPING_OVERLAP");
            overlap = !(e && e[0] == '0');
        }
        if (overlap) ms = map_side_for(st);
    }
    for (int it = 0; it < iters; ++it) {
        gp.draws_hist = a->draws_hist ? a->draws_hist + (int64_t)it * gp.bs_hist : nullptr;
        gp.draws_new = a->draws_new ? a->draws_new + (int64_t)it * a->bs_new : nullptr;
        gp.eik_count = a->eik_count ? a->eik_count + it : nullptr;
        k_map_gather<<<grid_for(n, 256), 256, 0, st>>>(gp);
        int rc = check_launch("k_map_gather");
        if (rc != PINQ_OK) return rc;
        if (ms && it > 0) {  // gather + binning + search now, the decode after the previous iteration's exchange + optimizer
            pinq_query_args q = a->query;
            q.decode_wait_event = ms->exchanged;
            rc = pinq_query_fused_fwd(&q, stream);
        } else {
            rc = pinq_query_fused_fwd(&a->query, stream);
        }
        if (rc != PINQ_OK) return rc;
        lp.eik_count = gp.eik_count;
        lp.loss = a->loss ? a->loss + 2 * it : nullptr;
        k_map_loss_grad<<<grid_for(a->bs + ng, 256), 256, 0, st>>>(lp);
        rc = check_launch("k_map_loss_grad");
        if (rc != PINQ_OK) return rc;
        rc = pinq_query_fused_bwd(&a->bwd, stream);
        if (rc != PINQ_OK) return rc;
        const int step = it + 1;
        if (a->peer) {  // exchange + optimizer as ONE kernel over peer memory (csrc/peer.cu): sum of the ranks' gradient blocks,
                        // Adam on the owned slice, parameters broadcast, gradients zeroed
            pinq_peer_args pa = *a->peer;
            pa.epoch = a->peer->epoch + (uint32_t)it;
            pa.adam_step = step;
            pa.lr = a->lr; pa.beta1 = a->beta1; pa.beta2 = a->beta2; pa.adam_eps = a->adam_eps;
            if (ms) {
                // plain launch: a cooperative grid would wait for an empty GPU, i.e. behind the next iteration's search; its
                // CTAs take SM slots as they free up and meet at the flag barrier (nothing the search needs depends on them)
                pa.cooperative = 0;
                cudaEventRecord(ms->bwd_done, st);
                cudaStreamWaitEvent(ms->side, ms->bwd_done, 0);
                rc = pinq_peer_exchange(&pa, ms->side);
                cudaEventRecord(ms->exchanged, ms->side);
            } else {
                rc = pinq_peer_exchange(&pa, stream);
            }
            if (rc != PINQ_OK) return rc;
            continue;
        }
        const double bc1 = 1.0 - pow((double)a->beta1, (double)step);
        const double bc2 = 1.0 - pow((double)a->beta2, (double)step);
        const float step_size = (float)((double)a->lr / bc1), bc2_sqrt = (float)sqrt(bc2);
        if (aligned)
            k_adam_multi<true><<<grid_for(groups, 256), 256, 0, st>>>(pk, groups, a->beta1, a->beta2, a->adam_eps, step_size,
                                                                      bc2_sqrt, a->weight_decay, it == 0);
        else
            k_adam_multi<false><<<grid_for(groups, 256), 256, 0, st>>>(pk, groups, a->beta1, a->beta2, a->adam_eps, step_size,
                                                                       bc2_sqrt, a->weight_decay, it == 0);
        rc = check_launch("k_adam_multi");
        if (rc != PINQ_OK) return rc;
    }
    if (ms) cudaStreamWaitEvent(st, ms->exchanged, 0);  // the call is complete on `stream` when the last exchange has landed
    return PINQ_OK;
}

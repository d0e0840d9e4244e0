// query.cu -- neighbour search (API form), fused neural-point SDF query forward (+ d sdf/dx, std),
// un-fused query_feature, mesher dense-grid query.
// Reference: model/neural_points.py:459-674, model/decoder.py:66-88, utils/tracker.py:245-253,
//            utils/mapper.py:466-468, utils/mesher.py:93-180.
#include "query_common.cuh"

namespace pinq {

// ---------------------------------------------------------------------------------------------
// Standalone API form: dist2[n,kn], idx[n,kn] (global int64).  One thread per (query, cell).
__global__ void k_neighbor_search(MapDev map, const float* __restrict__ xyz, int64_t n, bool recip,
                                  const float* __restrict__ map_xyz, const int64_t* __restrict__ ts_create,
                                  const float* __restrict__ travel_dist, int64_t cur_ts, float diff_td,
                                  bool time_filtering, float* __restrict__ out_d2, int64_t* __restrict__ out_idx) {
    extern __shared__ uint32_t s_off_ns[];
    load_stencil(map, s_off_ns);
    __syncthreads();
    const int64_t total = n * map.kn;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t q = t / map.kn;
        const int c = (int)(t - q * map.kn);
        const float x = xyz[3 * q], y = xyz[3 * q + 1], z = xyz[3 * q + 2];
        const uint32_t base = base_slot(map, x, y, z, recip);
        int32_t idx = map.table[probe_slot(base, s_off_ns[c], map.B)];
        if (time_filtering && map.m > 0) {
            const int64_t j = idx < 0 ? map.m - 1 : idx;  // -1 is a live wrapping index (hazard H8)
            const float dt = fabsf(__fsub_rn(travel_dist[cur_ts], travel_dist[ts_create[j]]));
            if (!(dt < diff_td)) idx = -1;
        }
        float d2 = map.max_d2;
        if (idx >= 0) {
            const int64_t j = idx;
            d2 = dist2_rn(map_xyz[3 * j], map_xyz[3 * j + 1], map_xyz[3 * j + 2], x, y, z);
            if (d2 > map.max_d2) idx = -1;  // the distance itself is kept (neural_points.py:495-499)
        }
        out_d2[t] = d2;
        out_idx[t] = (int64_t)idx;
    }
}

__global__ void k_query_certainty(MapDev map, const float* __restrict__ xyz, int64_t n, bool recip,
                                  const float* __restrict__ map_xyz, const float* __restrict__ cert,
                                  float* __restrict__ out) {
    extern __shared__ uint32_t s_off_qc[];
    load_stencil(map, s_off_qc);
    __syncthreads();
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    const float x = xyz[3 * q], y = xyz[3 * q + 1], z = xyz[3 * q + 2];
    const uint32_t base = base_slot(map, x, y, z, recip);
    // torch.max over the row of certainties with invalid entries set to 0 (neural_points.py:518-521)
    float best = -INFINITY;
    for (int c = 0; c < map.kn; ++c) {
        const int32_t idx = __ldg(&map.table[probe_slot(base, s_off_qc[c], map.B)]);
        float v = 0.f;
        if (idx >= 0) {
            const int64_t j = idx;
            const float d2 = dist2_rn(map_xyz[3 * j], map_xyz[3 * j + 1], map_xyz[3 * j + 2], x, y, z);
            if (!(d2 > map.max_d2)) v = cert[j];
        }
        best = fmaxf(best, v);
    }
    out[q] = best;
}

// ---------------------------------------------------------------------------------------------
struct FwdParams {
    MapDev map;
    pinq_decoder_view dec;
    const float* xyz;
    int64_t n;
    uint32_t flags;
    const float* features;
    float* certainties;
    const float4* quats;
    long long* ts_update;
    const int64_t* query_ts;
    float *sdf, *grad_x, *sdf_std;
    int32_t* nn_count;
    float* certainty;
    int32_t* knn_idx;
    float *knn_w, *knn_d2;
    float* feat_out;  // MODE 1: [n,K,IN]
    bool scatter_in_kernel;
    // MODE 2 (mesher grid)
    float ox, oy, oz, voxel;
    int ny, nz;
    int64_t first;
    int min_nn;
    uint8_t* mask_out;
};

// MODE 0: fused forward   MODE 1: un-fused query_feature (no MLP)   MODE 2: mesher grid (forward only)
template <bool LOCAL, bool GRAD, int MODE>
__global__ void __launch_bounds__(kBlock) k_query_fwd(const FwdParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    float4* s_w = reinterpret_cast<float4*>(smem);                     // H*16 floats
    uint32_t* s_off = reinterpret_cast<uint32_t*>(smem + H * 16 * 4);  // kn, padded to a multiple of 4
    int32_t* s_list = reinterpret_cast<int32_t*>(s_off + ((p.map.kn + 3) & ~3));
    load_stencil(p.map, s_off);
    if (MODE != 1) load_decoder(p.dec, s_w);
    __syncthreads();

    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= p.n) return;
    const bool recip = (p.flags & PINQ_F_RECIP_DIV) != 0;
    const bool rotate = (p.flags & PINQ_F_ROTATE) != 0;
    float x, y, z;
    if (MODE == 2) {
        // utils/mesher.py:168-178: int16 meshgrid -> float -> *= voxel -> += origin; z fastest
        const int64_t g = p.first + q;
        const int iz = (int)(g % p.nz);
        const int64_t r = g / p.nz;
        const int iy = (int)(r % p.ny);
        const int ix = (int)(r / p.ny);
        x = __fadd_rn(__fmul_rn((float)ix, p.voxel), p.ox);
        y = __fadd_rn(__fmul_rn((float)iy, p.voxel), p.oy);
        z = __fadd_rn(__fmul_rn((float)iz, p.voxel), p.oz);
    } else {
        x = p.xyz[3 * q];
        y = p.xyz[3 * q + 1];
        z = p.xyz[3 * q + 2];
    }

    float kd2[K];
    int32_t kg[K];
    int nn;
    search_topk<LOCAL>(p.map, s_off, s_list, x, y, z, recip, kd2, kg, nn);

    // resolve the K winners: position (for x - p_k) and the index space of the feature arrays
    int32_t kidx[K];
    float in[K][IN];
    float ux[K], uy[K], uz[K];  // un-rotated x - p_k, needed by the IDW gradient
#pragma unroll
    for (int k = 0; k < K; ++k) {
        kidx[k] = -1;
        ux[k] = uy[k] = uz[k] = 0.f;
#pragma unroll
        for (int i = 0; i < IN; ++i) in[k][i] = 0.f;
        if (kg[k] >= 0) {
            const float4 r = __ldg(&p.map.cand[kg[k]]);
            const int32_t li = LOCAL ? __float_as_int(r.w) : kg[k];
            kidx[k] = li;
            const float4* fr = reinterpret_cast<const float4*>(p.features + (int64_t)li * F);
            const float4 a = __ldg(fr), b = __ldg(fr + 1);
            in[k][0] = a.x; in[k][1] = a.y; in[k][2] = a.z; in[k][3] = a.w;
            in[k][4] = b.x; in[k][5] = b.y; in[k][6] = b.z; in[k][7] = b.w;
            float vx = __fsub_rn(x, r.x), vy = __fsub_rn(y, r.y), vz = __fsub_rn(z, r.z);
            ux[k] = vx; uy[k] = vy; uz[k] = vz;
            if (rotate) quat_rotate(__ldg(&p.quats[li]), false, vx, vy, vz);
            in[k][8] = vx; in[k][9] = vy; in[k][10] = vz;
        }
    }

    float wt[K], w[K], wsum;
    idw_weights(kd2, kidx, nn, wt, w, wsum);

    if (p.nn_count) p.nn_count[q] = nn;
    if (p.knn_idx) {
#pragma unroll
        for (int k = 0; k < K; ++k) p.knn_idx[q * K + k] = kidx[k];
    }
    if (p.knn_w) {
#pragma unroll
        for (int k = 0; k < K; ++k) p.knn_w[q * K + k] = w[k];
    }
    if (p.knn_d2) {
#pragma unroll
        for (int k = 0; k < K; ++k) p.knn_d2[q * K + k] = kd2[k];
    }
    if (p.certainty) {  // sum_k w_k c_k, neural_points.py:651-656
        float c = 0.f;
#pragma unroll
        for (int k = 0; k < K; ++k)
            if (kidx[k] >= 0) c += p.certainties[kidx[k]] * w[k];
        p.certainty[q] = c;
    }
    if (p.scatter_in_kernel) {  // training side effects, neural_points.py:637-648 (hazard H9)
        const long long ts = p.query_ts ? (long long)p.query_ts[q] : 0;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (kidx[k] >= 0) {
                atomicAdd(&p.certainties[kidx[k]], w[k]);
                if (p.query_ts && p.ts_update) atomicMax(&p.ts_update[kidx[k]], ts);
            }
        }
    }

    if (MODE == 1) {
        float* o = p.feat_out + q * (K * IN);
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int i = 0; i < IN; ++i) o[k * IN + i] = in[k][i];
        return;
    }
    if (MODE == 2) {
        if (p.mask_out) p.mask_out[q] = (uint8_t)(nn >= p.min_nn);
        if (nn == 0) {  // MLP skipped where nothing is near (mesher.py:96-103)
            p.sdf[q] = 0.f;
            return;
        }
    }

    // phase D: MLP.  acc[k] = sum_j W2[j] relu(pre_kj);  gv[k] = sum_j W2[j] 1[pre>0] W1[j][8..10]
    float acc[K];
    float gvx[K], gvy[K], gvz[K];
#pragma unroll
    for (int k = 0; k < K; ++k) acc[k] = gvx[k] = gvy[k] = gvz[k] = 0.f;
#pragma unroll 2
    for (int j = 0; j < H; ++j) {
        const float4 wa = s_w[4 * j], wb = s_w[4 * j + 1], wc = s_w[4 * j + 2], wd = s_w[4 * j + 3];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const float pre = mlp_pre(in[k], wa, wb, wc);
            acc[k] = fmaf(fmaxf(pre, 0.f), wd.x, acc[k]);
            if (GRAD) {
                const float m = pre > 0.f ? wd.x : 0.f;
                gvx[k] = fmaf(m, wc.x, gvx[k]);
                gvy[k] = fmaf(m, wc.y, gvy[k]);
                gvz[k] = fmaf(m, wc.z, gvz[k]);
            }
        }
    }
    const float b2 = __ldg(p.dec.b2), scale = p.dec.sdf_scale;
    float s[K], sdf = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        s[k] = (acc[k] + b2) * scale;  // decoder.py:80-81
        sdf = fmaf(s[k], w[k], sdf);
    }
    if (p.sdf) p.sdf[q] = sdf;
    if (p.sdf_std) {  // tracker.py:246-247
        float var = 0.f;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const float d = s[k] - sdf;
            var = fmaf(w[k], d * d, var);
        }
        p.sdf_std[q] = sqrtf(var);
    }
    if (GRAD && p.grad_x) {
        // d sdf/dx = sum_k w_k R_k^T (ds_k/dv_k) + (1/S) sum_k (s_k - sdf) d wt_k/dx,  d wt_k/dx = -2 wt_k^2 (x - p_k)
        float gx = 0.f, gy = 0.f, gz = 0.f;
        const float inv_s = 1.0f / wsum;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (kidx[k] >= 0) {
                float ax = gvx[k] * scale, ay = gvy[k] * scale, az = gvz[k] * scale;
                if (rotate) quat_rotate(__ldg(&p.quats[kidx[k]]), true, ax, ay, az);
                const float c = (s[k] - sdf) * (-2.f * wt[k] * wt[k]) * inv_s;
                gx += w[k] * ax + c * ux[k];
                gy += w[k] * ay + c * uy[k];
                gz += w[k] * az + c * uz[k];
            }
        }
        p.grad_x[3 * q] = gx;
        p.grad_x[3 * q + 1] = gy;
        p.grad_x[3 * q + 2] = gz;
    }
}

// second pass of the exact training path: scatter after every query has read its certainties
__global__ void k_certainty_scatter(const int32_t* __restrict__ knn_idx, const float* __restrict__ knn_w,
                                    const int64_t* __restrict__ query_ts, int64_t n, float* certainties,
                                    long long* ts_update) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * K) return;
    const int32_t i = knn_idx[t];
    if (i < 0) return;
    atomicAdd(&certainties[i], knn_w[t]);
    if (query_ts && ts_update) atomicMax(&ts_update[i], (long long)query_ts[t / K]);
}

static size_t fwd_smem_bytes(int kn) { return (size_t)H * 16 * 4 + (size_t)((kn + 3) & ~3) * 4 + (size_t)kn * kBlock * 4; }

template <bool LOCAL, bool GRAD, int MODE>
static int launch_fwd(const FwdParams& p, cudaStream_t st) {
    const size_t smem = fwd_smem_bytes(p.map.kn);
    auto kern = k_query_fwd<LOCAL, GRAD, MODE>;
    static size_t configured = 0;  // per instantiation
    if (smem > 48 * 1024 && smem > configured) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) {
            set_error("cudaFuncSetAttribute(%zu B smem): %s", smem, cudaGetErrorString(e));
            return PINQ_ERR_CUDA;
        }
        configured = smem;
    }
    kern<<<grid_for(p.n, kBlock), kBlock, smem, st>>>(p);
    return check_launch("k_query_fwd");
}

static int fill_params(const pinq_query_args* a, FwdParams& p, bool need_decoder) {
    int rc = check_map(a->map);
    if (rc) return rc;
    if (need_decoder) {
        rc = check_decoder(a->dec, a->k, a->feature_dim);
        if (rc) return rc;
    } else if (a->k != K || a->feature_dim != F) {
        set_error("unsupported specialisation: need query_nn_k=6, feature_dim=8");
        return PINQ_ERR_UNSUPPORTED;
    }
    if (!a->map.cand4 || !a->features) {
        set_error("null cand4/features");
        return PINQ_ERR_BAD_ARG;
    }
    if ((a->flags & PINQ_F_ROTATE) && !a->quats) {
        set_error("PINQ_F_ROTATE needs quats");
        return PINQ_ERR_BAD_ARG;
    }
    if (((a->flags & PINQ_F_TRAINING) || a->certainty) && !a->certainties) {
        set_error("certainties pointer required");
        return PINQ_ERR_BAD_ARG;
    }
    p.map = to_dev(a->map);
    p.dec = a->dec;
    p.xyz = a->xyz; p.n = a->n; p.flags = a->flags;
    p.features = a->features; p.certainties = a->certainties;
    p.quats = reinterpret_cast<const float4*>(a->quats);
    p.ts_update = reinterpret_cast<long long*>(a->ts_update);
    p.query_ts = a->query_ts;
    p.sdf = a->sdf; p.grad_x = a->grad_x; p.sdf_std = a->sdf_std; p.nn_count = a->nn_count;
    p.certainty = a->certainty; p.knn_idx = a->knn_idx; p.knn_w = a->knn_w; p.knn_d2 = a->knn_d2;
    p.feat_out = nullptr;
    // Exactness of the queried certainty in training mode needs all reads before any scatter: if the
    // caller wants it (and gave knn buffers), scatter in a second kernel; otherwise fuse the atomics.
    const bool training = (a->flags & PINQ_F_TRAINING) != 0;
    p.scatter_in_kernel = training && !(a->certainty && a->knn_idx && a->knn_w);
    p.ox = p.oy = p.oz = p.voxel = 0.f; p.ny = p.nz = 1; p.first = 0; p.min_nn = 0; p.mask_out = nullptr;
    return PINQ_OK;
}

static int post_scatter(const pinq_query_args* a, const FwdParams& p, cudaStream_t st) {
    const bool training = (a->flags & PINQ_F_TRAINING) != 0;
    if (training && !p.scatter_in_kernel && a->n > 0) {
        k_certainty_scatter<<<grid_for(a->n * K, 256), 256, 0, st>>>(a->knn_idx, a->knn_w, a->query_ts, a->n,
                                                                      a->certainties,
                                                                      reinterpret_cast<long long*>(a->ts_update));
        return check_launch("k_certainty_scatter");
    }
    return PINQ_OK;
}

}  // namespace pinq

using namespace pinq;

extern "C" {

int pinq_neighbor_search(const pinq_map_view* map, const float* xyz, int64_t n, uint32_t flags, const float* map_xyz,
                         const int64_t* ts_create, const float* travel_dist, int64_t cur_ts,
                         float diff_travel_dist_local, int time_filtering, float* out_dist2, int64_t* out_idx,
                         void* stream) {
    PINQ_REQUIRE(map && out_dist2 && out_idx, "null pointer");
    int rc = check_map(*map);
    if (rc) return rc;
    PINQ_REQUIRE(!time_filtering || (ts_create && travel_dist), "time filtering needs ts_create and travel_dist");
    if (n == 0) return PINQ_OK;
    MapDev d = to_dev(*map);
    int64_t total = n * d.kn;
    unsigned g = grid_for(total, 256);
    if (g > 148u * 64u) g = 148u * 64u;
    k_neighbor_search<<<g, 256, d.kn * sizeof(uint32_t), as_stream(stream)>>>(
        d, xyz, n, (flags & PINQ_F_RECIP_DIV) != 0, map_xyz, ts_create, travel_dist, cur_ts, diff_travel_dist_local,
        time_filtering != 0, out_dist2, out_idx);
    return check_launch("pinq_neighbor_search");
}

int pinq_query_certainty(const pinq_map_view* map, const float* xyz, int64_t n, uint32_t flags, const float* map_xyz,
                         const float* certainties, float* out, void* stream) {
    PINQ_REQUIRE(map && out, "null pointer");
    int rc = check_map(*map);
    if (rc) return rc;
    if (n == 0) return PINQ_OK;
    MapDev d = to_dev(*map);
    k_query_certainty<<<grid_for(n, 128), 128, d.kn * sizeof(uint32_t), as_stream(stream)>>>(
        d, xyz, n, (flags & PINQ_F_RECIP_DIV) != 0, map_xyz, certainties, out);
    return check_launch("pinq_query_certainty");
}

int pinq_query_fused_fwd(const pinq_query_args* a, void* stream) {
    PINQ_REQUIRE(a != nullptr, "null args");
    FwdParams p;
    int rc = fill_params(a, p, true);
    if (rc) return rc;
    if (a->n == 0) return PINQ_OK;
    cudaStream_t st = as_stream(stream);
    const bool local = (a->flags & PINQ_F_LOCAL) != 0;
    const bool grad = a->grad_x != nullptr;
    if (local) rc = grad ? launch_fwd<true, true, 0>(p, st) : launch_fwd<true, false, 0>(p, st);
    else rc = grad ? launch_fwd<false, true, 0>(p, st) : launch_fwd<false, false, 0>(p, st);
    if (rc) return rc;
    return post_scatter(a, p, st);
}

int pinq_query_feature(const pinq_query_args* a, float* feat_out, void* stream) {
    PINQ_REQUIRE(a != nullptr && feat_out != nullptr, "null args");
    FwdParams p;
    int rc = fill_params(a, p, false);
    if (rc) return rc;
    if (a->n == 0) return PINQ_OK;
    p.feat_out = feat_out;
    cudaStream_t st = as_stream(stream);
    rc = (a->flags & PINQ_F_LOCAL) ? launch_fwd<true, false, 1>(p, st) : launch_fwd<false, false, 1>(p, st);
    if (rc) return rc;
    return post_scatter(a, p, st);
}

int pinq_mesher_grid_query(const pinq_query_args* a, const float* origin3_host, const int32_t* dims3_host, float voxel,
                           int64_t first, int64_t count, int32_t min_nn, float* sdf_out, uint8_t* mask_out,
                           void* stream) {
    PINQ_REQUIRE(a && origin3_host && dims3_host && sdf_out, "null pointer");
    FwdParams p;
    int rc = fill_params(a, p, true);
    if (rc) return rc;
    PINQ_REQUIRE(!(a->flags & PINQ_F_TRAINING), "mesher query is inference only");
    const int64_t total = (int64_t)dims3_host[0] * dims3_host[1] * dims3_host[2];
    PINQ_REQUIRE(first >= 0 && count >= 0 && first + count <= total, "grid range out of bounds");
    if (count == 0) return PINQ_OK;
    p.n = count; p.xyz = nullptr;
    p.ox = origin3_host[0]; p.oy = origin3_host[1]; p.oz = origin3_host[2]; p.voxel = voxel;
    p.ny = dims3_host[1]; p.nz = dims3_host[2]; p.first = first; p.min_nn = min_nn;
    p.sdf = sdf_out; p.mask_out = mask_out; p.grad_x = nullptr; p.sdf_std = nullptr;
    p.scatter_in_kernel = false;
    cudaStream_t st = as_stream(stream);
    return (a->flags & PINQ_F_LOCAL) ? launch_fwd<true, false, 2>(p, st) : launch_fwd<false, false, 2>(p, st);
}

}  // extern "C"

// query.cu -- neighbour search (API form), neural-point SDF query forward (+ d sdf/dx, std), un-fused
// query_feature, mesher dense-grid query.
//
// The forward is two launches that share 28 B/query of workspace:
//   k_search : 81 probes of the occupancy index + candidate records, top-K in registers.  Few registers
//              -> high occupancy, which is what hides the L2 latency of the random probes.
//   k_decode : feature gather + IDW + 11->64->1 MLP + weighted combine (+ d sdf/dx, std, side effects).
//              FP32-FMA bound; the decoder lives in shared memory, activations in registers.
// Reference: model/neural_points.py:459-674, model/decoder.py:66-88, utils/tracker.py:245-253,
//            utils/mapper.py:466-468, utils/mesher.py:93-180.
#include "query_common.cuh"

namespace pinq {

// ---------------------------------------------------------------------------------------------
// Standalone API form: dist2[n,kn], idx[n,kn] (global int64).  One thread per (query, cell).
__global__ void k_neighbor_search(MapDev map, const float* __restrict__ xyz, int64_t n, bool recip,
                                  const int64_t* __restrict__ ts_create, const float* __restrict__ travel_dist,
                                  int64_t cur_ts, float diff_td, bool time_filtering, float* __restrict__ out_d2,
                                  int64_t* __restrict__ out_idx) {
    extern __shared__ uint32_t s_off_ns[];
    load_stencil(map, s_off_ns);
    __syncthreads();
    const int64_t total = n * map.kn;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t q = t / map.kn;
        const int c = (int)(t - q * map.kn);
        const float x = xyz[3 * q], y = xyz[3 * q + 1], z = xyz[3 * q + 2];
        const uint32_t slot = probe_slot(base_slot(map, x, y, z, recip), s_off_ns[c], map.B);
        const int32_t r = rank_of(map.occ[slot >> 5], slot);
        int32_t idx = -1;
        float4 rc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r >= 0) {
            rc = map.rec[r];
            idx = __float_as_int(rc.w);
        }
        if (time_filtering && map.m > 0) {
            const int64_t j = idx < 0 ? map.m - 1 : idx;  // -1 is a live wrapping index (hazard H8)
            const float dt = fabsf(__fsub_rn(travel_dist[cur_ts], travel_dist[ts_create[j]]));
            if (!(dt < diff_td)) idx = -1;
        }
        float d2 = map.max_d2;
        if (idx >= 0) {
            d2 = dist2_rn(rc.x, rc.y, rc.z, x, y, z);
            if (d2 > map.max_d2) idx = -1;  // the distance itself is kept (neural_points.py:495-499)
        }
        out_d2[t] = d2;
        out_idx[t] = (int64_t)idx;
    }
}

__global__ void k_query_certainty(MapDev map, const float* __restrict__ xyz, int64_t n, bool recip,
                                  const float* __restrict__ cert, float* __restrict__ out) {
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
        const uint32_t slot = probe_slot(base, s_off_qc[c], map.B);
        const int32_t r = rank_of(__ldg(&map.occ[slot >> 5]), slot);
        float v = 0.f;
        if (r >= 0) {
            const float4 rc = __ldg(&map.rec[r]);
            const float d2 = dist2_rn(rc.x, rc.y, rc.z, x, y, z);
            if (!(d2 > map.max_d2)) v = cert[__float_as_int(rc.w)];
        }
        best = fmaxf(best, v);
    }
    out[q] = best;
}

// ---------------------------------------------------------------------------------------------
// workspace layout: int32 ws[n][K+1] = K record ranks (ascending d2, -1 = none) + nn_count
constexpr int kWs = K + 1;

__global__ void __launch_bounds__(kBlock) k_search(MapDev map, PointSrc src, int64_t n, bool recip,
                                                   int32_t* __restrict__ ws) {
    extern __shared__ uint32_t s_off_s[];
    load_stencil(map, s_off_s);
    __syncthreads();
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    float x, y, z;
    load_point(src, q, x, y, z);
    float kd2[K];
    int32_t kr[K];
    int nn;
    search_topk(map, s_off_s, x, y, z, recip, kd2, kr, nn);
    int32_t* o = ws + q * kWs;
#pragma unroll
    for (int k = 0; k < K; ++k) o[k] = kr[k];
    o[K] = nn;
}

struct DecodeParams {
    const float4* rec;
    float max_d2;
    pinq_decoder_view dec;
    PointSrc src;
    int64_t n;
    bool rotate, scatter_in_kernel;
    const int32_t* ws;
    const float* features;
    const float* point_xyz;
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
    int min_nn;       // MODE 2
    uint8_t* mask_out;
};

// MODE 0: fused decode   MODE 1: un-fused query_feature (no MLP)   MODE 2: mesher grid (forward only)
template <bool GRAD, int MODE>
__global__ void __launch_bounds__(kBlock) k_decode(const DecodeParams p) {
    __shared__ __align__(16) float4 s_w[MODE == 1 ? 1 : H * 4];
    if (MODE != 1) {
        load_decoder(p.dec, s_w);
        __syncthreads();
    }
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= p.n) return;
    float x, y, z;
    load_point(p.src, q, x, y, z);

    const int32_t* wsq = p.ws + q * kWs;
    const int nn = wsq[K];
    int32_t kidx[K];
    float kd2[K];
    float in[K][IN];
    float ux[K], uy[K], uz[K];  // un-rotated x - p_k, needed by the IDW gradient
#pragma unroll
    for (int k = 0; k < K; ++k) {
        kidx[k] = -1;
        kd2[k] = kInvalidD2;
        ux[k] = uy[k] = uz[k] = 0.f;
#pragma unroll
        for (int i = 0; i < IN; ++i) in[k][i] = 0.f;
        const int32_t r = wsq[k];
        if (r >= 0) {
            const float4 rc = __ldg(&p.rec[r]);
            int32_t li = __float_as_int(rc.w);
            float px = rc.x, py = rc.y, pz = rc.z;
            kd2[k] = dist2_rn(rc.x, rc.y, rc.z, x, y, z);  // same ops as the search -> same bits
            if (li == -2) {  // hazard H13: candidate outside the local map -> the reference gathers local row 1
                li = 1;
                px = p.point_xyz[3]; py = p.point_xyz[4]; pz = p.point_xyz[5];
            }
            kidx[k] = li;
            const float4* fr = reinterpret_cast<const float4*>(p.features + (int64_t)li * F);
            const float4 a = __ldg(fr), b = __ldg(fr + 1);
            in[k][0] = a.x; in[k][1] = a.y; in[k][2] = a.z; in[k][3] = a.w;
            in[k][4] = b.x; in[k][5] = b.y; in[k][6] = b.z; in[k][7] = b.w;
            float vx = __fsub_rn(x, px), vy = __fsub_rn(y, py), vz = __fsub_rn(z, pz);
            // the IDW weights (and their x-gradient) come from the TRUE candidate position
            ux[k] = __fsub_rn(x, rc.x); uy[k] = __fsub_rn(y, rc.y); uz[k] = __fsub_rn(z, rc.z);
            if (p.rotate) quat_rotate(__ldg(&p.quats[li]), false, vx, vy, vz);
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

    // MLP.  acc[k] = sum_j W2[j] relu(pre_kj);  gv[k] = sum_j W2[j] 1[pre>0] W1[j][8..10]
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
                if (p.rotate) quat_rotate(__ldg(&p.quats[kidx[k]]), true, ax, ay, az);
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

static int run_query(const pinq_query_args* a, int mode, float* feat_out, const PointSrc* grid_src, int min_nn,
                     uint8_t* mask_out, cudaStream_t st) {
    int rc = check_map(a->map);
    if (rc) return rc;
    if (mode != 1) {
        rc = check_decoder(a->dec, a->k, a->feature_dim);
        if (rc) return rc;
    } else if (a->k != K || a->feature_dim != F) {
        set_error("unsupported specialisation: need query_nn_k=6, feature_dim=8");
        return PINQ_ERR_UNSUPPORTED;
    }
    if (!a->features || !a->workspace || ((a->flags & PINQ_F_LOCAL) && !a->point_xyz)) {
        set_error("null features/workspace (or point_xyz for a local query)");
        return PINQ_ERR_BAD_ARG;
    }
    if ((a->flags & PINQ_F_ROTATE) && !a->quats) {
        set_error("PINQ_F_ROTATE needs quats");
        return PINQ_ERR_BAD_ARG;
    }
    const bool training = (a->flags & PINQ_F_TRAINING) != 0;
    if ((training || a->certainty) && !a->certainties) {
        set_error("certainties pointer required");
        return PINQ_ERR_BAD_ARG;
    }
    const int64_t n = grid_src ? a->n : a->n;
    if (n == 0) return PINQ_OK;
    const MapDev map = to_dev(a->map);
    PointSrc src;
    if (grid_src) src = *grid_src;
    else {
        src.xyz = a->xyz; src.grid = false;
        src.ox = src.oy = src.oz = src.voxel = 0.f; src.ny = src.nz = 1; src.first = 0;
    }
    const bool recip = (a->flags & PINQ_F_RECIP_DIV) != 0;
    int32_t* ws = reinterpret_cast<int32_t*>(a->workspace);
    k_search<<<grid_for(n, kBlock), kBlock, map.kn * sizeof(uint32_t), st>>>(map, src, n, recip, ws);
    rc = check_launch("k_search");
    if (rc) return rc;

    DecodeParams p;
    p.rec = map.rec; p.max_d2 = map.max_d2; p.dec = a->dec; p.src = src; p.n = n;
    p.rotate = (a->flags & PINQ_F_ROTATE) != 0;
    // Exactness of the queried certainty in training mode needs all reads before any scatter: if the
    // caller wants it (and gave knn buffers), scatter in a second kernel; otherwise fuse the atomics.
    const bool two_pass = training && a->certainty && a->knn_idx && a->knn_w;
    p.scatter_in_kernel = training && !two_pass;
    p.ws = ws; p.features = a->features; p.point_xyz = a->point_xyz; p.certainties = a->certainties;
    p.quats = reinterpret_cast<const float4*>(a->quats);
    p.ts_update = reinterpret_cast<long long*>(a->ts_update); p.query_ts = a->query_ts;
    p.sdf = a->sdf; p.grad_x = a->grad_x; p.sdf_std = a->sdf_std; p.nn_count = a->nn_count;
    p.certainty = a->certainty; p.knn_idx = a->knn_idx; p.knn_w = a->knn_w; p.knn_d2 = a->knn_d2;
    p.feat_out = feat_out; p.min_nn = min_nn; p.mask_out = mask_out;
    const unsigned g = grid_for(n, kBlock);
    if (mode == 1) k_decode<false, 1><<<g, kBlock, 0, st>>>(p);
    else if (mode == 2) k_decode<false, 2><<<g, kBlock, 0, st>>>(p);
    else if (a->grad_x) k_decode<true, 0><<<g, kBlock, 0, st>>>(p);
    else k_decode<false, 0><<<g, kBlock, 0, st>>>(p);
    rc = check_launch("k_decode");
    if (rc) return rc;
    if (two_pass) {
        k_certainty_scatter<<<grid_for(n * K, 256), 256, 0, st>>>(a->knn_idx, a->knn_w, a->query_ts, n, a->certainties,
                                                                  reinterpret_cast<long long*>(a->ts_update));
        return check_launch("k_certainty_scatter");
    }
    return PINQ_OK;
}

}  // namespace pinq

using namespace pinq;

extern "C" {

int64_t pinq_query_workspace_bytes(int64_t n) { return (n > 0 ? n : 1) * (int64_t)kWs * (int64_t)sizeof(int32_t); }

int pinq_neighbor_search(const pinq_map_view* map, const float* xyz, int64_t n, uint32_t flags, const float* map_xyz,
                         const int64_t* ts_create, const float* travel_dist, int64_t cur_ts,
                         float diff_travel_dist_local, int time_filtering, float* out_dist2, int64_t* out_idx,
                         void* stream) {
    (void)map_xyz;
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
        d, xyz, n, (flags & PINQ_F_RECIP_DIV) != 0, ts_create, travel_dist, cur_ts, diff_travel_dist_local,
        time_filtering != 0, out_dist2, out_idx);
    return check_launch("pinq_neighbor_search");
}

int pinq_query_certainty(const pinq_map_view* map, const float* xyz, int64_t n, uint32_t flags, const float* map_xyz,
                         const float* certainties, float* out, void* stream) {
    (void)map_xyz;
    PINQ_REQUIRE(map && out && certainties, "null pointer");
    int rc = check_map(*map);
    if (rc) return rc;
    if (n == 0) return PINQ_OK;
    MapDev d = to_dev(*map);
    k_query_certainty<<<grid_for(n, 128), 128, d.kn * sizeof(uint32_t), as_stream(stream)>>>(
        d, xyz, n, (flags & PINQ_F_RECIP_DIV) != 0, certainties, out);
    return check_launch("pinq_query_certainty");
}

int pinq_query_fused_fwd(const pinq_query_args* a, void* stream) {
    PINQ_REQUIRE(a != nullptr, "null args");
    return run_query(a, 0, nullptr, nullptr, 0, nullptr, as_stream(stream));
}

int pinq_query_feature(const pinq_query_args* a, float* feat_out, void* stream) {
    PINQ_REQUIRE(a != nullptr && feat_out != nullptr, "null args");
    return run_query(a, 1, feat_out, nullptr, 0, nullptr, as_stream(stream));
}

int pinq_mesher_grid_query(const pinq_query_args* a, const float* origin3_host, const int32_t* dims3_host, float voxel,
                           int64_t first, int64_t count, int32_t min_nn, float* sdf_out, uint8_t* mask_out,
                           void* stream) {
    PINQ_REQUIRE(a && origin3_host && dims3_host && sdf_out, "null pointer");
    PINQ_REQUIRE(!(a->flags & PINQ_F_TRAINING), "mesher query is inference only");
    const int64_t total = (int64_t)dims3_host[0] * dims3_host[1] * dims3_host[2];
    PINQ_REQUIRE(first >= 0 && count >= 0 && first + count <= total, "grid range out of bounds");
    pinq_query_args b = *a;
    b.n = count; b.xyz = nullptr; b.sdf = sdf_out; b.grad_x = nullptr; b.sdf_std = nullptr;
    PointSrc src;
    src.xyz = nullptr; src.grid = true;
    src.ox = origin3_host[0]; src.oy = origin3_host[1]; src.oz = origin3_host[2]; src.voxel = voxel;
    src.ny = dims3_host[1]; src.nz = dims3_host[2]; src.first = first;
    return run_query(&b, 2, nullptr, &src, min_nn, mask_out, as_stream(stream));
}

}  // extern "C"

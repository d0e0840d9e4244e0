// query_common.cuh -- device building blocks shared by the forward, backward and mesher kernels.
// One thread owns one query point end to end:
//   phase A  kn (81) independent 4-byte probes of the dense slot table; non-empty hits are compacted
//            into a per-thread shared-memory list (bank == lane -> conflict free);
//   phase B  one 16-byte candidate record per hit (xyz + filtered local index), exact fp32 d^2,
//            running top-K kept sorted in registers (replaces torch.sort over [N,kn]);
//   phase C  gather K feature rows (2 x 16 B each), IDW weights;
//   phase D  11 -> 64 -> 1 MLP, decoder staged once per CTA in shared memory (broadcast LDS.128),
//            hidden activations never leave registers.
// Reference: model/neural_points.py:459-674, model/decoder.py:66-88.
#pragma once
#include "common.cuh"

namespace pinq {

constexpr int K = 6;    // query_nn_k
constexpr int F = 8;    // feature_dim
constexpr int IN = 11;  // F + 3
constexpr int H = 64;   // geo_mlp_hidden_dim
constexpr int kBlock = 128;
constexpr float kInvalidD2 = 9e3f;  // model/neural_points.py:561
constexpr float kEps = 1e-15f;      // model/neural_points.py:618

struct MapDev {
    const int32_t* table;
    uint32_t B;
    float res, inv_res, max_d2;
    int kn;
    const int32_t* stencil_dx;
    const float4* cand;
    int64_t m;
};

static inline MapDev to_dev(const pinq_map_view& v) {
    MapDev d;
    d.table = v.table;
    d.B = v.buffer_size;
    d.res = v.resolution;
    d.inv_res = 1.0f / v.resolution;
    d.max_d2 = v.max_valid_dist2;
    d.kn = v.kn;
    d.stencil_dx = v.stencil_dx;
    d.cand = reinterpret_cast<const float4*>(v.cand4);
    d.m = v.m;
    return d;
}

static inline int check_map(const pinq_map_view& v) {
    if (!v.table || !v.stencil_dx || v.kn <= 0 || v.kn > PINQ_MAX_STENCIL || v.buffer_size == 0 ||
        v.buffer_size >= 0x7fffffffu || !(v.resolution > 0.f)) {
        set_error("bad pinq_map_view (table/stencil null, kn out of range or bad buffer_size/resolution)");
        return PINQ_ERR_BAD_ARG;
    }
    return PINQ_OK;
}

// (dx . primes) mod B as uint32 -- every probe becomes add / compare / subtract instead of an int64 modulo
__device__ __forceinline__ void load_stencil(const MapDev& map, uint32_t* s_off) {
    for (int c = threadIdx.x; c < map.kn; c += blockDim.x)
        s_off[c] = slot_of_cell(map.stencil_dx[3 * c], map.stencil_dx[3 * c + 1], map.stencil_dx[3 * c + 2], map.B);
}

__device__ __forceinline__ uint32_t probe_slot(uint32_t base, uint32_t off, uint32_t B) {
    uint32_t s = base + off;
    return s >= B ? s - B : s;
}

__device__ __forceinline__ uint32_t base_slot(const MapDev& map, float x, float y, float z, bool recip) {
    return slot_of_cell(cell_of(x, map.res, map.inv_res, recip), cell_of(y, map.res, map.inv_res, recip),
                        cell_of(z, map.res, map.inv_res, recip), map.B);
}

// phase A + B.  kg[k] = GLOBAL index of the k-th nearest valid candidate (-1 = none), kd2 ascending.
// Validity (neural_points.py:480-499, 555-561): local filter (cand.w >= 0) and d2 <= max_valid_dist2.
template <bool LOCAL>
__device__ __forceinline__ void search_topk(const MapDev& map, const uint32_t* s_off, int32_t* s_list, float x, float y,
                                            float z, bool recip, float (&kd2)[K], int32_t (&kg)[K], int& nn) {
    const uint32_t base = base_slot(map, x, y, z, recip);
    const int kn = map.kn;
    int cnt = 0;
    constexpr int U = 9;  // 81 = 9 * 9: nine independent probes in flight per thread
    for (int c0 = 0; c0 < kn; c0 += U) {
        int32_t v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            int c = c0 + u;
            v[u] = (c < kn) ? __ldg(&map.table[probe_slot(base, s_off[c], map.B)]) : -1;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (v[u] >= 0) {
                s_list[cnt * blockDim.x + threadIdx.x] = v[u];
                ++cnt;
            }
        }
    }
#pragma unroll
    for (int k = 0; k < K; ++k) {
        kd2[k] = kInvalidD2;
        kg[k] = -1;
    }
    nn = 0;
    for (int j = 0; j < cnt; ++j) {
        const int32_t gi = s_list[j * blockDim.x + threadIdx.x];
        const float4 r = __ldg(&map.cand[gi]);
        const float d2 = dist2_rn(r.x, r.y, r.z, x, y, z);
        const bool ok = (!LOCAL || __float_as_int(r.w) >= 0) && !(d2 > map.max_d2);
        if (ok) {
            ++nn;
            if (d2 < kd2[K - 1]) {  // strict <: equal distances keep stencil order (stable)
                kd2[K - 1] = d2;
                kg[K - 1] = gi;
#pragma unroll
                for (int k = K - 1; k > 0; --k) {
                    if (kd2[k] < kd2[k - 1]) {
                        float td = kd2[k]; kd2[k] = kd2[k - 1]; kd2[k - 1] = td;
                        int32_t ti = kg[k]; kg[k] = kg[k - 1]; kg[k - 1] = ti;
                    }
                }
            }
        }
    }
}

// IDW weights, model/neural_points.py:618-632.  wt = un-normalised, w = normalised (0 for invalid).
__device__ __forceinline__ void idw_weights(const float (&kd2)[K], const int32_t (&kidx)[K], int nn, float (&wt)[K],
                                            float (&w)[K], float& wsum) {
    wsum = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        float t = (kidx[k] >= 0) ? __fdiv_rn(1.0f, __fadd_rn(kd2[k], kEps)) : 0.f;
        if (nn == 0) t = kEps;
        wt[k] = t;
        wsum = __fadd_rn(wsum, t);
    }
#pragma unroll
    for (int k = 0; k < K; ++k) w[k] = (kidx[k] >= 0) ? __fdiv_rn(wt[k], wsum) : 0.f;
}

// passive rotation of utils/tools.py:316-323 (conj = false) and its transpose (conj = true);
// quaternion stored (w, x, y, z) -> float4 {x=w, y=x, z=y, w=z}
__device__ __forceinline__ void quat_rotate(const float4 q, bool conj, float& vx, float& vy, float& vz) {
    const float s = conj ? 1.f : -1.f;
    const float ux = s * q.y, uy = s * q.z, uz = s * q.w, qw = q.x;
    const float tx = 2.f * (uy * vz - uz * vy), ty = 2.f * (uz * vx - ux * vz), tz = 2.f * (ux * vy - uy * vx);
    const float rx = vx + qw * tx + (uy * tz - uz * ty);
    const float ry = vy + qw * ty + (uz * tx - ux * tz);
    const float rz = vz + qw * tz + (ux * ty - uy * tx);
    vx = rx; vy = ry; vz = rz;
}

// decoder staged in shared memory, 16 floats per hidden unit j:
//   [0..7] W1[j][0..7]   [8..10] W1[j][8..10]   [11] b1[j]   [12] W2[j]   [13..15] zero
__device__ __forceinline__ void load_decoder(const pinq_decoder_view& d, float4* s_w) {
    float* s = reinterpret_cast<float*>(s_w);
    for (int t = threadIdx.x; t < H * 16; t += blockDim.x) {
        const int j = t >> 4, c = t & 15;
        float v = 0.f;
        if (c < IN) v = d.W1[j * IN + c];
        else if (c == 11) v = d.b1[j];
        else if (c == 12) v = d.W2[j];
        s[t] = v;
    }
}

__device__ __forceinline__ float mlp_pre(const float (&in)[IN], const float4 wa, const float4 wb, const float4 wc) {
    float pre = wc.w;  // b1[j]
    pre = fmaf(in[0], wa.x, pre); pre = fmaf(in[1], wa.y, pre); pre = fmaf(in[2], wa.z, pre); pre = fmaf(in[3], wa.w, pre);
    pre = fmaf(in[4], wb.x, pre); pre = fmaf(in[5], wb.y, pre); pre = fmaf(in[6], wb.z, pre); pre = fmaf(in[7], wb.w, pre);
    pre = fmaf(in[8], wc.x, pre); pre = fmaf(in[9], wc.y, pre); pre = fmaf(in[10], wc.z, pre);
    return pre;
}

static inline int check_decoder(const pinq_decoder_view& d, int k, int feature_dim) {
    if (k != K || feature_dim != F || d.hidden != H || d.in_dim != IN) {
        set_error("unsupported specialisation: need query_nn_k=6, feature_dim=8, hidden=64, in_dim=11 (got %d,%d,%d,%d)",
                  k, feature_dim, d.hidden, d.in_dim);
        return PINQ_ERR_UNSUPPORTED;
    }
    if (!d.W1 || !d.b1 || !d.W2 || !d.b2) {
        set_error("null decoder pointer");
        return PINQ_ERR_BAD_ARG;
    }
    return PINQ_OK;
}

}  // namespace pinq

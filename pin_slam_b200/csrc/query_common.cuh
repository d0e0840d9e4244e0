// query_common.cuh -- device building blocks shared by the forward, backward and mesher kernels.
// One thread owns one query point end to end:
//   search   kn (81) probes of the L2-resident occupancy index (8 B each), one 16-byte candidate record
//            per occupied slot, exact fp32 d^2, running top-K kept sorted in registers (replaces
//            torch.sort over [N,kn]); hands K record ranks + nn_count to the decode kernel;
//   decode   gather K feature rows (2 x 16 B each), IDW weights, 11 -> 64 -> 1 MLP with the decoder
//            staged once per CTA in shared memory (broadcast LDS.128); hidden activations never leave
//            registers; d sdf/dx in the same pass.
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

// Search -> decode -> backward hand-off ("stage"): one block per 32 consecutive processing positions,
// struct-of-arrays inside the block so that lane = query (resolve, epilogue) AND lane = (g,t) fragment
// accesses are coalesced.  124 B/query; written once by k_resolve, read by the decode and backward kernels.
struct StageBlock {
    int32_t idx[K][32];   // neighbour index (local or global), -1 = invalid
    float w[K][32];       // normalised IDW weight (0 for invalid)
    float v[K][3][32];    // decoder inputs 8..10: x - p_k, rotated into the point frame when after_pgo
    int32_t nn[32];       // valid candidates over the whole stencil
};
static_assert(sizeof(StageBlock) == 3968, "124 bytes per query");

struct MapDev {
    const uint2* occ;    // occupancy index: {bits, prefix} per 32 slots
    const float4* rec;   // rank-ordered candidate records {x, y, z, bits(index)}
    uint32_t B;
    float res, inv_res, max_d2;
    int kn;
    const int32_t* stencil_dx;
    int64_t m;
};

static inline MapDev to_dev(const pinq_map_view& v) {
    MapDev d;
    d.occ = reinterpret_cast<const uint2*>(v.occ);
    d.rec = reinterpret_cast<const float4*>(v.rec4);
    d.B = v.buffer_size;
    d.res = v.resolution;
    d.inv_res = 1.0f / v.resolution;
    d.max_d2 = v.max_valid_dist2;
    d.kn = v.kn;
    d.stencil_dx = v.stencil_dx;
    d.m = v.m;
    return d;
}

static inline int check_map(const pinq_map_view& v) {
    if (!v.occ || !v.rec4 || !v.stencil_dx || v.kn <= 0 || v.kn > PINQ_MAX_STENCIL || v.buffer_size == 0 ||
        v.buffer_size >= 0x7fffffffu || !(v.resolution > 0.f)) {
        set_error("bad pinq_map_view (occ/rec/stencil null, kn out of range or bad buffer_size/resolution)");
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

// table[slot] through the occupancy index: rank of the slot's record, or -1 when the slot is empty
__device__ __forceinline__ int32_t rank_of(const uint2 w, uint32_t slot) {
    const uint32_t bit = slot & 31u;
    return ((w.x >> bit) & 1u) ? (int32_t)(w.y + __popc(w.x & ((1u << bit) - 1u))) : -1;
}

// where the query point comes from: an [n,3] array, or generated in-kernel for the mesher grid
// (utils/mesher.py:168-178: int16 meshgrid -> float -> *= voxel -> += origin; z fastest)
struct PointSrc {
    const float* xyz;
    float ox, oy, oz, voxel;
    int ny, nz;
    int64_t first;
    bool grid;
};

__device__ __forceinline__ void load_point(const PointSrc& s, int64_t q, float& x, float& y, float& z) {
    if (s.grid) {
        const int64_t g = s.first + q;
        const int iz = (int)(g % s.nz);
        const int64_t r = g / s.nz;
        const int iy = (int)(r % s.ny);
        const int ix = (int)(r / s.ny);
        x = __fadd_rn(__fmul_rn((float)ix, s.voxel), s.ox);
        y = __fadd_rn(__fmul_rn((float)iy, s.voxel), s.oy);
        z = __fadd_rn(__fmul_rn((float)iz, s.voxel), s.oz);
    } else {
        x = s.xyz[3 * q];
        y = s.xyz[3 * q + 1];
        z = s.xyz[3 * q + 2];
    }
}

// Neighbour search of one query: kr[k] = record rank of the k-th nearest valid candidate (-1 = none),
// kd2 ascending, nn = valid candidates over the whole stencil.  Probes go out in batches of U independent
// 8-byte index loads followed by U predicated 16-byte record loads (memory-level parallelism without a
// shared-memory hit list, so occupancy is bounded by registers only).
// Validity (neural_points.py:480-499, 555-561): index != -1 (local filter baked into rec.w) and d2 <= max.
__device__ __forceinline__ void search_topk(const MapDev& map, const uint32_t* s_off, float x, float y, float z,
                                            bool recip, float (&kd2)[K], int32_t (&kr)[K], int& nn) {
    const uint32_t base = base_slot(map, x, y, z, recip);
    const int kn = map.kn;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        kd2[k] = kInvalidD2;
        kr[k] = -1;
    }
    nn = 0;
    constexpr int U = 9;  // 81 = 9 * 9
    for (int c0 = 0; c0 < kn; c0 += U) {
        int32_t r[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int c = c0 + u;
            r[u] = -1;
            if (c < kn) {
                const uint32_t slot = probe_slot(base, s_off[c], map.B);
                r[u] = rank_of(__ldg(&map.occ[slot >> 5]), slot);
            }
        }
        float4 rc[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (r[u] >= 0) rc[u] = __ldg(&map.rec[r[u]]);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (r[u] < 0) continue;
            const float d2 = dist2_rn(rc[u].x, rc[u].y, rc[u].z, x, y, z);
            if (__float_as_int(rc[u].w) == -1 || d2 > map.max_d2) continue;  // -2 = hazard H13, still a candidate
            ++nn;
            if (d2 < kd2[K - 1]) {  // strict <: equal distances keep stencil order (stable)
                kd2[K - 1] = d2;
                kr[K - 1] = r[u];
#pragma unroll
                for (int k = K - 1; k > 0; --k) {
                    if (kd2[k] < kd2[k - 1]) {
                        const float td = kd2[k]; kd2[k] = kd2[k - 1]; kd2[k - 1] = td;
                        const int32_t ti = kr[k]; kr[k] = kr[k - 1]; kr[k - 1] = ti;
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

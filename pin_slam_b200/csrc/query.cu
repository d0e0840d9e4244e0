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
#include <stdlib.h>

#include "brick.cuh"
#include "mma.cuh"

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
                                                   const int32_t* __restrict__ perm, int32_t* __restrict__ ws,
                                                   const int32_t* __restrict__ skip) {
    if (skip && *skip) return;
    extern __shared__ uint32_t s_off_s[];
    load_stencil(map, s_off_s);
    __syncthreads();
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // position in processing order
    if (i >= n) return;
    float x, y, z;
    load_point(src, perm ? perm[i] : i, x, y, z);
    float kd2[K];
    int32_t kr[K];
    int nn;
    search_topk(map, s_off_s, x, y, z, recip, kd2, kr, nn);
    int32_t* o = ws + i * kWs;
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
    const int32_t* perm;  // processing order -> query index (spatial binning, csrc/sort.cu); NULL = identity.
                          // ws / stage / relu_mask are indexed by processing position, everything else by query.
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
    StageBlock* stage;    // [ceil(n/32)] search -> decode -> backward hand-off (processing order)
    uint32_t* relu_mask;  // [ceil(n/32)][12][32] bit masks of relu'(pre), fragment order (consumed by the backward)
    float* feat_out;      // un-fused query_feature: [n,K,IN]
    int min_nn;           // mesher
    uint8_t* mask_out;
    const int32_t* skip;  // device flag: non-zero = this launch is a no-op (device-side Gauss-Newton loop, gn.cu)
    float4* zero_fill = nullptr;  // pinq_query_args.zero_fill: thread i of the brick search clears float4 i of it
    int64_t zero_fill4 = 0;
};

// Per-query part shared by the fused decode and the un-fused query_feature: resolve the K record ranks
// to (index, d^2, x - p_k), IDW weights, API outputs and the training-mode side effects.
struct QueryLane {
    int32_t kidx[K];
    float kd2[K], w[K], wt[K], wsum;
    float vx[K], vy[K], vz[K];  // decoder inputs 8..10 (rotated into the point frame when after_pgo)
    float ux[K], uy[K], uz[K];  // un-rotated x - p_k of the TRUE candidate position (IDW gradient)
    int nn;
};

// EFFECTS: write the per-query API outputs and apply the training-mode side effects (done exactly once per query:
// by k_resolve / k_query_feature; the gradient flavour of the decode re-derives the lane state with EFFECTS = false)
template <bool WANT_U, bool EFFECTS, bool STORE_KNN>
__device__ __forceinline__ void resolve_ranks(const DecodeParams& p, const int32_t (&r)[K], int nn, int64_t q, bool live,
                                              float x, float y, float z, QueryLane& L) {
    L.nn = 0;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        L.kidx[k] = -1;
        L.kd2[k] = kInvalidD2;
        L.vx[k] = L.vy[k] = L.vz[k] = 0.f;
        if (WANT_U) L.ux[k] = L.uy[k] = L.uz[k] = 0.f;
    }
    if (live) {
        L.nn = nn;
        float4 rc[K];
#pragma unroll
        for (int k = 0; k < K; ++k)
            if (r[k] >= 0) rc[k] = __ldg(&p.rec[r[k]]);
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (r[k] < 0) continue;
            int32_t li = __float_as_int(rc[k].w);
            float px = rc[k].x, py = rc[k].y, pz = rc[k].z;
            L.kd2[k] = dist2_rn(rc[k].x, rc[k].y, rc[k].z, x, y, z);  // same ops as the search -> same bits
            if (li == -2) {  // hazard H13: candidate outside the local map -> the reference gathers local row 1
                li = 1;
                px = p.point_xyz[3]; py = p.point_xyz[4]; pz = p.point_xyz[5];
            }
            L.kidx[k] = li;
            float vx = __fsub_rn(x, px), vy = __fsub_rn(y, py), vz = __fsub_rn(z, pz);
            if (WANT_U) {  // the IDW weights (and their x-gradient) come from the TRUE candidate position
                L.ux[k] = __fsub_rn(x, rc[k].x); L.uy[k] = __fsub_rn(y, rc[k].y); L.uz[k] = __fsub_rn(z, rc[k].z);
            }
            if (p.rotate) quat_rotate(__ldg(&p.quats[li]), false, vx, vy, vz);
            L.vx[k] = vx; L.vy[k] = vy; L.vz[k] = vz;
        }
    }
    idw_weights(L.kd2, L.kidx, L.nn, L.wt, L.w, L.wsum);
    if (!live || !EFFECTS) return;
    if (p.nn_count) p.nn_count[q] = L.nn;
    if (STORE_KNN && p.knn_idx) {
#pragma unroll
        for (int k = 0; k < K; ++k) p.knn_idx[q * K + k] = L.kidx[k];
    }
    if (STORE_KNN && p.knn_w) {
#pragma unroll
        for (int k = 0; k < K; ++k) p.knn_w[q * K + k] = L.w[k];
    }
    if (p.knn_d2) {
#pragma unroll
        for (int k = 0; k < K; ++k) p.knn_d2[q * K + k] = L.kd2[k];
    }
    if (p.certainty) {  // sum_k w_k c_k, neural_points.py:651-656
        float c = 0.f;
#pragma unroll
        for (int k = 0; k < K; ++k)
            if (L.kidx[k] >= 0) c += p.certainties[L.kidx[k]] * L.w[k];
        p.certainty[q] = c;
    }
    if (p.scatter_in_kernel) {  // training side effects, neural_points.py:637-648 (hazard H9)
        const long long ts = p.query_ts ? (long long)p.query_ts[q] : 0;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (L.kidx[k] >= 0) {
                atomicAdd(&p.certainties[L.kidx[k]], L.w[k]);
                if (p.query_ts && p.ts_update) atomicMax(&p.ts_update[L.kidx[k]], ts);
            }
        }
    }
}

// the same from the search workspace (K record ranks + count of query position i)
template <bool WANT_U, bool EFFECTS, bool STORE_KNN>
__device__ __forceinline__ void resolve_query(const DecodeParams& p, const int32_t* wsq, int64_t i, int64_t q, bool live,
                                              float x, float y, float z, QueryLane& L) {
    (void)i;
    int32_t r[K];
    int nn = 0;
#pragma unroll
    for (int k = 0; k < K; ++k) r[k] = live ? wsq[k] : -1;
    if (live) nn = wsq[K];
    resolve_ranks<WANT_U, EFFECTS, STORE_KNN>(p, r, nn, q, live, x, y, z, L);
}

// lane = query: the lane's column of its warp's stage block (the hand-off to the decode and the backward)
__device__ __forceinline__ void stage_store(const DecodeParams& p, int64_t i, int lane, const QueryLane& L) {
    StageBlock& sb = p.stage[i >> 5];
#pragma unroll
    for (int k = 0; k < K; ++k) {
        sb.idx[k][lane] = L.kidx[k];
        sb.w[k][lane] = L.w[k];
        sb.v[k][0][lane] = L.vx[k];
        sb.v[k][1][lane] = L.vy[k];
        sb.v[k][2][lane] = L.vz[k];
    }
    sb.nn[lane] = L.nn;
}

// ---------------------------------------------------------------------------------------------
// Neighbour search through the brick index (csrc/brick.cuh): the stencil of a query touches 2 x 2 x 2 bricks;
// per brick one probe, then only the OCCUPIED voxels that belong to the stencil are visited (bit scan of
// occupancy & stencil mask).  Same candidates, same exact d^2, same (d^2, stencil position) order as k_search.
//
// Stencil masks: for a query whose stencil box starts at voxel (sx, sy, sz) of brick b0 (s = (cell - R) & 3)
// and for brick b0 + (ix, iy, iz), the 64-bit set of brick voxels that are stencil members.  64 x 8 masks,
// built once per launch (k_stencil_masks) and staged in shared memory.
constexpr int kStencilMasks = 64 * 8;
constexpr uint32_t kRankBits = 24;  // low key word = stencil position << 24 | record rank (host guards m < 2^24)

__global__ void k_stencil_masks(const int32_t* __restrict__ stencil_dx, int kn, int R, unsigned long long* __restrict__ out) {
    __shared__ uint8_t member[128];
    const int W = 2 * R + 1;
    for (int c = threadIdx.x; c < 128; c += blockDim.x) member[c] = 0;
    __syncthreads();
    for (int c = threadIdx.x; c < kn; c += blockDim.x)
        member[((stencil_dx[3 * c] + R) * W + (stencil_dx[3 * c + 1] + R)) * W + (stencil_dx[3 * c + 2] + R)] = 1;
    __syncthreads();
    for (int e = threadIdx.x; e < kStencilMasks; e += blockDim.x) {
        const int sx = (e >> 7) & 3, sy = (e >> 5) & 3, sz = (e >> 3) & 3, ix = (e >> 2) & 1, iy = (e >> 1) & 1, iz = e & 1;
        unsigned long long m = 0ull;
        for (int bit = 0; bit < 64; ++bit) {
            // offset of brick voxel (lx,ly,lz) from the box origin, per axis: 4 i + l - s  in [0, 2R] for members
            const int ox = 4 * ix + (bit >> 4) - sx, oy = 4 * iy + ((bit >> 2) & 3) - sy, oz = 4 * iz + (bit & 3) - sz;
            if (ox < 0 || oy < 0 || oz < 0 || ox >= W || oy >= W || oz >= W) continue;
            if (member[(ox * W + oy) * W + oz]) m |= 1ull << bit;
        }
        out[e] = m;
    }
}

// sorted insertion of a 64-bit key (d^2 bits : stencil position : rank) into the ascending top-K: K compares,
// each shared by two neighbouring slots
__device__ __forceinline__ void topk_insert64(unsigned long long key, unsigned long long (&kk)[K]) {
    bool lt[K];
#pragma unroll
    for (int k = 0; k < K; ++k) lt[k] = key < kk[k];
#pragma unroll
    for (int k = K - 1; k > 0; --k) kk[k] = lt[k - 1] ? kk[k - 1] : (lt[k] ? key : kk[k]);
    kk[0] = lt[0] ? key : kk[0];
}

// Per-lane walk list staged in shared memory ([entry][thread], conflict-free).  An entry is one non-empty HALF brick
// (x-slab of 2 x 4 x 4 voxels = 32 bits, so the walk is all 32-bit integer work); a (2R+1)^3 stencil with R <= 2
// touches at most 3 x 2 x 2 of them.
constexpr int kWalkEntries = 12;
constexpr uint32_t kWalkIdMask = 0x1FFu;  // entry (4 bits) << 5 | voxel bit (5 bits): identity of a candidate within a lane's walk
#ifndef PINQ_WALK_SEL
#define PINQ_WALK_SEL 0
#endif
struct BrickLane {
#if PINQ_WALK_SEL
    uint32_t sel[kWalkEntries][kBlock];
#endif
    uint32_t msk[kWalkEntries][kBlock];  // occupancy of the half brick
    uint32_t pk[kWalkEntries][kBlock];   // rank of its first record | (2 * brick + half) << 24
};
// (the stencil members still to visit are NOT stored: `msk & half-brick stencil mask` is recomputed when the walk reaches the
// entry -- two words per entry instead of three keep the kernel's shared memory at 17 KB per CTA, i.e. 90 instead of 56 KB of
// L1 for the candidate records next to 8 resident CTAs)
constexpr size_t kBrickSmem = sizeof(BrickLane) + kStencilMasks * 8 + 128 + 32 + 64;

// RESOLVE: the walk ends with the lane's K winners in registers and their records hot in L1, so the resolve step
// (stage block, per-query outputs, training side effects -- what k_resolve does after the slot-probing search)
// runs in the tail of this kernel; the rank workspace is then only written for the gradient flavour of the decode.
#ifndef PINQ_SEARCH_MINB
#define PINQ_SEARCH_MINB 9
#endif
// DTEST = false: the caller has proved that every stencil cell lies within max_valid_dist2 of every point of the query's own cell
// (PINQ_F_STENCIL_IN_RANGE), so `d2 > max` never fires and the per-candidate comparison is compiled out.
template <bool RESOLVE, bool DTEST>
__global__ void __launch_bounds__(kBlock, PINQ_SEARCH_MINB) k_search_brick(MapDev map, BrickDev bd, PointSrc src, int64_t n, bool recip,
                                                         const int32_t* __restrict__ perm,
                                                         const unsigned long long* __restrict__ stencil_masks,
                                                         int32_t* __restrict__ ws, const DecodeParams p) {
    if (p.skip && *p.skip) return;
    {   // pinq_query_args.zero_fill: the gradient block of the backward that follows, cleared with one streaming store per thread
        const int64_t zi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
        if (zi < p.zero_fill4) __stcs(p.zero_fill + zi, make_float4(0.f, 0.f, 0.f, 0.f));
    }
    extern __shared__ __align__(16) unsigned char s_brick_raw[];
    BrickLane* bl = reinterpret_cast<BrickLane*>(s_brick_raw);
    // stencil masks staged as 32-bit HALF-brick masks, [2 b + h][s] with s = box origin inside its brick (0..63): a lane reads
    // word s of a row, so lanes with different origins hit different banks (the [s][b] 64-bit rows of the global table put
    // every lane on one of two bank groups: ncu r4a counted 17 wavefronts per LDS.64, 64 % of the kernel's shared wavefronts)
    uint32_t* s_hm = reinterpret_cast<uint32_t*>(s_brick_raw + sizeof(BrickLane));
    uint8_t* s_sid = reinterpret_cast<uint8_t*>(s_hm + 2 * kStencilMasks);  // stencil position of offset (dx,dy,dz)
    uint8_t* s_bitoff = s_sid + 128;  // linear stencil offset of voxel `bit` of a half brick relative to its voxel 0
    uint8_t* s_need = s_bitoff + 32;  // per origin s: bit b = brick b holds stencil members
    const int R = bd.radius, W = 2 * R + 1, tid = threadIdx.x;
    for (int c = tid; c < kStencilMasks; c += blockDim.x) {  // c = b * 64 + s: consecutive lanes write consecutive words
        const int b = c >> 6, so = c & 63;
        const unsigned long long m = stencil_masks[so * 8 + b];
        s_hm[(2 * b) * 64 + so] = (uint32_t)m;
        s_hm[(2 * b + 1) * 64 + so] = (uint32_t)(m >> 32);
    }
    for (int c = tid; c < 64; c += blockDim.x) {
        uint32_t nb = 0u;
#pragma unroll
        for (int b = 0; b < 8; ++b) nb |= (stencil_masks[c * 8 + b] != 0ull ? 1u : 0u) << b;
        s_need[c] = (uint8_t)nb;
    }
    for (int c = tid; c < 128; c += blockDim.x) s_sid[c] = 0xFF;
    for (int c = tid; c < 32; c += blockDim.x) s_bitoff[c] = (uint8_t)(((c >> 4) * W + ((c >> 2) & 3)) * W + (c & 3));
    __syncthreads();
    for (int c = tid; c < map.kn; c += blockDim.x) {
        const int dx = map.stencil_dx[3 * c], dy = map.stencil_dx[3 * c + 1], dz = map.stencil_dx[3 * c + 2];
        s_sid[((dx + R) * W + (dy + R)) * W + (dz + R)] = (uint8_t)c;
    }
    __syncthreads();

    const int64_t i = (int64_t)blockIdx.x * blockDim.x + tid;
    const bool live = i < n;
    const int64_t q = (live && perm) ? (int64_t)perm[i] : i;
    float x = 0.f, y = 0.f, z = 0.f;
    if (live) load_point(src, q, x, y, z);
    const int cx = cell_of(x, map.res, map.inv_res, recip), cy = cell_of(y, map.res, map.inv_res, recip),
              cz = cell_of(z, map.res, map.inv_res, recip);
    const int ax = cx - R, ay = cy - R, az = cz - R;  // box origin
    const int b0x = ax >> 2, b0y = ay >> 2, b0z = az >> 2;
    const int so = ((ax & 3) * 4 + (ay & 3)) * 4 + (az & 3);  // box origin inside brick b0

    // phase 1: one probe per brick, in batches of PINQ_PROBE_BATCH: the 32-byte entries of a batch are requested back to
    // back (one 256-bit load each) before the first key is compared, so a batch costs ONE exposed L2 latency instead of
    // chained ones (ncu r3b: the eight sequential key waits were 11 % of the kernel's stall samples); batches of two keep
    // the kernel at 64 registers = 8 CTAs per SM (batches of four: 80 registers, slower).  An entry that is not the
    // brick's own (linear-probing displacement, rare at load factor <= 0.5) falls back to the probe loop
    int total = 0, cnt = 0;
    const uint32_t needbits = live ? (uint32_t)s_need[so] : 0u;
#ifndef PINQ_PROBE_BATCH
#define PINQ_PROBE_BATCH 2
#endif
#ifndef PINQ_PROBE_UNROLL
#define PINQ_PROBE_UNROLL 1
#endif
    // A brick's entry goes straight into the walk list (no per-brick register arrays: the loop over the batches is a real
    // loop, which keeps both the register count and the code size of this phase down)
#define PINQ_PRAGMA_(x) _Pragma(#x)
#define PINQ_UNROLL_(n) PINQ_PRAGMA_(unroll n)
    PINQ_UNROLL_(PINQ_PROBE_UNROLL)
    for (int b0 = 0; b0 < 8; b0 += PINQ_PROBE_BATCH) {
        unsigned long long want[PINQ_PROBE_BATCH];
        uint32_t ew[PINQ_PROBE_BATCH][8];
        bool need[PINQ_PROBE_BATCH];
#pragma unroll
        for (int u = 0; u < PINQ_PROBE_BATCH; ++u) {
            const int b = b0 + u;
            need[u] = (needbits >> b) & 1u;
            want[u] = brick_key(b0x + (b >> 2), b0y + ((b >> 1) & 1), b0z + (b & 1));
#pragma unroll
            for (int w = 0; w < 8; ++w) ew[u][w] = 0u;
            if (need[u]) ldg_entry(&bd.tab[brick_hash(want[u], bd.cap_mask)], ew[u]);
        }
#pragma unroll
        for (int u = 0; u < PINQ_PROBE_BATCH; ++u) {
            const int b = b0 + u;
            unsigned long long pmb = 0ull;
            uint32_t pbb = 0u;
            if (need[u]) {
                const unsigned long long got = ((unsigned long long)ew[u][5] << 32) | ew[u][4];
                if (got == want[u]) {
                    pmb = ((unsigned long long)ew[u][1] << 32) | ew[u][0];
                    pbb = ew[u][2];
                } else if (got != kBrickEmpty) {  // displaced entry: walk the probe sequence
                    const int e = brick_find(bd.tab, bd.cap_mask, want[u]);
                    if (e >= 0) {
                        const uint4 v = __ldg(reinterpret_cast<const uint4*>(&bd.tab[e]));
                        pmb = ((unsigned long long)v.y << 32) | v.x;
                        pbb = v.z;
                    }
                }
            }
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const uint32_t occ32 = (uint32_t)(pmb >> (32 * h));
                const uint32_t sel32 = occ32 & s_hm[(2 * b + h) * 64 + so];
                if (sel32) {
                    const uint32_t base = pbb + (h ? (uint32_t)__popc((uint32_t)pmb) : 0u);
#if PINQ_WALK_SEL
                    bl->sel[cnt][tid] = sel32;
#endif
                    bl->msk[cnt][tid] = occ32;
                    bl->pk[cnt][tid] = base | ((uint32_t)(2 * b + h) << kRankBits);
                    ++cnt;
                    total += __popc(sel32);
                }
            }
        }
    }

    // phase 2: ONE loop over the lane's candidates of all its half bricks (trip count = max over the warp of `total`,
    // bodies predicated so the warp stays converged), software-pipelined: the record of the next candidate is in
    // flight while the current one is ranked.
    //
    // Fast pass: 32-bit keys = top 23 bits of d^2 (sign-free fp32 order = unsigned order) | walk identity (entry << 5 | bit),
    // sorted top-(K+1) by min/max pairs (2 instructions per slot instead of 7 for the 64-bit key).  A key with fewer d^2 bits
    // orders two candidates correctly unless their truncated d^2 are EQUAL; if no two neighbours of the final sorted top-(K+1)
    // share their 23 bits, every candidate outside it has a strictly larger d^2 than slot K-1 and the K winners and their
    // order are those of the exact (d^2, stencil position) sort.  Otherwise (about 1 query in 10^4 on the bench map, exact
    // ties included) the lane repeats the walk with the full 64-bit keys.
    const int total0 = total;
    uint32_t k7[K + 1];
#pragma unroll
    for (int k = 0; k <= K; ++k) k7[k] = ~0u;
    int nn = 0, e = -1;
    uint32_t m = 0u, msk = 0u, bas = 0u;
    bool nact;
    float4 nrc;
    uint32_t nid;  // walk identity of the candidate in flight
    auto advance32 = [&]() {
        nact = total > 0;
        if (nact && m == 0u) {  // next half brick of the list
            ++e;
            msk = bl->msk[e][tid];
            const uint32_t pk = bl->pk[e][tid];
            bas = pk & ((1u << kRankBits) - 1u);
#if PINQ_WALK_SEL
            m = bl->sel[e][tid];
#else
            m = msk & s_hm[(pk >> kRankBits) * 64 + so];
#endif
        }
        --total;
        const int bit = (__ffs((int)m) - 1) & 31;
        m &= m - 1u;  // 0 stays 0
        const uint32_t rank = bas + (uint32_t)__popc(msk & ((1u << bit) - 1u));
        nrc = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
        if (nact) nrc = __ldg(&bd.rec[rank]);
        nid = ((uint32_t)e << 5) | (uint32_t)bit;
    };
    // U candidates per trip: their records are requested back to back and ranked afterwards, so the loads of a trip
    // overlap each other (with one record in flight the copy of the loaded registers at the loop end waited for a load
    // issued ~45 instructions earlier: 20 % of the kernel's stall samples sat on that one instruction, ncu r3b)
#ifndef PINQ_SEARCH_BATCH
#define PINQ_SEARCH_BATCH 3
#endif
    while (__any_sync(0xffffffffu, total > 0)) {
        float4 rcs[PINQ_SEARCH_BATCH];
        uint32_t ids[PINQ_SEARCH_BATCH];
        bool acts[PINQ_SEARCH_BATCH];
#pragma unroll
        for (int u = 0; u < PINQ_SEARCH_BATCH; ++u) {
            advance32();
            rcs[u] = nrc;
            ids[u] = nid;
            acts[u] = nact;
        }
#pragma unroll
        for (int u = 0; u < PINQ_SEARCH_BATCH; ++u) {
            const float d2 = dist2_rn(rcs[u].x, rcs[u].y, rcs[u].z, x, y, z);
            // valid: index != -1 (-2 = hazard H13, still a candidate) and not d2 > max (neural_points.py:499)
            const bool ok = acts[u] && __float_as_int(rcs[u].w) != -1 && (!DTEST || !(d2 > map.max_d2));
            nn += ok ? 1 : 0;
            uint32_t key = ok ? ((__float_as_uint(d2) & ~kWalkIdMask) | ids[u]) : ~0u;
#ifndef PINQ_SEARCH_SERIAL_INSERT
            // ascending insertion, every slot from the OLD list: new[k] = min(old[k], max(old[k-1], key)) -- the same 2
            // instructions per slot as a bubble pass, but no slot waits for its neighbour (dependency depth 2 instead of 2 K)
            uint32_t nk[K + 1];
            nk[0] = min(k7[0], key);
#pragma unroll
            for (int k = 1; k <= K; ++k) nk[k] = min(k7[k], max(k7[k - 1], key));
#pragma unroll
            for (int k = 0; k <= K; ++k) k7[k] = nk[k];
#else
#pragma unroll
            for (int k = 0; k <= K; ++k) {  // ascending insertion: slot keeps the smaller, the larger moves on
                const uint32_t lo = min(k7[k], key);
                key = max(k7[k], key);
                k7[k] = lo;
            }
#endif
        }
    }
    bool tie = false;
#pragma unroll
    for (int k = 0; k < K; ++k) tie |= k7[k + 1] != ~0u && ((k7[k] ^ k7[k + 1]) <= kWalkIdMask);
    tie |= bd.exact_keys;
    int32_t r[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
        r[k] = -1;
        if (k7[k] != ~0u) {
            const uint32_t we = (k7[k] & kWalkIdMask) >> 5, wb = k7[k] & 31u;
            r[k] = (int32_t)((bl->pk[we][tid] & ((1u << kRankBits) - 1u)) + (uint32_t)__popc(bl->msk[we][tid] & ((1u << wb) - 1u)));
        }
    }
    if (__any_sync(0xffffffffu, tie)) {  // exact pass for the lanes with a truncated-key tie
        unsigned long long kk[K];
#pragma unroll
        for (int k = 0; k < K; ++k) kk[k] = ~0ull;
        total = tie ? total0 : 0;
        e = -1;
        m = 0u;
        int lin0 = 0;
        uint32_t nlow;  // stencil position << 24 | record rank of the candidate in flight
        auto advance = [&]() {
            nact = total > 0;
            if (nact && m == 0u) {
                ++e;
                msk = bl->msk[e][tid];
                const uint32_t pk = bl->pk[e][tid];
                m = msk & s_hm[(pk >> kRankBits) * 64 + so];
                bas = pk & ((1u << kRankBits) - 1u);
                // stencil offset of the half brick's voxel 0 from the box origin, linearised like s_sid
                const int bh = (int)(pk >> kRankBits), b = bh >> 1;
                lin0 = ((4 * (b >> 2) - (ax & 3)) * W + (4 * ((b >> 1) & 1) - (ay & 3))) * W + (4 * (b & 1) - (az & 3)) +
                       (bh & 1) * 2 * W * W;
            }
            --total;
            const int bit = (__ffs((int)m) - 1) & 31;
            m &= m - 1u;
            const uint32_t rank = bas + (uint32_t)__popc(msk & ((1u << bit) - 1u));
            nrc = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
            if (nact) nrc = __ldg(&bd.rec[rank]);
            nlow = ((uint32_t)s_sid[(lin0 + s_bitoff[bit]) & 127] << kRankBits) | rank;
        };
        advance();
        while (__any_sync(0xffffffffu, nact)) {
            const bool act = nact;
            const float4 rc = nrc;
            const uint32_t low = nlow;
            advance();
            const float d2 = dist2_rn(rc.x, rc.y, rc.z, x, y, z);
            const bool ok = act && __float_as_int(rc.w) != -1 && (!DTEST || !(d2 > map.max_d2));
            const unsigned long long key = ok ? ((unsigned long long)__float_as_uint(d2) << 32) | low : ~0ull;
            topk_insert64(key, kk);
        }
        if (tie) {
#pragma unroll
            for (int k = 0; k < K; ++k) r[k] = kk[k] == ~0ull ? -1 : (int32_t)((uint32_t)kk[k] & ((1u << kRankBits) - 1u));
        }
    }
    if (live && (!RESOLVE || p.grad_x)) {
        int32_t* o = ws + i * kWs;
#pragma unroll
        for (int k = 0; k < K; ++k) o[k] = r[k];
        o[K] = nn;
    }
    if (RESOLVE) {
        QueryLane L;
        resolve_ranks<false, true, true>(p, r, nn, q, live, x, y, z, L);
        if (i - (tid & 31) < n) stage_store(p, i, tid & 31, L);
    }
}

// Un-fused API form (model/neural_points.py:574-632): writes the [n,K,F+3] feature tensor, no MLP.
__global__ void __launch_bounds__(kBlock) k_query_feature(const DecodeParams p) {
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= p.n) return;
    float x, y, z;
    load_point(p.src, q, x, y, z);
    QueryLane L;
    resolve_query<false, true, true>(p, p.ws + q * kWs, q, q, true, x, y, z, L);
    float* o = p.feat_out + q * (K * IN);
#pragma unroll
    for (int k = 0; k < K; ++k) {
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
        if (L.kidx[k] >= 0) {
            const float4* fr = reinterpret_cast<const float4*>(p.features + (int64_t)L.kidx[k] * F);
            a = __ldg(fr);
            b = __ldg(fr + 1);
        }
        float* ok = o + k * IN;
        ok[0] = a.x; ok[1] = a.y; ok[2] = a.z; ok[3] = a.w;
        ok[4] = b.x; ok[5] = b.y; ok[6] = b.z; ok[7] = b.w;
        ok[8] = L.vx[k]; ok[9] = L.vy[k]; ok[10] = L.vz[k];
    }
}

// ---------------------------------------------------------------------------------------------
// k_resolve: lane = query.  Turns the K record ranks of the search into the stage block (index, IDW weight,
// x - p_k), writes the per-query API outputs and applies the training-mode side effects.  A light kernel
// (few registers, many resident warps) so that its dependent gathers are hidden by occupancy instead of
// stalling the tensor-core kernel (ncu r1l: that part ran at IPC 0.2 inside the decode).
__global__ void __launch_bounds__(kBlock) k_resolve(const DecodeParams p) {
    if (p.skip && *p.skip) return;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // processing position
    const int lane = threadIdx.x & 31;
    const bool live = i < p.n;
    const int64_t q = (live && p.perm) ? (int64_t)p.perm[i] : i;
    float x = 0.f, y = 0.f, z = 0.f;
    if (live) load_point(p.src, q, x, y, z);
    QueryLane L;
    resolve_query<false, true, true>(p, p.ws + i * kWs, i, q, live, x, y, z, L);
    if (i - lane >= p.n) return;  // whole warp past the end
    stage_store(p, i, lane, L);
}

// ---------------------------------------------------------------------------------------------
// Fused decode on the tensor cores.  One warp owns 32 queries = 192 (query, neighbour) rows = 12 m-tiles:
// m-tile mt = 2 k + half holds neighbour k of queries 16 half + {g, g + 8}.
//   lane = (g,t) : A fragments straight from the stage block and the feature table (each quad reads one
//                  32-byte row), software-pipelined one m-tile ahead; 3xTF32 layer 1 with the B fragments
//                  resident in registers; relu . W2 in the accumulator layout; relu' mask -> saved bits
//                  (+ mask x V product for d s_k / d v_k when the gradient is wanted)
//   lane = query : weighted combine, std, d sdf/dx, outputs
template <bool GRAD>
struct WarpStage {
    float sk[K][32];  // sum_j W2[j] relu(pre_j) of row (k, query)
    float4 gv[GRAD ? K : 1][32];
};

// shared-memory bytes of the decoder fragments a decode variant needs (vf / vv only for the mask products)
__host__ __device__ constexpr size_t dec_frag_bytes(bool grad) { return grad ? sizeof(DecFrags) : offsetof(DecFrags, vf); }

// min(bits(h), 1) as opaque PTX: nvcc otherwise expands the float-class reasoning of fmaxf + bit cast into ~10 instructions
__device__ __forceinline__ uint32_t umin1(float h) {
    uint32_t m;
    asm("min.u32 %0, %1, 1;" : "=r"(m) : "r"(__float_as_uint(h)));
    return m;
}

template <bool GRAD, bool MESH>
__global__ void __launch_bounds__(kBlock, GRAD ? 3 : 4) k_decode_mma(const DecodeParams p) {
    if (p.skip && *p.skip) return;
    extern __shared__ __align__(16) unsigned char s_raw[];
    DecFrags* fr = reinterpret_cast<DecFrags*>(s_raw);
    WarpStage<GRAD>* stages = reinterpret_cast<WarpStage<GRAD>*>(s_raw + dec_frag_bytes(GRAD));
    build_dec_frags(p.dec, fr, true, false, GRAD);
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    WarpStage<GRAD>& st = stages[warp];
    const float b2 = __ldg(p.dec.b2), scale = p.dec.sdf_scale;
    const int64_t nblocks = (p.n + kBlock - 1) / kBlock;
    // layer-1 B fragments live in registers for the CTA's lifetime: re-reading them from shared memory for every
    // m-tile (16 x LDS.128) made the shared-memory pipe the busiest unit of the kernel
    uint32_t breg[8][6];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
        const float4 wa = fr->w1a[nt][lane];
        const float4 wb = fr->w1b[nt][lane];
        // pairs as the k-steps consume them: (hi B[t], hi B[t+4]) (lo B[t], lo B[t+4]) (hi B[8+t], lo B[8+t])
        breg[nt][0] = __float_as_uint(wa.x); breg[nt][1] = __float_as_uint(wa.y); breg[nt][2] = __float_as_uint(wa.w);
        breg[nt][3] = __float_as_uint(wb.x); breg[nt][4] = __float_as_uint(wa.z); breg[nt][5] = __float_as_uint(wb.y);
    }

    for (int64_t blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
        const int64_t q0 = blk * kBlock + warp * 32;  // first processing position of this warp
        if (q0 >= p.n) continue;                      // warp-uniform
        const StageBlock& sb = p.stage[q0 >> 5];
        QueryLane L;  // gradient flavour only: lane = query state for the IDW gradient
        int64_t qg = 0;
        if (GRAD) {
            const int64_t i = q0 + lane;
            const bool live = i < p.n;
            qg = (live && p.perm) ? (int64_t)p.perm[i] : i;
            float x = 0.f, y = 0.f, z = 0.f;
            if (live) load_point(p.src, qg, x, y, z);
            resolve_query<true, false, false>(p, p.ws + i * kWs, i, qg, live, x, y, z, L);
        }
        __syncwarp();  // the previous iteration's epilogue reads of st are done

        // ---- lane = (g,t): software-pipelined over the 12 m-tiles ----
        // two-stage pipeline: the neighbour indices are loaded two m-tiles ahead, the rows they address one m-tile
        // ahead, so that neither load waits on the other inside the loop (ncu r1o: long-scoreboard was the top stall)
        float nraw[6];
        bool nany;
        int nla, nlb;
        auto fetch_idx = [&](int mt) {
            const int qa = 16 * (mt & 1) + g;
            nla = __ldg(&sb.idx[mt >> 1][qa]);
            nlb = __ldg(&sb.idx[mt >> 1][qa + 8]);
        };
        auto fetch = [&](int mt) {
            const int k = mt >> 1, half = mt & 1;
            const int qa = 16 * half + g, qb = qa + 8;
            const int la = nla, lb = nlb;
#pragma unroll
            for (int i = 0; i < 6; ++i) nraw[i] = 0.f;
            nany = __any_sync(0xffffffffu, (la >= 0) | (lb >= 0));
            if (la >= 0) {
                const float* f = p.features + (int64_t)la * F;
                nraw[0] = __ldg(f + t); nraw[2] = __ldg(f + t + 4);
            }
            if (lb >= 0) {
                const float* f = p.features + (int64_t)lb * F;
                nraw[1] = __ldg(f + t); nraw[3] = __ldg(f + t + 4);
            }
            // inputs 8..11: x - p_k and the constant 1 of the bias column
            nraw[4] = t < 3 ? __ldg(&sb.v[k][t < 3 ? t : 0][qa]) : 1.f;
            nraw[5] = t < 3 ? __ldg(&sb.v[k][t < 3 ? t : 0][qb]) : 1.f;
        };
        fetch_idx(0);
        fetch(0);
        fetch_idx(1);
#pragma unroll 1
        for (int mt = 0; mt < kMTiles; ++mt) {
            const int k = mt >> 1, half = mt & 1;
            const int qa = 16 * half + g, qb = qa + 8;
            float raw[6];
#pragma unroll
            for (int i = 0; i < 6; ++i) raw[i] = nraw[i];
            const bool any = nany;
            if (mt + 1 < kMTiles) fetch(mt + 1);
            if (mt + 2 < kMTiles) fetch_idx(mt + 2);
            uint32_t mA0 = 0u, mA1 = 0u, mB0 = 0u, mB1 = 0u;  // relu' words of rows qa and qb (this lane's share)
            float sA = 0.f, sB = 0.f, c2[4] = {0.f, 0.f, 0.f, 0.f};
            if (any) {  // warp-uniform: tiles without a single valid neighbour carry weight 0 everywhere
                uint32_t ah[6], al[6];
#pragma unroll
                for (int i = 0; i < 6; ++i) tf32_split_trunc(raw[i], ah[i], al[i]);
                // The three products of 3xTF32 in 5 k-steps with four DISJOINT A quads (a value shared by two quads
                // makes the register allocator rebuild the aligned tuples per n-tile: 60 MOVs per m-tile in SASS):
                //   q0 = lo(cols t, t+4)   x hi B          q1 = hi(cols t, t+4) x lo B, then x hi B
                //   q2 = lo | hi (col 8+t) x (hi B | lo B)  q3 = hi(col 8+t), 0 x (hi B | -)
                const uint32_t q2a = al[4], q2b = al[5], q2c = ah[4], q2d = ah[5];
                uint32_t q3a = ah[4], q3b = ah[5];
                asm volatile("" : "+r"(q3a), "+r"(q3b));  // distinct registers for the second use of hi(col 8+t)
                float acc[8][4];
#pragma unroll
                for (int nt = 0; nt < 8; ++nt) {
                    acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
                    const uint32_t h0 = breg[nt][0], h1 = breg[nt][1], l0 = breg[nt][2], l1 = breg[nt][3];
                    const uint32_t h2 = breg[nt][4], l2 = breg[nt][5];
                    mma_tf32(acc[nt], al[0], al[1], al[2], al[3], h0, h1);  // small terms first
                    mma_tf32(acc[nt], ah[0], ah[1], ah[2], ah[3], l0, l1);
                    mma_tf32(acc[nt], q2a, q2b, q2c, q2d, h2, l2);
                    mma_tf32(acc[nt], ah[0], ah[1], ah[2], ah[3], h0, h1);
                    mma_tf32(acc[nt], q3a, q3b, 0u, 0u, h2, l2);
                }
#pragma unroll
                for (int nt = 0; nt < 8; ++nt) {
                    const float2 w2 = fr->w2[4 * nt + t];
                    const float h0 = fmaxf(acc[nt][0], 0.f), h1 = fmaxf(acc[nt][1], 0.f);
                    const float h2 = fmaxf(acc[nt][2], 0.f), h3 = fmaxf(acc[nt][3], 0.f);
                    sA = fmaf(h0, w2.x, sA);
                    sA = fmaf(h1, w2.y, sA);
                    sB = fmaf(h2, w2.x, sB);
                    sB = fmaf(h3, w2.y, sB);
                    // relu'(pre) = [h > 0] = min(bits(h), 1) for h >= 0: one IMNMX + one IMAD per element, at the
                    // t = 0 positions of mma.cuh's row layout (shifted by t below)
                    uint32_t& wa = nt < 4 ? mA0 : mA1;
                    uint32_t& wb = nt < 4 ? mB0 : mB1;
                    wa = umin1(h0) * (1u << mask_row_bit0(nt, 0)) + wa;
                    wa = umin1(h1) * (1u << mask_row_bit0(nt, 1)) + wa;
                    wb = umin1(h2) * (1u << mask_row_bit0(nt, 2)) + wb;
                    wb = umin1(h3) * (1u << mask_row_bit0(nt, 3)) + wb;
                }
                sA += __shfl_xor_sync(0xffffffffu, sA, 1);
                sB += __shfl_xor_sync(0xffffffffu, sB, 1);
                sA += __shfl_xor_sync(0xffffffffu, sA, 2);
                sB += __shfl_xor_sync(0xffffffffu, sB, 2);
                if (GRAD || p.relu_mask) {  // the four lanes of a quad hold disjoint bits of the same two rows
                    mA0 <<= t; mA1 <<= t; mB0 <<= t; mB1 <<= t;
#pragma unroll
                    for (int o = 1; o <= 2; o <<= 1) {
                        mA0 |= __shfl_xor_sync(0xffffffffu, mA0, o);
                        mA1 |= __shfl_xor_sync(0xffffffffu, mA1, o);
                        mB0 |= __shfl_xor_sync(0xffffffffu, mB0, o);
                        mB1 |= __shfl_xor_sync(0xffffffffu, mB1, o);
                    }
                }
                if (GRAD) {  // d s_k / d v_k = sum_j relu'(pre_j) W2[j] W1[j][8..10]
                    uint32_t am[4][4];
                    mask_frags_rows(make_uint2(mA0, mA1), make_uint2(mB0, mB1), t, am);
#pragma unroll
                    for (int s = 2; s >= 0; --s)
#pragma unroll
                        for (int pp = 0; pp < 4; ++pp) {
                            const uint2 bv = fr->vv[pp][s][lane];
                            mma_bf16(c2, am[pp][0], am[pp][1], am[pp][2], am[pp][3], bv.x, bv.y);
                        }
                }
            }
            if (t == 0) {
                st.sk[k][qa] = sA;
                st.sk[k][qb] = sB;
            }
            if (GRAD) {
                float* ga = reinterpret_cast<float*>(&st.gv[k][qa]);
                float* gb = reinterpret_cast<float*>(&st.gv[k][qb]);
                if (t == 0) { ga[0] = c2[0]; ga[1] = c2[1]; gb[0] = c2[2]; gb[1] = c2[3]; }
                if (t == 1) { ga[2] = c2[0]; gb[2] = c2[2]; }
            }
            if (p.relu_mask && t < 2) {
                uint2* m2 = reinterpret_cast<uint2*>(p.relu_mask) + ((q0 >> 5) * K + k) * 32;
                if (t == 0) m2[qa] = make_uint2(mA0, mA1);
                else m2[qb] = make_uint2(mB0, mB1);
            }
        }
        __syncwarp();

        // ---- lane = query: weighted combine and outputs ----
        const int64_t i = q0 + lane;
        if (i >= p.n) continue;
        const int64_t q = GRAD ? qg : (p.perm ? (int64_t)p.perm[i] : i);
        if (MESH) {
            const int nnq = __ldg(&sb.nn[lane]);
            if (p.mask_out) p.mask_out[q] = (uint8_t)(nnq >= p.min_nn);
            if (nnq == 0) {  // MLP result unused where nothing is near (mesher.py:96-103)
                p.sdf[q] = 0.f;
                continue;
            }
        }
        float s[K], wq[K], sdf = 0.f;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            wq[k] = __ldg(&sb.w[k][lane]);
            s[k] = (st.sk[k][lane] + b2) * scale;  // decoder.py:80-81
            sdf = fmaf(s[k], wq[k], sdf);
        }
        if (p.sdf) p.sdf[q] = sdf;
        if (p.sdf_std) {  // tracker.py:246-247
            float var = 0.f;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                const float d = s[k] - sdf;
                var = fmaf(wq[k], d * d, var);
            }
            p.sdf_std[q] = sqrtf(var);
        }
        if (GRAD && p.grad_x) {
            // d sdf/dx = sum_k w_k R_k^T (ds_k/dv_k) + (1/S) sum_k (s_k - sdf) d wt_k/dx,  d wt_k/dx = -2 wt_k^2 (x - p_k)
            float gx = 0.f, gy = 0.f, gz = 0.f;
            const float inv_s = 1.0f / L.wsum;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                if (L.kidx[k] >= 0) {
                    const float4 gvk = st.gv[k][lane];
                    float ax = gvk.x * scale, ay = gvk.y * scale, az = gvk.z * scale;
                    if (p.rotate) quat_rotate(__ldg(&p.quats[L.kidx[k]]), true, ax, ay, az);
                    const float c = (s[k] - sdf) * (-2.f * L.wt[k] * L.wt[k]) * inv_s;
                    gx += wq[k] * ax + c * L.ux[k];
                    gy += wq[k] * ay + c * L.uy[k];
                    gz += wq[k] * az + c * L.uz[k];
                }
            }
            p.grad_x[3 * q] = gx;
            p.grad_x[3 * q + 1] = gy;
            p.grad_x[3 * q + 2] = gz;
        }
    }
}

#include "decode_tc.cuh"


// second pass of the exact training path: scatter after every query has read its certainties
__global__ void k_certainty_scatter(const StageBlock* __restrict__ stage, const int64_t* __restrict__ query_ts,
                                    const int32_t* __restrict__ perm, int64_t n, float* certainties,
                                    long long* ts_update) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // (k, position) pairs, position fastest
    const int64_t i = t % (((n + 31) / 32) * 32);
    const int k = (int)(t / (((n + 31) / 32) * 32));
    if (k >= K || i >= n) return;
    const StageBlock& sb = stage[i >> 5];
    const int32_t idx = sb.idx[k][i & 31];
    if (idx < 0) return;
    atomicAdd(&certainties[idx], sb.w[k][i & 31]);
    if (query_ts && ts_update) atomicMax(&ts_update[idx], (long long)query_ts[perm ? perm[i] : i]);
}

// the same second pass for the un-fused form, from its [n,K] outputs
__global__ void k_certainty_scatter_knn(const int32_t* __restrict__ knn_idx, const float* __restrict__ knn_w,
                                        const int64_t* __restrict__ query_ts, int64_t n, float* certainties,
                                        long long* ts_update) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * K) return;
    const int32_t i = knn_idx[t];
    if (i < 0) return;
    atomicAdd(&certainties[i], knn_w[t]);
    if (query_ts && ts_update) atomicMax(&ts_update[i], (long long)query_ts[t / K]);
}

static inline int64_t ws_bytes(int64_t n) { return (((n > 0 ? n : 1) * (int64_t)kWs * 4 + 255) / 256) * 256; }
static inline int64_t stage_bytes(int64_t n) { return (((n > 0 ? n : 1) + 31) / 32) * (int64_t)sizeof(StageBlock); }
// workspace layout: [search ranks | binning scratch | stencil masks | stage (when the caller keeps none)]
static inline int64_t ws_off_masks(int64_t n) { return ws_bytes(n) + bin_scratch_bytes(n); }
static inline int64_t ws_off_stage(int64_t n) { return ((ws_off_masks(n) + kStencilMasks * 8 + 255) / 256) * 256; }

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
    if (n == 0) {
        if (mode == 0 && a->zero_fill && a->zero_fill_bytes > 0) cudaMemsetAsync(a->zero_fill, 0, (size_t)a->zero_fill_bytes, st);
        return PINQ_OK;
    }
    const MapDev map = to_dev(a->map);
    PointSrc src;
    if (grid_src) src = *grid_src;
    else {
        src.xyz = a->xyz; src.grid = false;
        src.ox = src.oy = src.oz = src.voxel = 0.f; src.ny = src.nz = 1; src.first = 0;
    }
    const bool recip = (a->flags & PINQ_F_RECIP_DIV) != 0;
    int32_t* ws = reinterpret_cast<int32_t*>(a->workspace);
    const int32_t* perm = nullptr;
    if (a->perm && !grid_src && mode != 1) {  // spatial binning: processing order only, results unchanged
        ktime_mark(PINQ_KT_BIN, st);
        rc = launch_spatial_bin(a->xyz, n, map.res, recip, reinterpret_cast<unsigned char*>(a->workspace) + ws_bytes(n),
                                a->perm, st);
        if (rc) return rc;
        perm = a->perm;
    }
    const bool brick = a->map.brick_tab != nullptr;
    BrickDev bd;
    if (brick) {
        if (!a->map.brick_rec4 || a->map.brick_cap < 1024 || (a->map.brick_cap & (a->map.brick_cap - 1)) ||
            a->map.stencil_radius < 0 || a->map.stencil_radius > 2 || map.kn > 125 || a->map.m >= (1ll << kRankBits)) {
            set_error("bad brick view (needs records, power-of-two capacity, stencil radius <= 2, m < 2^24)");
            return PINQ_ERR_BAD_ARG;
        }
        bd.tab = reinterpret_cast<const BrickEntry*>(a->map.brick_tab);
        bd.cap_mask = a->map.brick_cap - 1;
        bd.rec = reinterpret_cast<const float4*>(a->map.brick_rec4);
        bd.radius = a->map.stencil_radius;
        bd.exact_keys = (a->flags & PINQ_F_EXACT_KEYS) != 0;
    }

    DecodeParams p;
    p.rec = brick ? bd.rec : map.rec; p.max_d2 = map.max_d2; p.dec = a->dec; p.src = src; p.n = n;
    p.rotate = (a->flags & PINQ_F_ROTATE) != 0;
    // Exactness of the queried certainty in training mode needs all reads before any scatter: if the
    // caller wants it (and gave knn buffers), scatter in a second kernel; otherwise fuse the atomics.
    const bool two_pass = training && a->certainty && (mode != 1 || (a->knn_idx && a->knn_w));
    p.scatter_in_kernel = training && !two_pass;
    p.ws = ws; p.perm = perm; p.features = a->features; p.point_xyz = a->point_xyz; p.certainties = a->certainties;
    p.quats = reinterpret_cast<const float4*>(a->quats);
    p.ts_update = reinterpret_cast<long long*>(a->ts_update); p.query_ts = a->query_ts;
    p.sdf = a->sdf; p.grad_x = a->grad_x; p.sdf_std = a->sdf_std; p.nn_count = a->nn_count;
    p.certainty = a->certainty; p.knn_idx = a->knn_idx; p.knn_w = a->knn_w; p.knn_d2 = a->knn_d2;
    p.feat_out = feat_out; p.min_nn = min_nn; p.mask_out = mask_out; p.skip = a->skip;
    if (a->zero_fill && a->zero_fill_bytes > 0 && mode == 0) {
        if ((a->zero_fill_bytes & 15) || (reinterpret_cast<uintptr_t>(a->zero_fill) & 15)) {
            set_error("zero_fill: 16-byte aligned buffer, size a multiple of 16");
            return PINQ_ERR_BAD_ARG;
        }
        if (brick && a->zero_fill_bytes / 16 <= (int64_t)grid_for(n, kBlock) * kBlock && !a->skip) {
            p.zero_fill = reinterpret_cast<float4*>(a->zero_fill);
            p.zero_fill4 = a->zero_fill_bytes / 16;
        } else if (cudaMemsetAsync(a->zero_fill, 0, (size_t)a->zero_fill_bytes, st) != cudaSuccess) {
            return check_launch("zero_fill memset");
        }
    }
    p.relu_mask = reinterpret_cast<uint32_t*>(a->relu_mask);
    p.stage = reinterpret_cast<StageBlock*>(a->stage ? a->stage
                                                     : reinterpret_cast<unsigned char*>(a->workspace) + ws_off_stage(n));

    ktime_mark(PINQ_KT_SEARCH, st);
    const bool resolve_in_search = brick && mode != 1;
    if (brick) {
        const unsigned long long* smasks = reinterpret_cast<const unsigned long long*>(a->map.stencil_masks);
        if (!smasks) {
            unsigned long long* built = reinterpret_cast<unsigned long long*>(
                reinterpret_cast<unsigned char*>(a->workspace) + ws_off_masks(n));
            k_stencil_masks<<<1, 256, 0, st>>>(map.stencil_dx, map.kn, bd.radius, built);
            smasks = built;
        }
        const size_t smem = kBrickSmem;
        static bool search_carve = false;
        if (!search_carve) {
            // shared memory for 8 CTAs and no more: what is left of the 256 KB stays L1 for the candidate records
            const char* ce = getenv("PINQ_SEARCH_CARVEOUT");
            const int carve = ce ? atoi(ce) : (int)((PINQ_SEARCH_MINB * (smem + 1024) * 100 + 232448 - 1) / 232448);
            cudaFuncSetAttribute(k_search_brick<true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            cudaFuncSetAttribute(k_search_brick<false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            cudaFuncSetAttribute(k_search_brick<true, false>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            cudaFuncSetAttribute(k_search_brick<false, false>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            search_carve = true;
        }
        const bool dtest = !(a->flags & PINQ_F_STENCIL_IN_RANGE);
        const unsigned gsb = grid_for(n, kBlock);
        if (resolve_in_search) {
            if (dtest) k_search_brick<true, true><<<gsb, kBlock, smem, st>>>(map, bd, src, n, recip, perm, smasks, ws, p);
            else k_search_brick<true, false><<<gsb, kBlock, smem, st>>>(map, bd, src, n, recip, perm, smasks, ws, p);
        } else {
            if (dtest) k_search_brick<false, true><<<gsb, kBlock, smem, st>>>(map, bd, src, n, recip, perm, smasks, ws, p);
            else k_search_brick<false, false><<<gsb, kBlock, smem, st>>>(map, bd, src, n, recip, perm, smasks, ws, p);
        }
    } else {
        k_search<<<grid_for(n, kBlock), kBlock, map.kn * sizeof(uint32_t), st>>>(map, src, n, recip, perm, ws, a->skip);
    }
    rc = check_launch("k_search");
    if (rc) return rc;
    ktime_mark(PINQ_KT_RESOLVE, st);
    if (mode == 1) {
        k_query_feature<<<grid_for(n, kBlock), kBlock, 0, st>>>(p);
    } else {
        if (!resolve_in_search) {
            k_resolve<<<grid_for(n, kBlock), kBlock, 0, st>>>(p);
            rc = check_launch("k_resolve");
            if (rc) return rc;
        }
        if (a->decode_wait_event) cudaStreamWaitEvent(st, reinterpret_cast<cudaEvent_t>(a->decode_wait_event), 0);
        ktime_mark(PINQ_KT_DECODE, st);
        // persistent-ish grid: the decoder fragments are built once per CTA and amortised over its blocks
        const unsigned nb = grid_for(n, kBlock);
        const size_t sm_plain = dec_frag_bytes(false) + (kBlock / 32) * sizeof(WarpStage<false>);
        const size_t sm_grad = dec_frag_bytes(true) + (kBlock / 32) * sizeof(WarpStage<true>);
        static bool carve = false;
        if (!carve) {
            cudaFuncSetAttribute(k_decode_mma<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_grad);
            carve = true;
        }
        // one wave of resident CTAs, each striding over the blocks (a partial second wave would idle SMs)
        static int occ_plain = 0, occ_mesh = 0, occ_grad = 0, sms = 0;
        if (!sms) {
            int dev = 0;
            cudaGetDevice(&dev);
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_plain, k_decode_mma<false, false>, kBlock, sm_plain);
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_mesh, k_decode_mma<false, true>, kBlock, sm_plain);
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_grad, k_decode_mma<true, false>, kBlock, sm_grad);
            // shared memory for the resident CTAs and no more: the rest of the 256 KB stays L1 for the gathers
            auto pct = [](int occ, size_t smem) { return (int)(((size_t)(occ > 0 ? occ : 1) * (smem + 1024) * 100 + 232448 - 1) / 232448); };
            cudaFuncSetAttribute(k_decode_mma<false, false>, cudaFuncAttributePreferredSharedMemoryCarveout, pct(occ_plain, sm_plain));
            cudaFuncSetAttribute(k_decode_mma<false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, pct(occ_mesh, sm_plain));
            cudaFuncSetAttribute(k_decode_mma<true, false>, cudaFuncAttributePreferredSharedMemoryCarveout, pct(occ_grad, sm_grad));
        }
        // all flavours: tcgen05 / TMEM kernel (decode_tc.cuh); PINQ_DECODE=mma keeps the mma.sync ones, PINQ_DECODE_GRAD=mma only
        // for the gradient flavour (A/B partner of k_decode_tc_grad)
        static int use_tc = -1, use_tc_grad = 1;
        static size_t dec_smem = tc::kSmemRequest;
        if (use_tc < 0) {
            const char* e = getenv("PINQ_DECODE");
            const char* eg = getenv("PINQ_DECODE_GRAD");
            use_tc = !(e && e[0] == 'm');
            use_tc_grad = !(eg && eg[0] == 'm');
            const char* se = getenv("PINQ_DEC_SMEM_KB");
            const char* ce = getenv("PINQ_DEC_CARVEOUT");
            dec_smem = se ? (size_t)atoi(se) * 1024 : tc::kSmemRequest;
            const int carve = ce ? atoi(ce) : tc::kCarveoutPct;
            cudaFuncSetAttribute(k_decode_tc<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dec_smem);
            cudaFuncSetAttribute(k_decode_tc<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dec_smem);
            cudaFuncSetAttribute(k_decode_tc_grad, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc::kSmemRequestGrad);
            cudaFuncSetAttribute(k_decode_tc<false>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            cudaFuncSetAttribute(k_decode_tc<true>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            cudaFuncSetAttribute(k_decode_tc_grad, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
        }
        if (use_tc && !(a->flags & PINQ_F_MMA_SYNC) && (!a->grad_x || (use_tc_grad && mode != 2))) {
            const unsigned nb128 = grid_for(n, tc::kRows);
            // TMEM columns per CTA bound the residency (96 -> 5 CTAs; gradient flavour 128 -> 4)
            const unsigned cap128 = (unsigned)sms * (a->grad_x ? tc::kCtasPerSmGrad : tc::kCtasPerSm);
            const unsigned g128 = nb128 < cap128 ? nb128 : cap128;
            if (a->grad_x) k_decode_tc_grad<<<g128, tc::kRows, tc::kSmemRequestGrad, st>>>(p);
            else if (mode == 2) k_decode_tc<true><<<g128, tc::kRows, dec_smem, st>>>(p);
            else k_decode_tc<false><<<g128, tc::kRows, dec_smem, st>>>(p);
        } else {
            const int occ = mode == 2 ? occ_mesh : (a->grad_x ? occ_grad : occ_plain);
            const unsigned cap = (unsigned)(sms * (occ > 0 ? occ : 1));
            const unsigned g = nb < cap ? nb : cap;
            if (mode == 2) k_decode_mma<false, true><<<g, kBlock, sm_plain, st>>>(p);
            else if (a->grad_x) k_decode_mma<true, false><<<g, kBlock, sm_grad, st>>>(p);
            else k_decode_mma<false, false><<<g, kBlock, sm_plain, st>>>(p);
        }
    }
    rc = check_launch("k_decode");
    if (rc) return rc;
    ktime_mark(two_pass ? PINQ_KT_SCATTER : PINQ_KT_FWD_END, st);
    if (two_pass && mode == 1) {
        k_certainty_scatter_knn<<<grid_for(n * K, 256), 256, 0, st>>>(a->knn_idx, a->knn_w, a->query_ts, n, a->certainties,
                                                                      reinterpret_cast<long long*>(a->ts_update));
        return check_launch("k_certainty_scatter");
    }
    if (two_pass) {
        const int64_t padded = ((n + 31) / 32) * 32;
        k_certainty_scatter<<<grid_for(padded * K, 256), 256, 0, st>>>(p.stage, a->query_ts, perm, n, a->certainties,
                                                                       reinterpret_cast<long long*>(a->ts_update));
        ktime_mark(PINQ_KT_FWD_END, st);
        return check_launch("k_certainty_scatter");
    }
    return PINQ_OK;
}

}  // namespace pinq

using namespace pinq;

extern "C" {

int64_t pinq_query_workspace_bytes(int64_t n) { return ws_off_stage(n > 0 ? n : 1) + stage_bytes(n); }
int64_t pinq_stage_bytes(int64_t n) { return stage_bytes(n); }

int64_t pinq_relu_mask_words(int64_t n) { return ((n > 0 ? n : 1) + 31) / 32 * (int64_t)kMTiles * 32; }

int pinq_stencil_masks_build(const int32_t* stencil_dx, int32_t kn, int32_t stencil_radius, uint64_t* out512, void* stream) {
    PINQ_REQUIRE(stencil_dx && out512, "null pointer");
    PINQ_REQUIRE(kn > 0 && kn <= PINQ_MAX_STENCIL && stencil_radius >= 0 && stencil_radius <= 2, "stencil radius 0..2 only");
    k_stencil_masks<<<1, 256, 0, as_stream(stream)>>>(stencil_dx, kn, stencil_radius,
                                                      reinterpret_cast<unsigned long long*>(out512));
    return check_launch("pinq_stencil_masks_build");
}

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
    PINQ_REQUIRE(((uintptr_t)a->features & 31) == 0, "features must be 32-byte aligned (rows are gathered with 256-bit loads)");
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

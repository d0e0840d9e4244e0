// query_bwd.cu -- hand-derived backward of the fused query w.r.t. neural-point features and decoder
// parameters.  Replaces torch.autograd through
// model/neural_points.py:562-632 + model/decoder.py:66-88 as triggered at utils/mapper.py:572.
//
//   a_k        = g * w_k * scale            (g = dL/dsdf; the IDW weights carry no parameter gradient)
//   m_kj       = relu'(pre_kj)              saved by the forward as one bit per (row, hidden unit)
//   dL/df_k    = a_k * (m_k . V)[0:8],  V[j][i] = W2[j] W1[j][i]      -> red.global.add.v2.f32 into grad_features[idx_k]
//   G[j][i]    = sum_rows m_kj a_k in_k[i]  (in_k = [f_k, v_k, 1])    -> reduction over rows on the tensor cores
//   dL/dW1[j][i] = W2[j] G[j][i]   dL/db1[j] = W2[j] G[j][11]
//   dL/dW2[j]  = sum_rows a_k relu(pre_kj) = sum_i W1b[j][i] G[j][i]  (relu(pre) = m * pre, pre = W1b . in)
//   dL/db2     = sum_rows a_k
// Nothing of the hidden layer is recomputed: the backward needs the mask (48 B/query), the K indices and
// weights, and -- only while the decoder is trainable -- the gathered inputs.  Both products have one EXACT
// operand (the 0/1 mask), so a 3-way bf16 split of the other one keeps fp32 accuracy (csrc/mma.cuh).
#include <stddef.h>
#include <stdlib.h>

#include "mma.cuh"

namespace pinq {

struct BwdParams {
    pinq_decoder_view dec;
    int64_t n;
    const float* features;
    const StageBlock* stage;
    const uint32_t* relu_mask;
    const int32_t* perm;
    const float* grad_sdf;
    float* grad_features;
    float *grad_W1, *grad_b1, *grad_W2, *grad_b2;
};

__device__ __forceinline__ void red_add_v2(float* addr, float a, float b) {
    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(addr), "f"(a), "f"(b) : "memory");
}

template <bool DEC>
struct BwdStage {  // per-warp copy of what phase B reads row-wise (shared-memory latency instead of L1/L2 latency)
    int32_t idx[K][32];
    float ab[K][32];  // a_k = g w_k scale of row (k, query); 0 for invalid rows
    float v[DEC ? K : 1][3][32];
};

constexpr int kGCols = IN + 1;  // 11 inputs + the bias column

// Same warp <-> row mapping as k_decode_mma: one warp = 32 queries = 12 m-tiles, m-tile mt = 2 k + half
// holds neighbour k of queries 16 half + {g, g + 8}.
template <bool DEC, bool FEAT>
__global__ void __launch_bounds__(kBlock, 4) k_bwd_mma(const BwdParams p) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    DecFrags* fr = reinterpret_cast<DecFrags*>(s_raw);
    BwdStage<DEC>* stages = reinterpret_cast<BwdStage<DEC>*>(s_raw + sizeof(DecFrags));
    __shared__ float s_G[DEC ? H * kGCols : 1];
    __shared__ float s_db2;
    build_dec_frags(p.dec, fr, false, FEAT, false);
    if (DEC) {
        for (int e = threadIdx.x; e < H * kGCols; e += blockDim.x) s_G[e] = 0.f;
        if (threadIdx.x == 0) s_db2 = 0.f;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    BwdStage<DEC>& st = stages[warp];
    const float scale = p.dec.sdf_scale;
    const int64_t nblocks = (p.n + kBlock - 1) / kBlock;

    float G[4][2][4];  // G[16 jt + g (+8)][8 nt + 2t (+1)], C-fragment layout
#pragma unroll
    for (int jt = 0; jt < 4; ++jt)
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) G[jt][nt][0] = G[jt][nt][1] = G[jt][nt][2] = G[jt][nt][3] = 0.f;
    float db2 = 0.f;

    for (int64_t blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
        const int64_t q0 = blk * kBlock + warp * 32;
        if (q0 >= p.n) continue;  // warp-uniform
        const int64_t i = q0 + lane;  // processing position; q = the query it holds
        const bool live = i < p.n;
        const int64_t q = (live && p.perm) ? (int64_t)p.perm[i] : i;
        const StageBlock& sb = p.stage[q0 >> 5];
        // ---- phase A: lane = query (coalesced, independent loads) ----
        __syncwarp();
        {
            const float gq = live ? p.grad_sdf[q] : 0.f;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                const int32_t li = __ldg(&sb.idx[k][lane]);
                const float a = (live && li >= 0) ? gq * __ldg(&sb.w[k][lane]) * scale : 0.f;
                st.idx[k][lane] = li;
                st.ab[k][lane] = a;
                if (DEC) {
                    db2 += a;
#pragma unroll
                    for (int c = 0; c < 3; ++c) st.v[k][c][lane] = __ldg(&sb.v[k][c][lane]);
                }
            }
        }
        __syncwarp();

        // ---- phase B: lane = (g, t); the mask bits and the gathered feature column of the next m-tile are in
        // flight while this one is multiplied ----
        uint2 nra, nrb;  // relu' words of rows qa, qb of the next m-tile
        float nf[DEC ? 4 : 1];
        auto fetch = [&](int mt) {
            const uint2* m2 = reinterpret_cast<const uint2*>(p.relu_mask) + ((q0 >> 5) * K + (mt >> 1)) * 32 + 16 * (mt & 1) + g;
            nra = __ldg(m2);
            nrb = __ldg(m2 + 8);
            if (DEC) {
#pragma unroll
                for (int r4 = 0; r4 < 4; ++r4) {
                    const int lr = st.idx[mt >> 1][16 * (mt & 1) + 2 * t + (r4 & 1) + 8 * (r4 >> 1)];
                    nf[r4] = lr >= 0 ? __ldg(p.features + (int64_t)lr * F + g) : 0.f;
                }
            }
        };
        fetch(0);
#pragma unroll 1
        for (int mt = 0; mt < kMTiles; ++mt) {
            const int k = mt >> 1, half = mt & 1;
            const int qa = 16 * half + g, qb = qa + 8;
            const int la = st.idx[k][qa], lb = st.idx[k][qb];
            const uint2 ra = nra, rb = nrb;
            float fcol[DEC ? 4 : 1];
            if (DEC) {
#pragma unroll
                for (int r4 = 0; r4 < 4; ++r4) fcol[r4] = nf[r4];
            }
            if (mt + 1 < kMTiles) fetch(mt + 1);
            if (!__any_sync(0xffffffffu, (la >= 0) | (lb >= 0))) continue;  // warp-uniform
            uint32_t am[4][4];
            mask_frags_rows(ra, rb, t, am);
            if (FEAT) {
                float c2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int s = 2; s >= 0; --s)
#pragma unroll
                    for (int pp = 0; pp < 4; ++pp) {
                        const uint2 bv = fr->vf[pp][s][lane];
                        mma_bf16(c2, am[pp][0], am[pp][1], am[pp][2], am[pp][3], bv.x, bv.y);
                    }
                // c2 = (m . V)[row g | g+8][2t, 2t+1]: the quad covers one 32-byte gradient row
                if (la >= 0) {
                    const float aa = st.ab[k][qa];
                    red_add_v2(p.grad_features + (int64_t)la * F + 2 * t, c2[0] * aa, c2[1] * aa);
                }
                if (lb >= 0) {
                    const float aa = st.ab[k][qb];
                    red_add_v2(p.grad_features + (int64_t)lb * F + 2 * t, c2[2] * aa, c2[3] * aa);
                }
            }
            if (DEC) {
                // A = mask^T (hidden unit x row): transpose the 8x8 blocks of the mask fragments
                uint32_t at[4][4];
#pragma unroll
                for (int pp = 0; pp < 4; ++pp) {
                    at[pp][0] = movmatrix_trans(am[pp][0]);  // (j = 16p + g     ; rows 2t, 2t+1)
                    at[pp][2] = movmatrix_trans(am[pp][1]);  // (j = 16p + g     ; rows 2t+8, 2t+9)
                    at[pp][1] = movmatrix_trans(am[pp][2]);  // (j = 16p + 8 + g ; rows 2t, 2t+1)
                    at[pp][3] = movmatrix_trans(am[pp][3]);  // (j = 16p + 8 + g ; rows 2t+8, 2t+9)
                }
                // B = X (row x input column): rows 2t, 2t+1, 2t+8, 2t+9 of this m-tile, column g (and 8 + g)
                float xf[4], xv[4];
#pragma unroll
                for (int r4 = 0; r4 < 4; ++r4) {
                    const int qr = 16 * half + 2 * t + (r4 & 1) + 8 * (r4 >> 1);
                    const float ar = st.ab[k][qr];
                    xf[r4] = ar * fcol[r4];
                    // columns 8..10 = x - p_k from the stage, column 11 = the bias input 1
                    xv[r4] = g < 3 ? ar * st.v[k][g < 3 ? g : 0][qr] : (g == 3 ? ar : 0.f);
                }
                uint32_t pf[2][3], pv[2][3];  // B fragments {rows 2t, 2t+1} {rows 2t+8, 2t+9} of the three bf16 pieces
                bf16_split3_pair(xf[0], xf[1], pf[0]);
                bf16_split3_pair(xf[2], xf[3], pf[1]);
                bf16_split3_pair(xv[0], xv[1], pv[0]);
                bf16_split3_pair(xv[2], xv[3], pv[1]);
#pragma unroll
                for (int jt = 0; jt < 4; ++jt) {
                    float tf[4] = {0.f, 0.f, 0.f, 0.f}, tv[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for (int s = 2; s >= 0; --s) {
                        mma_bf16(tf, at[jt][0], at[jt][1], at[jt][2], at[jt][3], pf[0][s], pf[1][s]);
                        mma_bf16(tv, at[jt][0], at[jt][1], at[jt][2], at[jt][3], pv[0][s], pv[1][s]);
                    }
                    // accumulate outside the tensor core (round-to-nearest fp32 adds over the long row reduction)
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        G[jt][0][c] += tf[c];
                        G[jt][1][c] += tv[c];
                    }
                }
            }
        }
    }

    if (DEC) {
#pragma unroll
        for (int jt = 0; jt < 4; ++jt)
#pragma unroll
            for (int nt = 0; nt < 2; ++nt)
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int j = 16 * jt + g + 8 * (c >> 1), i = 8 * nt + 2 * t + (c & 1);
                    if (i < kGCols) atomicAdd(&s_G[j * kGCols + i], G[jt][nt][c]);
                }
        db2 = warp_sum(db2);
        if (lane == 0) atomicAdd(&s_db2, db2);
        __syncthreads();
        for (int e = threadIdx.x; e < H * kGCols; e += blockDim.x) {
            const int j = e / kGCols, i = e - j * kGCols;
            const float v = __ldg(&p.dec.W2[j]) * s_G[e];
            if (i < IN) atomicAdd(&p.grad_W1[j * IN + i], v);
            else atomicAdd(&p.grad_b1[j], v);
        }
        for (int j = threadIdx.x; j < H; j += blockDim.x) {
            float acc = __ldg(&p.dec.b1[j]) * s_G[j * kGCols + IN];
#pragma unroll
            for (int i = 0; i < IN; ++i) acc = fmaf(__ldg(&p.dec.W1[j * IN + i]), s_G[j * kGCols + i], acc);
            atomicAdd(&p.grad_W2[j], acc);
        }
        if (threadIdx.x == 0) atomicAdd(p.grad_b2, s_db2);
    }
}

#include "bwd_tc.cuh"

}  // namespace pinq

using namespace pinq;

extern "C" int pinq_query_fused_bwd(const pinq_bwd_args* a, void* stream) {
    PINQ_REQUIRE(a != nullptr, "null args");
    int rc = check_decoder(a->dec, a->k, a->feature_dim);
    if (rc) return rc;
    PINQ_REQUIRE(a->features && a->stage && a->grad_sdf && a->relu_mask,
                 "null input (the backward needs the stage and the relu mask saved by pinq_query_fused_fwd)");
    PINQ_REQUIRE(((uintptr_t)a->features & 31) == 0 && ((uintptr_t)a->grad_features & 15) == 0,
                 "features must be 32-byte aligned (256-bit row loads), grad_features 16-byte aligned (vector reductions)");
    const bool dec = a->grad_W1 != nullptr;
    PINQ_REQUIRE(!dec || (a->grad_b1 && a->grad_W2 && a->grad_b2), "decoder gradients: all four or none");
    const bool feat = a->grad_features != nullptr;
    if (a->n == 0 || (!dec && !feat)) return PINQ_OK;
    BwdParams p;
    p.dec = a->dec; p.n = a->n;
    p.features = a->features; p.stage = reinterpret_cast<const StageBlock*>(a->stage);
    p.relu_mask = a->relu_mask; p.perm = a->perm;
    p.grad_sdf = a->grad_sdf; p.grad_features = a->grad_features;
    p.grad_W1 = a->grad_W1; p.grad_b1 = a->grad_b1; p.grad_W2 = a->grad_W2; p.grad_b2 = a->grad_b2;
    cudaStream_t st = as_stream(stream);
    const unsigned nb = grid_for(a->n, kBlock);
    const size_t sm_dec = sizeof(DecFrags) + (kBlock / 32) * sizeof(BwdStage<true>);
    const size_t sm_feat = sizeof(DecFrags) + (kBlock / 32) * sizeof(BwdStage<false>);
    const size_t sm = dec ? sm_dec : sm_feat;
    // one wave of resident CTAs striding over the blocks (every CTA ends in 833 global atomics when dec)
    static int occ_a = 0, occ_b = 0, occ_c = 0, sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_a, k_bwd_mma<true, true>, kBlock, sm_dec);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_b, k_bwd_mma<true, false>, kBlock, sm_dec);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_c, k_bwd_mma<false, true>, kBlock, sm_feat);
    }
    // tcgen05 / TMEM kernel (bwd_tc.cuh); PINQ_BWD=mma keeps the mma.sync one
    static int use_tc = -1;
    if (use_tc < 0) {
        const char* e = getenv("PINQ_BWD");
        use_tc = !(e && e[0] == 'm');
        cudaFuncSetAttribute(k_bwd_tc<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tcb::kSmemRequest);
        cudaFuncSetAttribute(k_bwd_tc<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tcb::kSmemRequest);
        cudaFuncSetAttribute(k_bwd_tc<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tcb::kSmemRequest);
        const char* ce = getenv("PINQ_BWD_CARVEOUT");
        // 4 CTAs x (31 + 1 KB) of shared memory; the rest of the 256 KB stays L1 for the stage blocks, relu' masks and feature
        // rows the tiles are built from (100 % carve-out: 0.594 ms, 66 %: 0.475 ms per 4 Mi queries)
        const int carve = ce ? atoi(ce) : (int)((tcb::kCtasPerSm * (tcb::kSmemRequest + 1024) * 100 + 232448 - 1) / 232448);
        cudaFuncSetAttribute(k_bwd_tc<true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
        cudaFuncSetAttribute(k_bwd_tc<true, false>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
        cudaFuncSetAttribute(k_bwd_tc<false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
    }
    ktime_mark(PINQ_KT_BWD, st);
    if (use_tc && !(a->flags & PINQ_F_MMA_SYNC)) {
        const unsigned nb128 = grid_for(a->n, tcb::kRows);
        const unsigned cap128 = (unsigned)sms * tcb::kCtasPerSm;  // resident CTAs: bounded by TMEM columns and the tile buffers
        const unsigned g128 = nb128 < cap128 ? nb128 : cap128;
        if (dec && feat) k_bwd_tc<true, true><<<g128, tcb::kRows, tcb::kSmemRequest, st>>>(p);
        else if (dec) k_bwd_tc<true, false><<<g128, tcb::kRows, tcb::kSmemRequest, st>>>(p);
        else k_bwd_tc<false, true><<<g128, tcb::kRows, tcb::kSmemRequest, st>>>(p);
    } else {
        const int occ = dec ? (feat ? occ_a : occ_b) : occ_c;
        const unsigned cap = (unsigned)(sms * (occ > 0 ? occ : 1));
        const unsigned g = nb < cap ? nb : cap;
        if (dec && feat) k_bwd_mma<true, true><<<g, kBlock, sm, st>>>(p);
        else if (dec) k_bwd_mma<true, false><<<g, kBlock, sm, st>>>(p);
        else k_bwd_mma<false, true><<<g, kBlock, sm, st>>>(p);
    }
    ktime_mark(PINQ_KT_BWD_END, st);
    return check_launch("pinq_query_fused_bwd");
}

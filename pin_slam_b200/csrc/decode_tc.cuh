// decode_tc.cuh -- fused decode on the 5th-generation tensor cores: tcgen05.mma with the accumulator in TMEM.
// (included by query.cu after DecodeParams / StageBlock; sm_100a only)
//
// model/decoder.py:66-88 for the K neighbours of a block of 128 queries = K products [128 x 12] x [12 x 64]:
// tile k holds neighbour k of the 128 queries, so THREAD = QUERY = accumulator row (TMEM lane) in every tile and the
// whole epilogue (relu, . W2, relu' bits, IDW combine, std) is thread-local -- no fragment shuffles, no MMA issue slots
// in the CUDA-core instruction stream (ncu on the mma.sync formulation: tensor pipe and issue both ~50 % busy, stalled on
// each other's latencies at 16 warps/SM).
//
//   stage   thread = row: 32-byte feature row + (x - p_k) from the stage block -> 3xTF32 pieces -> ONE tcgen05.st of the
//           32 A columns of the thread's TMEM lane (the A operand never touches shared memory: ncu on the shared-memory
//           version showed L1TEX 80 % busy with the tile writes, the tensor core's operand reads and the gathers)
//   mma     one thread: 5 x tcgen05.mma.kind::tf32 M=128 N=64 K=8, A from TMEM, B from shared memory (small terms first)
//              [lo(f0..7)] x [hi W]   [hi(f0..7)] x [lo W]   [lo(v,0) | hi(v,1)] x [hi W ; lo W]
//              [hi(f0..7)] x [hi W]   [hi(v,1) | 0] x [hi W ; -]          then tcgen05.commit -> mbarrier
//           `hi` is the raw fp32 word (the tensor core ignores the low 13 mantissa bits), lo = x - trunc(x) is exact
//   epilog  tcgen05.ld 32x32b: 64 pre-activations x_j of the thread's row.  relu(x) W2 = |x| W2/2 + x W2/2: the first half is
//           one FFMA per element (|x| as operand modifier, W2/2 broadcast from shared memory), the second is linear in the
//           decoder inputs and added once per tile (in . u, u = sum_j W2_j/2 [W1 | b1]_j, 12 FFMAs at staging time); relu'
//           bits collected from the SIGN of the pre-activation with one funnel shift per element -- 2 instructions per
//           element, one on the FMA pipe and one on the (half-rate) ALU pipe
//   grad    (k_decode_tc_grad, the tracker's flavour) the relu' bits, expanded to bytes in 16 more TMEM columns, are the A
//           operand of a second product (tcgen05.mma.kind::i8) against W2 W1[:, 8:11] in 32-bit fixed point: d s_k / d v_k
// 96 TMEM columns per CTA (64 accumulator + 32 A) -> 5 CTAs per SM; the other CTAs' work hides the MMA latency
// (single-buffered tiles measured as fast as double-buffered ones at 4 CTAs).  Gradient flavour: 128 columns, 3 CTAs.
//
// Deviation (documented in DESIGN.md): relu'(+0.0) is taken as 1 (torch: 0); reachable only when a hidden
// pre-activation is exactly +0.0.
#pragma once
#include "tc.cuh"

// (inside namespace pinq)
namespace tc {

constexpr int kRows = 128;                          // UMMA M = queries per CTA block = threads per CTA
constexpr int kChunksB = 6;                         // [h0 h1 l0 l1 h2 l2] per hidden unit, 16 bytes (4 tf32) each
constexpr uint32_t kChunkBBytes = H * 16;           // 1024: K-direction (leading) byte offset of B
constexpr uint32_t kTmemColsD = H;                  // accumulator: 64 fp32 columns
constexpr uint32_t kTmemColsA = 32;                 // A operand: 4 k-steps of 8 tf32 columns
constexpr unsigned kCtasPerSm = 5;                  // 5 x 96 TMEM columns <= 512
constexpr int kGradN = 16;                          // GRAD: mask x (W2 W1[:, 8:11]) in 4 digits -> 12 columns, padded to 16
constexpr uint32_t kTmemColsG = 32;                 // GRAD: 16 columns of mask bytes (A operand) + 16 of s32 results
#ifndef PINQ_DEC_GRAD_CTAS
#define PINQ_DEC_GRAD_CTAS 3
#endif
constexpr unsigned kCtasPerSmGrad = PINQ_DEC_GRAD_CTAS;  // 147 registers: 3 CTAs without spills beat 4 with (1.66 vs 1.75 ms per 4 Mi queries)
constexpr size_t kSmemRequestGrad = (100 / PINQ_DEC_GRAD_CTAS - 1) * 1024;  // CTAS x (request + 1 KB) fill the same 100 KB configuration
// Residency is capped at 5 CTAs per SM (TMEM columns) through shared memory: 5 x (19 + 1 KB) fill the 100 KB carve-out exactly
// and a sixth CTA does not fit.  The rest of the 256 KB stays L1 for the feature-row gathers (a 40 KB request under a 100 %
// carve-out did the same capping but left 28 KB of L1: decode 0.417 -> 0.407 ms).
constexpr size_t kSmemRequest = 19 * 1024;
constexpr int kCarveoutPct = 44;                    // of 227 KB -> the 100 KB configuration

struct Smem {
    alignas(128) unsigned char b[kChunksB * kChunkBBytes];
    alignas(16) float w2[H];  // W2 / 2 in accumulator-column order: relu(x) W2 = |x| W2/2 + x W2/2
    alignas(16) float u[12];  // sum_j (W2_j / 2) [W1 | b1]_j: the linear half, sum_j x_j W2_j / 2 = in . u
    // GRAD: B operand of the mask product, [16 hidden (mask-bit order)][n = input * 4 + digit, 16] x 16 bytes of int8
    alignas(128) unsigned char cq[(H / 16) * kGradN * 16];
    float gscale[4];          // 2^-s of d s / d v_i
    alignas(8) unsigned long long bar, bar2;
    uint32_t tmem_d, tmem_a, tmem_g;
};
static_assert(sizeof(Smem) <= kSmemRequest, "decode_tc shared memory");

// Accumulator column c <-> hidden unit.  The relu' bits of a row are two words; column 32 w + i lands on bit 31 - i of
// word w (32 funnel shifts), and the backward wants hidden unit 16 p + 8 h + 2 t + e on bit 8 (p & 1) + 4 h + t + 16 e of
// word p >> 1 (mma.cuh, mask_frags_rows) -- so the hidden units are permuted when B and W2 are staged.
__host__ __device__ constexpr int hidden_of_col(int c) {
    const int w = c >> 5, b = 31 - (c & 31);
    const int e = b >> 4, r = b & 15, pp = r >> 3, h = (r >> 2) & 1, t = r & 3;
    return 16 * (2 * w + pp) + 8 * h + 2 * t + e;
}

// instruction descriptor: D = f32, A = B = tf32, both K-major, N = 64, M = 128
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(H >> 3) << 17) | ((uint32_t)(kRows >> 4) << 24);

__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(kIdesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
        :: "r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(kIdesc), "r"(accumulate) : "memory");
}

// mask product of the gradient flavour: D = s32, A = B = signed 8 bit, K-major, N = 16, M = 128; A from TMEM
constexpr uint32_t kIdescG = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kGradN >> 3) << 17) | ((uint32_t)(kRows >> 4) << 24);
__device__ __forceinline__ void umma_i8_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}"
        :: "r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(kIdescG), "r"(accumulate) : "memory");
}
// K index kk of the mask product = bit kk of the row's 64 relu' bits (word kk >> 5, bit kk & 31) <-> accumulator column
__host__ __device__ constexpr int col_of_maskbit(int kk) { return 32 * (kk >> 5) + 31 - (kk & 31); }

}  // namespace tc

// GRAD (tracking, utils/tracker.py:230-253 + utils/tools.py:174): d sdf/dx in the same pass.  d s_k / d v_k = sum_j relu'_j W2_j
// W1[j, 8:11] is a SECOND tensor-core product per tile: the row's 64 relu' bits, expanded to 0 / 1 bytes, go into 16 TMEM
// columns as the A operand of a tcgen05.mma.kind::i8 against W2 W1[:, 8:11] in 32-bit fixed point (four digits per input,
// tc.cuh), exact integer sums recombined in fp32.  It is issued together with the NEXT tile's layer-1 product (one barrier per
// tile; one extra at the end of a block) and read one tile later.  The IDW part of the gradient needs the un-normalised
// weights and the TRUE x - p_k (hazard H13): re-derived per query from the search's rank workspace in finalize(), exactly like
// k_decode_mma<GRAD>.
template <bool MESH, bool GRAD>
__device__ __forceinline__ void decode_tc_body(const DecodeParams& p) {
    using namespace tc;
    if (p.skip && *p.skip) return;
    extern __shared__ __align__(128) unsigned char s_tc_raw[];
    Smem& S = *reinterpret_cast<Smem*>(s_tc_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // ---- once per CTA: B operand and W2 in accumulator-column order, barrier, TMEM (accumulator + A operand) ----
    for (int c = tid; c < H; c += kRows) {
        const int j = hidden_of_col(c);
        S.w2[c] = 0.5f * __ldg(&p.dec.W2[j]);
        uint32_t hi[12], lo[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) tf32_split(l1_operand(p.dec, i, j), hi[i], lo[i]);
        uint4* brow = reinterpret_cast<uint4*>(S.b + c * 16);
        brow[0 * (kChunkBBytes / 16)] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        brow[1 * (kChunkBBytes / 16)] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
        brow[2 * (kChunkBBytes / 16)] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        brow[3 * (kChunkBBytes / 16)] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
        brow[4 * (kChunkBBytes / 16)] = make_uint4(hi[8], hi[9], hi[10], hi[11]);
        brow[5 * (kChunkBBytes / 16)] = make_uint4(lo[8], lo[9], lo[10], lo[11]);
    }
    if (tid < 12) {  // accumulated in double: u carries the linear half of every neighbour's sdf
        double acc = 0.0;
        for (int j = 0; j < H; ++j) acc += 0.5 * (double)__ldg(&p.dec.W2[j]) * (double)l1_operand(p.dec, tid, j);
        S.u[tid] = (float)acc;
    }
    if (GRAD) {
        if (tid < 3) {  // fixed-point scale of d s / d v_i: 2^(30 - e) with max_j |W2_j W1[j][8 + i]| < 2^e
            float m = 0.f;
            for (int j = 0; j < H; ++j) m = fmaxf(m, fabsf(__ldg(&p.dec.W2[j]) * __ldg(&p.dec.W1[j * IN + 8 + tid])));
            int ec = floor_log2(m) + 1;
            ec = ec < -90 ? -90 : ec;
            S.gscale[tid] = pow2i(ec - 30);
        }
        __syncthreads();
        for (int e = tid; e < (H / 16) * kGradN; e += kRows) {
            const int kc = e / kGradN, n = e % kGradN, i = n >> 2, dg = n & 3;
            uint32_t w[4] = {0u, 0u, 0u, 0u};
            if (i < 3) {
                const float q = 1.f / S.gscale[i];  // a power of two: exact
#pragma unroll
                for (int b = 0; b < 16; ++b) {
                    const int j = hidden_of_col(col_of_maskbit(16 * kc + b));
                    const float v = __ldg(&p.dec.W2[j]) * __ldg(&p.dec.W1[j * IN + 8 + i]);
                    const uint32_t d = (digits_of(__float2int_rn(v * q)) >> (8 * dg)) & 0xffu;
                    w[b >> 2] |= d << (8 * (b & 3));
                }
            }
            *reinterpret_cast<uint4*>(S.cq + e * 16) = make_uint4(w[0], w[1], w[2], w[3]);
        }
    }
    if (tid == 0) {
        mbar_init(smem_u32(&S.bar), 1);
        mbar_init(smem_u32(&S.bar2), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     :: "r"(smem_u32(&S.tmem_d)), "r"(kTmemColsD) : "memory");
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     :: "r"(smem_u32(&S.tmem_a)), "r"(kTmemColsA) : "memory");
        if (GRAD)
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                         :: "r"(smem_u32(&S.tmem_g)), "r"(kTmemColsG) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_d = S.tmem_d, tmem_a = S.tmem_a, tmem_g = GRAD ? S.tmem_g : 0u;
    const uint64_t cdesc = umma_desc(smem_u32(S.cq), kGradN * 16, 128);  // K-major: next 16 hidden = 256 bytes, next 8 n = 128
    const uint32_t lanes = (uint32_t)(32 * warp) << 16;  // this warp's 32 TMEM lanes
    const uint64_t bdesc = umma_desc(smem_u32(S.b), kChunkBBytes, 128);
    const float b2 = __ldg(p.dec.b2), scale = p.dec.sdf_scale;
    const int64_t nblocks = (p.n + kRows - 1) / kRows;
    uint2* mask2 = reinterpret_cast<uint2*>(p.relu_mask);

    uint32_t phase = 0u, phase2 = 0u;
    float sdf = 0.f, sk[K], wk[K];
    float gv[GRAD ? K : 1][3];  // d s_k / d v_k before the sdf scale
#pragma unroll
    for (int k = 0; k < K; ++k) sk[k] = wk[k] = 0.f;
    // result of the mask product issued for tile kp (one tile, or one barrier, earlier)
    auto read_grad = [&](int kp) {
        mbar_wait(smem_u32(&S.bar2), phase2);
        phase2 ^= 1u;
        tc_fence_after();
        uint32_t d[16];
        tmem_ld16(tmem_g + lanes + 16, d);
        tmem_ld_wait(d);
        tc_fence_before();
#pragma unroll
        for (int i = 0; i < 3; ++i) gv[GRAD ? kp : 0][i] = undigit(d[4 * i], d[4 * i + 1], d[4 * i + 2], d[4 * i + 3]) * S.gscale[i];
    };

    // epilogue of tile k: the accumulator row of this thread (processing position ei)
    auto epilogue = [&](int k, int64_t ei, bool ewarp, float lin) {
        const StageBlock& sb = p.stage[ewarp ? (ei >> 5) : 0];
        const float w = ewarp ? __ldg(&sb.w[k][lane]) : 0.f;
        mbar_wait(smem_u32(&S.bar), phase);
        phase ^= 1u;
        tc_fence_after();
        // 16 columns at a time, the next chunk in flight while this one is reduced (32 live registers instead of 64)
        uint32_t va[16], vb[16];
        const uint32_t t0 = tmem_d + lanes;
        float s0 = lin, s1 = 0.f;
        uint32_t neg0 = 0u, neg1 = 0u;
        auto reduce16 = [&](const uint32_t (&v)[16], int c0, float& s, uint32_t& neg) {
#pragma unroll
            for (int c = 0; c < 16; c += 4) {
                const float4 w4 = *reinterpret_cast<const float4*>(&S.w2[c0 + c]);
                const float wv[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    // relu(x) W2 = |x| W2/2 + x W2/2: the first half is ONE FFMA per element (|x| is an operand modifier), the
                    // second is linear in the decoder inputs and was added once per tile (lin = in . u, 12 FFMAs).  The max
                    // and the sign-bit funnel shift both run on the half-rate ALU pipe, which bounded this epilogue (ncu r4a:
                    // pipe_alu 61 %, issue 63 %); now it is 2 instructions per element, one per pipe
                    s = fmaf(fabsf(__uint_as_float(v[c + u])), wv[u], s);
                    neg = __funnelshift_l(v[c + u], neg, 1);  // sign bits: 1 <-> relu' = 0
                }
            }
        };
        tmem_ld16(t0, va);
        tmem_ld_wait(va);
        tmem_ld16(t0 + 16, vb);
        reduce16(va, 0, s0, neg0);
        tmem_ld_wait(vb);
        tmem_ld16(t0 + 32, va);
        reduce16(vb, 16, s1, neg0);
        tmem_ld_wait(va);
        tmem_ld16(t0 + 48, vb);
        reduce16(va, 32, s0, neg1);
        tmem_ld_wait(vb);
        tc_fence_before();  // the accumulator may be overwritten after the next block-wide barrier
        reduce16(vb, 48, s1, neg1);
        if (mask2 && ewarp) mask2[((ei >> 5) * K + k) * 32 + lane] = make_uint2(~neg0, ~neg1);
        if (GRAD) {
            if (k > 0) read_grad(k - 1);  // ... which also frees the mask operand for this tile's bits
            uint32_t mb[16];  // 64 relu' bits -> 64 bytes of 0 / 1 (4 bits -> 4 bytes by one multiply)
#pragma unroll
            for (int t4 = 0; t4 < 16; ++t4)
                mb[t4] = ((((t4 < 8 ? ~neg0 : ~neg1) >> (4 * (t4 & 7))) & 0xfu) * 0x00204081u) & 0x01010101u;
            tmem_st16(tmem_g + lanes, mb);
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            tc_fence_before();
        }
        const float s = ((s0 + s1) + b2) * scale;  // decoder.py:80-81
        sk[k] = s;
        wk[k] = w;
        sdf = k == 0 ? s * w : fmaf(s, w, sdf);
    };
    // per-query outputs once the K tiles of a block are in
    auto finalize = [&](int64_t ei) {
        if (ei >= p.n) return;
        const int64_t q = p.perm ? (int64_t)p.perm[ei] : ei;
        if (MESH) {
            const int nnq = __ldg(&p.stage[ei >> 5].nn[lane]);
            if (p.mask_out) p.mask_out[q] = (uint8_t)(nnq >= p.min_nn);
            if (nnq == 0) {  // MLP result unused where nothing is near (mesher.py:96-103)
                p.sdf[q] = 0.f;
                return;
            }
        }
        if (p.sdf) p.sdf[q] = sdf;
        if (p.sdf_std) {  // tracker.py:246-247
            float var = 0.f;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                const float d = sk[k] - sdf;
                var = fmaf(wk[k], d * d, var);
            }
            p.sdf_std[q] = sqrtf(var);
        }
        if (GRAD && p.grad_x) {
            // d sdf/dx = sum_k w_k R_k^T (ds_k/dv_k) + (1/S) sum_k (s_k - sdf) d wt_k/dx,  d wt_k/dx = -2 wt_k^2 (x - p_k)
            float x, y, z;
            load_point(p.src, q, x, y, z);
            QueryLane L;
            resolve_query<true, false, false>(p, p.ws + ei * kWs, ei, q, true, x, y, z, L);
            float ox = 0.f, oy = 0.f, oz = 0.f;
            const float inv_s = 1.0f / L.wsum;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                if (L.kidx[k] >= 0) {
                    float ax = gv[GRAD ? k : 0][0] * scale, ay = gv[GRAD ? k : 0][1] * scale, az = gv[GRAD ? k : 0][2] * scale;
                    if (p.rotate) quat_rotate(__ldg(&p.quats[L.kidx[k]]), true, ax, ay, az);
                    const float c = (sk[k] - sdf) * (-2.f * L.wt[k] * L.wt[k]) * inv_s;
                    ox += wk[k] * ax + c * L.ux[k];
                    oy += wk[k] * ay + c * L.uy[k];
                    oz += wk[k] * az + c * L.uz[k];
                }
            }
            p.grad_x[3 * q] = ox;
            p.grad_x[3 * q + 1] = oy;
            p.grad_x[3 * q + 2] = oz;
        }
    };

    // Software pipeline of the gathers: the neighbour indices of a block are loaded one block ahead (coalesced), the
    // feature row and x - p_k of a tile one tile ahead, so staging never waits on a dependent load.
    int cidx[K], nidx[K];
    float4 pfa, pfb, phv;  // inputs of the next tile to stage
    auto load_idx = [&](int64_t bi, int (&idx)[K]) {
        const bool wl = (bi - lane) < p.n;
        const StageBlock& sb = p.stage[wl ? (bi >> 5) : 0];
#pragma unroll
        for (int k = 0; k < K; ++k) idx[k] = wl ? __ldg(&sb.idx[k][lane]) : -1;
    };
    auto load_tile = [&](int64_t bi, int k, int li) {
        pfa = make_float4(0.f, 0.f, 0.f, 0.f);
        pfb = pfa;
        phv = make_float4(0.f, 0.f, 0.f, 1.f);  // inputs 8..10 = x - p_k, input 11 = 1 (bias column)
        if (li >= 0) {
            const StageBlock& sb = p.stage[bi >> 5];
            ldg_row8(p.features + (int64_t)li * F, pfa, pfb);
            phv.x = __ldg(&sb.v[k][0][lane]);
            phv.y = __ldg(&sb.v[k][1][lane]);
            phv.z = __ldg(&sb.v[k][2][lane]);
        }
    };
    if ((int64_t)blockIdx.x < nblocks) {
        load_idx((int64_t)blockIdx.x * kRows + tid, cidx);
        load_tile((int64_t)blockIdx.x * kRows + tid, 0, cidx[0]);
    }
    for (int64_t blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
        const int64_t i = blk * kRows + tid;
        const bool wlive = (i - lane) < p.n;  // this warp has a stage block
        const int64_t ni = i + (int64_t)gridDim.x * kRows;
        const bool has_next = blk + gridDim.x < nblocks;
        if (tid == 0 && blk + 2 * (int64_t)gridDim.x < nblocks) {  // stage blocks of the block after the next -> L2
            const int64_t sb0 = (blk + 2 * (int64_t)gridDim.x) * (kRows / 32), nsb = (p.n + 31) / 32;
            const int64_t cnt = nsb - sb0 < kRows / 32 ? nsb - sb0 : kRows / 32;
            l2_prefetch(&p.stage[sb0], (uint32_t)(cnt * sizeof(StageBlock)));
        }
#pragma unroll
        for (int k = 0; k < K; ++k) nidx[k] = -1;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            float lin = 0.f;  // linear half of this tile's sdf
            // ---- stage tile (blk, k): thread = row = TMEM lane; one tcgen05.st of the 32 A columns
            //      [lo f0..3 | lo f4..7 | lo v,0 | hi v,1 | hi v,1 | 0 | hi f0..3 | hi f4..7]
            {
                const float4 fa = pfa, fb = pfb, hv = phv;
                {
                    // (re-read per tile through a volatile access: hoisted out of the loop the 12 values cost the registers
                    // that keep 5 CTAs on an SM)
                    const volatile float* su = S.u;
                    const float4 u0 = make_float4(su[0], su[1], su[2], su[3]), u1 = make_float4(su[4], su[5], su[6], su[7]),
                                 u2 = make_float4(su[8], su[9], su[10], su[11]);
                    float l = hv.w * u2.w;
                    l = fmaf(fa.x, u0.x, l); l = fmaf(fa.y, u0.y, l); l = fmaf(fa.z, u0.z, l); l = fmaf(fa.w, u0.w, l);
                    l = fmaf(fb.x, u1.x, l); l = fmaf(fb.y, u1.y, l); l = fmaf(fb.z, u1.z, l); l = fmaf(fb.w, u1.w, l);
                    l = fmaf(hv.x, u2.x, l); l = fmaf(hv.y, u2.y, l); l = fmaf(hv.z, u2.z, l);
                    lin = l;
                }
                auto lo_of = [](float x) { return __float_as_uint(x - __uint_as_float(__float_as_uint(x) & 0xffffe000u)); };
                uint32_t r[32];
                r[0] = lo_of(fa.x); r[1] = lo_of(fa.y); r[2] = lo_of(fa.z); r[3] = lo_of(fa.w);
                r[4] = lo_of(fb.x); r[5] = lo_of(fb.y); r[6] = lo_of(fb.z); r[7] = lo_of(fb.w);
                r[8] = lo_of(hv.x); r[9] = lo_of(hv.y); r[10] = lo_of(hv.z); r[11] = 0u;
                r[12] = __float_as_uint(hv.x); r[13] = __float_as_uint(hv.y); r[14] = __float_as_uint(hv.z);
                r[15] = __float_as_uint(hv.w);
                r[16] = r[12]; r[17] = r[13]; r[18] = r[14]; r[19] = r[15];
                r[20] = r[21] = r[22] = r[23] = 0u;
                r[24] = __float_as_uint(fa.x); r[25] = __float_as_uint(fa.y); r[26] = __float_as_uint(fa.z);
                r[27] = __float_as_uint(fa.w);
                r[28] = __float_as_uint(fb.x); r[29] = __float_as_uint(fb.y); r[30] = __float_as_uint(fb.z);
                r[31] = __float_as_uint(fb.w);
                tmem_st32(tmem_a + lanes, r);
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            }
            tc_fence_before();
            __syncthreads();
            if (tid == 0) {
                tc_fence_after();
                constexpr uint64_t cb = kChunkBBytes >> 4;  // descriptor units per B chunk
                umma_tf32_ts(tmem_d, tmem_a + 0, bdesc + 0 * cb, 0u);   // lo(f) x hi W
                umma_tf32_ts(tmem_d, tmem_a + 24, bdesc + 2 * cb, 1u);  // hi(f) x lo W
                umma_tf32_ts(tmem_d, tmem_a + 8, bdesc + 4 * cb, 1u);   // lo(v,0) | hi(v,1)  x  hi W ; lo W
                umma_tf32_ts(tmem_d, tmem_a + 24, bdesc + 0 * cb, 1u);  // hi(f) x hi W
                umma_tf32_ts(tmem_d, tmem_a + 16, bdesc + 4 * cb, 1u);  // hi(v,1) | 0  x  hi W ; -
                umma_commit(smem_u32(&S.bar));
                if (GRAD && k > 0) {  // mask of tile k - 1 (staged by its epilogue) x W2 W1[:, 8:11]: 2 x K = 32
                    umma_i8_ts(tmem_g + 16, tmem_g, cdesc, 0u);
                    umma_i8_ts(tmem_g + 16, tmem_g + 8, cdesc + (2 * kGradN * 16 >> 4), 1u);
                    umma_commit(smem_u32(&S.bar2));
                }
            }
            // ---- loads of the next tile while the MMAs run, then the epilogue of this one ----
            if (k == 1 && has_next) load_idx(ni, nidx);
            if (k + 1 < K) load_tile(i, k + 1, cidx[k + 1 < K ? k + 1 : 0]);
            else if (has_next) load_tile(ni, 0, nidx[0]);
            epilogue(k, i, wlive, lin);
        }
        if (GRAD) {  // the last tile's mask product has no next tile to ride on
            __syncthreads();
            if (tid == 0) {
                tc_fence_after();
                umma_i8_ts(tmem_g + 16, tmem_g, cdesc, 0u);
                umma_i8_ts(tmem_g + 16, tmem_g + 8, cdesc + (2 * kGradN * 16 >> 4), 1u);
                umma_commit(smem_u32(&S.bar2));
            }
            read_grad(K - 1);
        }
        finalize(i);
#pragma unroll
        for (int k = 0; k < K; ++k) cidx[k] = nidx[k];
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_d), "r"(kTmemColsD) : "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_a), "r"(kTmemColsA) : "memory");
        if (GRAD) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_g), "r"(kTmemColsG) : "memory");
    }
}

// (two entry points: a min-blocks hint on the plain flavours makes ptxas spill at 96 registers where it needs 92 without one)
template <bool MESH>
__global__ void __launch_bounds__(tc::kRows) k_decode_tc(const DecodeParams p) {
    decode_tc_body<MESH, false>(p);
}
__global__ void __launch_bounds__(tc::kRows, tc::kCtasPerSmGrad) k_decode_tc_grad(const DecodeParams p) {
    decode_tc_body<false, true>(p);
}

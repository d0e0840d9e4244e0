// bwd_tc.cuh -- backward of the fused query on tcgen05 / TMEM (included by query_bwd.cu inside namespace pinq).
//
// Same algebra as k_bwd_mma (query_bwd.cu header), other mapping: tile k = neighbour k of a block of 128 queries, so
// THREAD = QUERY = row.  Per tile the thread expands its 64 relu' bits into one row of the MASK tile and writes
// a_k [f_k, v_k, 1] into the X tile; both tiles are stored [16-byte chunk][row], which is the UMMA canonical no-swizzle
// layout K-major for (row x hidden) AND MN-major for its transpose:
//   mask x V    D1[128 rows x 32]   = MASK (A, K-major)  x V4 (B, K-major; 8 feature columns x 4 digits)      2 MMAs K=32
//   G           D2[64 hidden x 48] += MASK^T (A, MN-major) x X (B, MN-major), reduction over the 128 rows       4 MMAs K=32
//
// Operand format (r4): 8-bit INTEGERS (tcgen05.mma.kind::i8, s32 accumulators), not fp16.  Shared-memory bandwidth bounds this
// kernel (ncu r4a: tensor-core operand reads 37 % + tile stores 22 % of the L1TEX data pipe, 42 + 22 KB per tile with fp16
// operands); MASK holds exact 0 / 1, so one byte per element halves both its store and its two operand reads (24 + 14 KB
// per tile).  The real-valued operands become FIXED POINT: x -> round(x 2^s) as a 32-bit integer, split into four
// balanced base-256 digits d3 d2 d1 d0 (each a signed byte; the bytes of (I + 0x808080) ^ 0x808080), one digit per operand
// column; the tensor core's integer sums are EXACT and the digits are recombined in fp32 after the sum:
//   V (= W2 W1, per feature column c): 2^s chosen per column from max_j |V_jc|  ->  |quantisation| <= 2^-31 max|V_c| per term
//   X (= 2^s a_k [f_k, v_k, 1]):       2^s per block of 6 tiles from the block's largest |dL/dsdf| (the weights are <= 1 and
//       the decoder inputs are assumed < 2^8, as the fp16 tiles assumed before: a larger input overflows the 32-bit fixed
//       point and the gradients of that block are wrong) -> |quantisation| <= 2^-23 max|a| per
//       element, against the 2^-24 |partial sum| an fp32 accumulation of the same rows commits per addition.
// D1 -> thread-local epilogue: dL/df_k = a_k (digits recombined) 2^-s_c -> red.global.add.v4.f32 x 2 into grad_features[idx_k].
// D2 is flushed into fp32 shared-memory sums once per block of 6 tiles (768 rows) with the block's 2^-s, so the long
// reduction over rows is done with round-to-nearest adds outside the tensor core as in k_bwd_mma.
#pragma once
#include "tc.cuh"

namespace tcb {

constexpr int kRows = 128;
constexpr int kDigits = 4;                  // signed base-256 digits of a 32-bit fixed-point operand
constexpr int kXCols = 12 * kDigits;        // column = input * 4 + digit
constexpr uint32_t kChunk = kRows * 16;     // 2048 bytes: one 16-byte chunk (16 int8 elements) of all 128 rows
constexpr int kMaskChunks = H / 16;         // 4
constexpr int kXChunks = kXCols / 16;       // 3
constexpr int kVN = 8 * kDigits;            // n = feature column * 4 + digit
constexpr uint32_t kTmemCols = 128;         // D1 buffers at columns 0 and 32, D2 at column 64 (48)
#ifndef PINQ_BWD_CTAS
#define PINQ_BWD_CTAS 4
#endif
constexpr unsigned kCtasPerSm = PINQ_BWD_CTAS;  // 4 x 128 TMEM columns = all 512
constexpr size_t kSmemRequest = 31 * 1024;     // 4 x (31 + 1) KB fit the 132 KB configuration: 32 KB more L1 than the next one

struct Smem {  // the tiles are double-buffered: tile t + 1 is staged while the MMAs of tile t run
    alignas(128) unsigned char x[2][kXChunks * kChunk];
    alignas(128) unsigned char mask[2][kMaskChunks * kChunk];
    alignas(128) unsigned char v4[kMaskChunks * kVN * 16];  // [hidden chunk][n] x 16 bytes
    float vscale[8];                                        // 2^-s of feature column c
    alignas(8) unsigned long long bar[2];
    uint32_t tmem_base;
    float db2;
    float wmax[2][kRows / 32];  // per-warp max |dL/dsdf| of a block, double-buffered by block parity
};
static_assert(sizeof(Smem) <= kSmemRequest, "bwd_tc shared memory");

// column c of the MASK tile = bit c of the row's 64 relu' bits (word c >> 5, bit c & 31) <-> hidden unit
// (mma.cuh: hidden 16 p + 8 h + 2 t + e <-> bit 8 (p & 1) + 4 h + t + 16 e of word p >> 1)
__host__ __device__ constexpr int hidden_of_maskcol(int col) {
    const int w = col >> 5, e = (col >> 4) & 1, s = col & 15;
    return 16 * (2 * w + (s >> 3)) + 8 * ((s >> 2) & 1) + 2 * (s & 3) + e;
}

// D = s32, A = B = signed 8 bit
constexpr uint32_t kIdescI8 = (2u << 4) | (1u << 7) | (1u << 10);
constexpr uint32_t kIdescMV = kIdescI8 | ((uint32_t)(kVN >> 3) << 17) | ((uint32_t)(kRows >> 4) << 24);  // K-major x K-major
// decoder-gradient product: M = 64 hidden units, N = 48, both operands MN-major (= transposed views of the row-major tiles)
constexpr uint32_t kIdescG = kIdescI8 | (1u << 15) | (1u << 16) | ((uint32_t)(kXCols >> 3) << 17) | ((uint32_t)(H >> 4) << 24);
constexpr int kFixBits = 30;   // V: |fixed point| < 2^30
constexpr int kXFixBits = 21;  // X: max |a| 2^s in [2^21, 2^22), inputs < 2^8

__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
        :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" :: "l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
}  // namespace tcb

template <bool DEC, bool FEAT>
__global__ void __launch_bounds__(tcb::kRows) k_bwd_tc(const BwdParams p) {
    using namespace tc;
    using namespace tcb;
    extern __shared__ __align__(128) unsigned char s_tcb_raw[];
    Smem& S = *reinterpret_cast<Smem*>(s_tcb_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // ---- once per CTA ----
    if (FEAT) {
        // fixed-point scale of feature column c: 2^(30 - e_c) with max_j |W2_j W1_jc| < 2^e_c
        if (tid < 8) {
            float m = 0.f;
            for (int j = 0; j < H; ++j) m = fmaxf(m, fabsf(__ldg(&p.dec.W2[j]) * __ldg(&p.dec.W1[j * IN + tid])));
            int ec = floor_log2(m) + 1;
            ec = ec < -90 ? -90 : ec;
            S.vscale[tid] = pow2i(ec - kFixBits);
        }
        __syncthreads();
        // V4[hidden chunk][n = c * 4 + digit]: 16 hidden units (mask-column order) x int8
        for (int e = tid; e < kMaskChunks * kVN; e += kRows) {
            const int kc = e / kVN, n = e % kVN, c = n >> 2, dg = n & 3;
            const float q = 1.f / S.vscale[c];  // a power of two: exact
            uint32_t w[4] = {0u, 0u, 0u, 0u};
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const int j = hidden_of_maskcol(16 * kc + i);
                const float v = __ldg(&p.dec.W2[j]) * __ldg(&p.dec.W1[j * IN + c]);
                const uint32_t d = (digits_of(__float2int_rn(v * q)) >> (8 * dg)) & 0xffu;
                w[i >> 2] |= d << (8 * (i & 3));
            }
            *reinterpret_cast<uint4*>(S.v4 + e * 16) = make_uint4(w[0], w[1], w[2], w[3]);
        }
    }
    if (DEC) {
        if (tid == 0) S.db2 = 0.f;
    }
    if (tid == 0) {
        mbar_init(smem_u32(&S.bar[0]), 1);
        mbar_init(smem_u32(&S.bar[1]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     :: "r"(smem_u32(&S.tmem_base)), "r"(kTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = S.tmem_base;
    const uint32_t tmem_row = tmem + ((uint32_t)(32 * warp) << 16);
    const uint32_t kD2 = 64;  // first TMEM column of the G^T accumulator
    // K-major: LBO = next 16-byte chunk along K, SBO = next 8 rows.  MN-major: LBO = next 8 rows (K), SBO = next chunk.
    const uint64_t v4K = umma_desc(smem_u32(S.v4), kVN * 16, 128);
    const float scale = p.dec.sdf_scale;
    const int64_t nblocks = (p.n + kRows - 1) / kRows;
    const uint2* mask2 = reinterpret_cast<const uint2*>(p.relu_mask);
    uint32_t phase[2] = {0u, 0u};
    float db2 = 0.f;
    float vsc[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) vsc[c] = FEAT ? S.vscale[c] : 0.f;

    // software pipeline: indices and dL/dsdf one block ahead, tile inputs one tile ahead
    int cidx[K], nidx[K];
    float cg = 0.f, ng = 0.f;
    // two register sets, tile t uses set t & 1 and is loaded while tile t - 2 is multiplied
    uint2 pm[2];
    float pw[2];
    float4 pfa[2], pfb[2], phv[2];
    auto load_block = [&](int64_t bi, int (&idx)[K], float& g) {
        const bool wl = (bi - lane) < p.n;
        const StageBlock& sb = p.stage[wl ? (bi >> 5) : 0];
#pragma unroll
        for (int k = 0; k < K; ++k) idx[k] = wl ? __ldg(&sb.idx[k][lane]) : -1;
        g = 0.f;
        if (bi < p.n) g = p.grad_sdf[p.perm ? (int64_t)p.perm[bi] : bi];
    };
    auto load_tile = [&](int64_t bi, int k, int li) {
        const int st = k & 1;
        pm[st] = make_uint2(0u, 0u);
        pw[st] = 0.f;
        pfa[st] = make_float4(0.f, 0.f, 0.f, 0.f);
        pfb[st] = pfa[st];
        phv[st] = make_float4(0.f, 0.f, 0.f, 1.f);
        if (li >= 0) {
            const StageBlock& sb = p.stage[bi >> 5];
            pm[st] = __ldg(&mask2[((bi >> 5) * K + k) * 32 + lane]);
            pw[st] = __ldg(&sb.w[k][lane]);
            if (DEC) {
                ldg_row8(p.features + (int64_t)li * F, pfa[st], pfb[st]);
                phv[st].x = __ldg(&sb.v[k][0][lane]);
                phv[st].y = __ldg(&sb.v[k][1][lane]);
                phv[st].z = __ldg(&sb.v[k][2][lane]);
            }
        }
    };
    if ((int64_t)blockIdx.x < nblocks) {
        load_block((int64_t)blockIdx.x * kRows + tid, cidx, cg);
        load_tile((int64_t)blockIdx.x * kRows + tid, 0, cidx[0]);
        load_tile((int64_t)blockIdx.x * kRows + tid, 1, cidx[1]);
    }
    // epilogue of the tile that ran in buffer `buf`: D1 row of this thread -> dL/df of neighbour `li`
    float a_prev = 0.f;
    int li_prev = -1;
    auto epilogue = [&](int buf) {
        mbar_wait(smem_u32(&S.bar[buf]), phase[buf]);
        phase[buf] ^= 1u;
        tc_fence_after();
        if (FEAT) {  // columns c * 4 + digit
            uint32_t d0[16], d1[16];
            tmem_ld16(tmem_row + 32 * buf, d0);
            tmem_ld16(tmem_row + 32 * buf + 16, d1);
            tmem_ld_wait(d0);
            tmem_ld_wait(d1);
            if (li_prev >= 0) {
                float o[8];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    o[c] = undigit(d0[4 * c], d0[4 * c + 1], d0[4 * c + 2], d0[4 * c + 3]) * (vsc[c] * a_prev);
                    o[4 + c] = undigit(d1[4 * c], d1[4 * c + 1], d1[4 * c + 2], d1[4 * c + 3]) * (vsc[4 + c] * a_prev);
                }
                float* gf = p.grad_features + (int64_t)li_prev * F;
                red_add_v4(gf, o[0], o[1], o[2], o[3]);
                red_add_v4(gf + 4, o[4], o[5], o[6], o[7]);
            }
        }
        tc_fence_before();
    };
    // G of a finished block (all its MMAs waited for) -> fp32 sums.  M = 64 accumulator layout: hidden row m sits in
    // TMEM lane (m % 16) + 32 (m / 16), i.e. lanes 0..15 of every warp (cute/atom/mma_traits_sm100.hpp, 1SM M_MMA == 64)
    float gacc[12];  // flushed G sums of hidden row 16 warp + lane (lanes 0..15), one per decoder input
#pragma unroll
    for (int i = 0; i < 12; ++i) gacc[i] = 0.f;
    auto flush = [&](float inv_scale) {
        if (DEC) {
#pragma unroll
            for (int c3 = 0; c3 < kXCols / 16; ++c3) {
                uint32_t v[16];
                tmem_ld16(tmem_row + kD2 + 16 * c3, v);
                tmem_ld_wait(v);
#pragma unroll
                for (int u = 0; u < 4; ++u)  // (meaningful in lanes 0..15: hidden row 16 warp + lane)
                    gacc[4 * c3 + u] += undigit(v[4 * u], v[4 * u + 1], v[4 * u + 2], v[4 * u + 3]) * inv_scale;
            }
            tc_fence_before();
        }
    };
    // Power-of-two scale of a block's X tiles: 2^s with 2^s * scale * max_rows |dL/dsdf| in [2^21, 2^22) -- |w_k| <= 1 and the
    // decoder inputs are below 2^8, so the fixed-point values stay below 2^30.  Each warp
    // publishes its max one block ahead (the value arrives with the block's index loads).
    auto publish_max = [&](float g, int parity) {
        float m = fabsf(g);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (lane == 0) S.wmax[parity][warp] = m;
    };
    auto block_scale = [&](int parity, float& up, float& down) {
        const float m = fmaxf(fmaxf(S.wmax[parity][0], S.wmax[parity][1]), fmaxf(S.wmax[parity][2], S.wmax[parity][3])) * fabsf(scale);
        const int e = floor_log2(m);
        up = pow2i(kXFixBits - e);
        down = pow2i(e - kXFixBits);
    };
    float xscale = 1.f, xinv = 1.f, xinv_prev = 1.f;
    int parity = 0;
    if (DEC) {
        publish_max(cg, 0);
        __syncthreads();
    }
    bool pending = false;  // the last tile of the previous block is still in flight
    for (int64_t blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
        if (DEC) {
            xinv_prev = xinv;
            block_scale(parity, xscale, xinv);
        }
        const int64_t i = blk * kRows + tid;
        const int64_t ni = i + (int64_t)gridDim.x * kRows;
        const bool has_next = blk + gridDim.x < nblocks;
        if (tid == 0 && blk + 2 * (int64_t)gridDim.x < nblocks) {  // stage blocks + relu' masks of the block after the next -> L2
            const int64_t sb0 = (blk + 2 * (int64_t)gridDim.x) * (kRows / 32), nsb = (p.n + 31) / 32;
            const int64_t cnt = nsb - sb0 < kRows / 32 ? nsb - sb0 : kRows / 32;
            l2_prefetch(&p.stage[sb0], (uint32_t)(cnt * sizeof(StageBlock)));
            l2_prefetch(&mask2[sb0 * K * 32], (uint32_t)(cnt * K * 32 * sizeof(uint2)));
        }
#pragma unroll
        for (int k = 0; k < K; ++k) nidx[k] = -1;
        ng = 0.f;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int li = cidx[k];
            const int st = k & 1;  // register set, tile buffer, D1 buffer and barrier of this tile (K is even)
            const float a = li >= 0 ? cg * pw[st] * scale : 0.f;  // rows without a neighbour carry no gradient
            // ---- MASK tile: 64 bits -> 64 bytes of 0 / 1 -> 4 chunks of 16 bytes (4 bits -> 4 bytes by one multiply) ----
            {
                const uint2 m = li >= 0 ? pm[st] : make_uint2(0u, 0u);
#pragma unroll
                for (int hc = 0; hc < kMaskChunks; ++hc) {
                    const uint32_t h16 = ((hc < 2 ? m.x : m.y) >> (16 * (hc & 1))) & 0xffffu;
                    uint32_t r[4];
#pragma unroll
                    for (int u4 = 0; u4 < 4; ++u4) r[u4] = (((h16 >> (4 * u4)) & 0xfu) * 0x00204081u) & 0x01010101u;
                    *reinterpret_cast<uint4*>(S.mask[st] + hc * kChunk + tid * 16) = make_uint4(r[0], r[1], r[2], r[3]);
                }
            }
            // ---- X tile: round(2^s a_k [f_k, v_k, 1]) as 4 signed digits, the 4 bytes of word `input` ----
            if (DEC) {
                db2 += a;
                const float as = a * xscale;
                const float ps[12] = {pfa[st].x, pfa[st].y, pfa[st].z, pfa[st].w, pfb[st].x, pfb[st].y, pfb[st].z, pfb[st].w,
                                      phv[st].x, phv[st].y, phv[st].z, phv[st].w};
                uint32_t wds[12];
#pragma unroll
                for (int c = 0; c < 12; ++c) wds[c] = digits_of(__float2int_rn(ps[c] * as));
#pragma unroll
                for (int c = 0; c < kXChunks; ++c)
                    *reinterpret_cast<uint4*>(S.x[st] + c * kChunk + tid * 16) =
                        make_uint4(wds[4 * c], wds[4 * c + 1], wds[4 * c + 2], wds[4 * c + 3]);
            }
            if (k == 0 && pending) {  // drain the previous block before its G^T accumulator is overwritten
                epilogue(1);
                flush(xinv_prev);
            }
            fence_async_smem();
            tc_fence_before();
            __syncthreads();
            if (tid == 0) {
                tc_fence_after();
                const uint32_t ma = smem_u32(S.mask[st]);
                if (FEAT) {
                    const uint64_t maskK = umma_desc(ma, kChunk, 128);
#pragma unroll
                    for (int ks = 0; ks < 2; ++ks)  // 32 hidden units = 2 chunks per step
                        umma_i8(tmem + 32 * st, maskK + (uint64_t)ks * (2 * kChunk >> 4), v4K + (uint64_t)ks * (2 * kVN * 16 >> 4),
                                kIdescMV, ks > 0);
                }
                if (DEC) {
                    const uint64_t xMN = umma_desc(smem_u32(S.x[st]), 128, kChunk);
                    const uint64_t maskMN = umma_desc(ma, 128, kChunk);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks)  // 32 rows = 512 bytes per step
                        umma_i8(tmem + kD2, maskMN + (uint64_t)ks * (512 >> 4), xMN + (uint64_t)ks * (512 >> 4), kIdescG,
                                (k > 0) | (ks > 0));
                }
                umma_commit(smem_u32(&S.bar[st]));
            }
            // ---- loads of the tile after the next, then the epilogue of the previous tile while the MMAs run ----
            if (k == 0 && has_next) load_block(ni, nidx, ng);
            if (DEC && k == 3) publish_max(ng, parity ^ 1);  // read at the start of the next block, two barriers from here
            if (k + 2 < K) load_tile(i, k + 2, cidx[k + 2 < K ? k + 2 : 0]);
            else if (has_next) load_tile(ni, k + 2 - K, nidx[k + 2 >= K ? k + 2 - K : 0]);
            if (k > 0) epilogue(st ^ 1);
            a_prev = a;
            li_prev = li;
        }
        pending = true;
        parity ^= 1;
#pragma unroll
        for (int k = 0; k < K; ++k) cidx[k] = nidx[k];
        cg = ng;
    }
    if (pending) {
        epilogue(1);
        flush(xinv);  // the last block's own scale
    }

    if (DEC) {
        db2 = warp_sum(db2);
        if (lane == 0) atomicAdd(&S.db2, db2);
    }
    __syncthreads();
    if (DEC) {  // G[j][i]; then dW1 = W2 G, db1 = W2 G[:,11], dW2[j] = W1b[j] . G[j], db2
        float* sG = reinterpret_cast<float*>(S.mask[0]);  // [H][12], the tiles are idle now
        if (lane < 16) {
            const int j = hidden_of_maskcol(16 * warp + lane);
#pragma unroll
            for (int i = 0; i < 12; ++i) sG[j * 12 + i] = gacc[i];
        }
        __syncthreads();
        for (int e = tid; e < H * 12; e += kRows) {
            const int j = e / 12, i = e - 12 * j;
            const float v = __ldg(&p.dec.W2[j]) * sG[e];
            if (i < IN) atomicAdd(&p.grad_W1[j * IN + i], v);
            else atomicAdd(&p.grad_b1[j], v);
        }
        for (int j = tid; j < H; j += kRows) {
            float acc = __ldg(&p.dec.b1[j]) * sG[j * 12 + IN];
#pragma unroll
            for (int i = 0; i < IN; ++i) acc = fmaf(__ldg(&p.dec.W1[j * IN + i]), sG[j * 12 + i], acc);
            atomicAdd(&p.grad_W2[j], acc);
        }
        if (tid == 0) atomicAdd(p.grad_b2, S.db2);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"(kTmemCols) : "memory");
    }
}

// decode_tc.cuh -- fused decode on the 5th-generation tensor cores: tcgen05.mma with the accumulator in TMEM.
// (included by query.cu after DecodeParams / StageBlock; sm_100a only)
//
// model/decoder.py:66-88 for the K neighbours of a block of 128 queries = K products [128 x 12] x [12 x 64]:
// tile k holds neighbour k of the 128 queries, so THREAD = QUERY = accumulator row (TMEM lane) in every tile and the
// whole epilogue (relu, . W2, relu' bits, IDW combine, std) is thread-local -- no fragment shuffles, no MMA issue slots
// in the CUDA-core instruction stream (ncu on the mma.sync formulation: tensor pipe and issue both ~50 % busy, stalled on
// each other's latencies at 16 warps/SM).
//
//   stage   thread = row: 32-byte feature row + (x - p_k) from the stage block -> 3xTF32 pieces -> six 16-byte chunks of
//           the A tile in shared memory (UMMA canonical K-major layout without swizzle: chunk c of row r at c*2048 + r*16)
//   mma     one thread: 5 x tcgen05.mma.kind::tf32 M=128 N=64 K=8 (small terms first), tcgen05.commit -> mbarrier
//              [lo(f0..7)] x [hi W]   [hi(f0..7)] x [lo W]   [lo(v,1) | hi(v,1)] x [hi W ; lo W]
//              [hi(f0..7)] x [hi W]   [hi(v,1) | 0] x [hi W ; -]
//           `hi` is the raw fp32 word (the tensor core ignores the low 13 mantissa bits), lo = x - trunc(x) is exact
//   epilog  tcgen05.ld 32x32b: 64 pre-activations of the thread's row; relu . W2 with W2 broadcast from shared memory;
//           relu' bits collected from the SIGN of the pre-activation with one funnel shift per element
// A tiles and accumulators are double-buffered: the MMAs of tile t run while the epilogue of tile t-1 executes.
//
// Deviation (documented in DESIGN.md): relu'(+0.0) is taken as 1 (torch: 0); reachable only when a hidden
// pre-activation is exactly +0.0.
#pragma once
#include "tc.cuh"

// (inside namespace pinq)
namespace tc {

constexpr int kRows = 128;                          // UMMA M = queries per CTA block = threads per CTA
constexpr int kChunksA = 7;                         // [L0 L1 L2 H2 Z H0 H1], 16 bytes (4 tf32) per row each
constexpr int kChunksB = 6;                         // [h0 h1 l0 l1 h2 l2] per hidden unit
constexpr uint32_t kChunkABytes = kRows * 16;       // 2048: K-direction (leading) byte offset of A
constexpr uint32_t kChunkBBytes = H * 16;           // 1024: K-direction (leading) byte offset of B
constexpr uint32_t kTmemCols = 2 * H;               // two accumulator buffers of 64 fp32 columns
constexpr size_t kSmemRequest = 48 * 1024;          // > 227 KB / 5: at most 4 CTAs (4 x 128 = 512 TMEM columns) per SM

struct Smem {
    alignas(128) unsigned char a[2][kChunksA * kChunkABytes];
    alignas(128) unsigned char b[kChunksB * kChunkBBytes];
    alignas(16) float w2[H];  // W2 in accumulator-column order
    alignas(8) unsigned long long bar[2];
    uint32_t tmem_base;
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
}  // namespace tc

template <bool MESH>
__global__ void __launch_bounds__(tc::kRows) k_decode_tc(const DecodeParams p) {
    using namespace tc;
    extern __shared__ __align__(128) unsigned char s_tc_raw[];
    Smem& S = *reinterpret_cast<Smem*>(s_tc_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // ---- once per CTA: B operand and W2 in accumulator-column order, zero chunk, barriers, TMEM ----
    for (int c = tid; c < H; c += kRows) {
        const int j = hidden_of_col(c);
        S.w2[c] = __ldg(&p.dec.W2[j]);
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
#pragma unroll
    for (int b = 0; b < 2; ++b) *reinterpret_cast<uint4*>(S.a[b] + 4 * kChunkABytes + tid * 16) = make_uint4(0u, 0u, 0u, 0u);
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
    const uint32_t tmem_row = tmem + ((uint32_t)(32 * warp) << 16);  // this warp's 32 TMEM lanes
    const uint64_t bdesc = umma_desc(smem_u32(S.b), kChunkBBytes, 128);
    const float b2 = __ldg(p.dec.b2), scale = p.dec.sdf_scale;
    const int64_t nblocks = (p.n + kRows - 1) / kRows;
    uint2* mask2 = reinterpret_cast<uint2*>(p.relu_mask);

    uint32_t phase[2] = {0u, 0u};
    float sdf = 0.f, sk[K], wk[K];
#pragma unroll
    for (int k = 0; k < K; ++k) sk[k] = wk[k] = 0.f;

    // epilogue of tile k of the block whose row of this thread is processing position ei
    auto epilogue = [&](int k, int64_t ei, bool ewarp) {
        const int buf = k & 1;
        const StageBlock& sb = p.stage[ewarp ? (ei >> 5) : 0];
        const float w = ewarp ? __ldg(&sb.w[k][lane]) : 0.f;
        mbar_wait(smem_u32(&S.bar[buf]), phase[buf]);
        phase[buf] ^= 1u;
        tc_fence_after();
        // 16 columns at a time, the next chunk in flight while this one is reduced (32 live registers instead of 64)
        uint32_t va[16], vb[16];
        const uint32_t t0 = tmem_row + (uint32_t)(buf * H);
        float s0 = 0.f, s1 = 0.f;
        uint32_t neg0 = 0u, neg1 = 0u;
        auto reduce16 = [&](const uint32_t (&v)[16], int c0, float& s, uint32_t& neg) {
#pragma unroll
            for (int c = 0; c < 16; c += 4) {
                const float4 w4 = *reinterpret_cast<const float4*>(&S.w2[c0 + c]);
                const float wv[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    s = fmaf(fmaxf(__uint_as_float(v[c + u]), 0.f), wv[u], s);
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
        tc_fence_before();  // the accumulator buffer may be overwritten after the next block-wide barrier
        reduce16(vb, 48, s1, neg1);
        if (mask2 && ewarp) mask2[((ei >> 5) * K + k) * 32 + lane] = make_uint2(~neg0, ~neg1);
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
            const float4* fr = reinterpret_cast<const float4*>(p.features + (int64_t)li * F);
            pfa = __ldg(fr);
            pfb = __ldg(fr + 1);
            phv.x = __ldg(&sb.v[k][0][lane]);
            phv.y = __ldg(&sb.v[k][1][lane]);
            phv.z = __ldg(&sb.v[k][2][lane]);
        }
    };
    int64_t pend_i = -1;  // processing position (of this thread) of the block whose last tile is still in flight
    bool pend_warp = false;
    if ((int64_t)blockIdx.x < nblocks) {
        load_idx((int64_t)blockIdx.x * kRows + tid, cidx);
        load_tile((int64_t)blockIdx.x * kRows + tid, 0, cidx[0]);
    }
    for (int64_t blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
        const int64_t i = blk * kRows + tid;
        const bool wlive = (i - lane) < p.n;  // this warp has a stage block
        const int64_t ni = i + (int64_t)gridDim.x * kRows;
        const bool has_next = blk + gridDim.x < nblocks;
#pragma unroll
        for (int k = 0; k < K; ++k) nidx[k] = -1;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int buf = k & 1;  // K is even: the tile parity is the same in every block
            // ---- stage tile (blk, k): thread = row ----
            const float4 fa = pfa, fb = pfb, hv = phv;
            auto lo_of = [](float x) { return x - __uint_as_float(__float_as_uint(x) & 0xffffe000u); };
            unsigned char* arow = S.a[buf] + tid * 16;
            *reinterpret_cast<float4*>(arow + 0 * kChunkABytes) = make_float4(lo_of(fa.x), lo_of(fa.y), lo_of(fa.z), lo_of(fa.w));
            *reinterpret_cast<float4*>(arow + 1 * kChunkABytes) = make_float4(lo_of(fb.x), lo_of(fb.y), lo_of(fb.z), lo_of(fb.w));
            *reinterpret_cast<float4*>(arow + 2 * kChunkABytes) = make_float4(lo_of(hv.x), lo_of(hv.y), lo_of(hv.z), 0.f);
            *reinterpret_cast<float4*>(arow + 3 * kChunkABytes) = hv;
            *reinterpret_cast<float4*>(arow + 5 * kChunkABytes) = fa;
            *reinterpret_cast<float4*>(arow + 6 * kChunkABytes) = fb;
            fence_async_smem();  // generic-proxy writes -> visible to the tensor core's async proxy
            tc_fence_before();
            __syncthreads();
            if (tid == 0) {
                tc_fence_after();
                const uint64_t adesc = umma_desc(smem_u32(S.a[buf]), kChunkABytes, 128);
                const uint32_t d = tmem + (uint32_t)(buf * H);
                constexpr uint64_t ca = kChunkABytes >> 4, cb = kChunkBBytes >> 4;  // descriptor units per chunk
                umma_tf32(d, adesc + 0 * ca, bdesc + 0 * cb, 0u);  // lo(f) x hi W
                umma_tf32(d, adesc + 5 * ca, bdesc + 2 * cb, 1u);  // hi(f) x lo W
                umma_tf32(d, adesc + 2 * ca, bdesc + 4 * cb, 1u);  // lo(v,1) | hi(v,1)  x  hi W ; lo W
                umma_tf32(d, adesc + 5 * ca, bdesc + 0 * cb, 1u);  // hi(f) x hi W
                umma_tf32(d, adesc + 3 * ca, bdesc + 4 * cb, 1u);  // hi(v,1) | 0  x  hi W ; -
                umma_commit(smem_u32(&S.bar[buf]));
            }
            // ---- loads of the next tile, then the epilogue of the previous one while the MMAs run ----
            if (k == 1 && has_next) load_idx(ni, nidx);
            if (k + 1 < K) load_tile(i, k + 1, cidx[k + 1 < K ? k + 1 : 0]);
            else if (has_next) load_tile(ni, 0, nidx[0]);
            if (k > 0) {
                epilogue(k - 1, i, wlive);
            } else if (pend_i >= 0) {
                epilogue(K - 1, pend_i, pend_warp);
                finalize(pend_i);
            }
        }
        pend_i = i;
        pend_warp = wlive;
#pragma unroll
        for (int k = 0; k < K; ++k) cidx[k] = nidx[k];
    }
    if (pend_i >= 0) {
        epilogue(K - 1, pend_i, pend_warp);
        finalize(pend_i);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"(kTmemCols) : "memory");
    }
}


// mma.cuh -- warp-level tensor-core building blocks of the decoder kernels.
//
// The 11 -> 64 -> 1 decoder (model/decoder.py:66-88) evaluated for 6 neighbours of 32 queries is a
// [192 x 12] x [12 x 64] product per warp; ncu showed the FP32-FMA formulation pipe-bound (profiles/r1c_*),
// so the three contractions of the path run on the tensor cores with fp32-equivalent accuracy:
//   layer 1      pre  = [f_k, v_k, 1] W1b^T          3xTF32 (hi*hi + hi*lo + lo*hi), m16n8k8
//   mask x V     d s_k / d in_k = relu'(pre) (W2 . W1)   the mask is exact in bf16, V split in 3 bf16, m16n8k16
//   mask^T x X   decoder gradients (reduction over rows)  mask exact, X = a_k in_k split in 3 bf16, m16n8k16
// Everything stays in registers: the layer-1 accumulator fragment IS the A fragment of the mask products
// (same trick as the P.V product of fused attention), so the hidden layer never touches memory.
//
// Fragment coordinates (PTX ISA, mma.m16n8k8 / m16n8k16): g = lane >> 2, t = lane & 3.
//   C (16x8 f32)      c0 (g, 2t)  c1 (g, 2t+1)  c2 (g+8, 2t)  c3 (g+8, 2t+1)
//   A tf32 (16x8)     a0 (g, t)   a1 (g+8, t)   a2 (g, t+4)   a3 (g+8, t+4)
//   B tf32 (8x8)      b0 (k=t, n=g)   b1 (k=t+4, n=g)
//   A bf16 (16x16)    a0 (g, 2t..2t+1)  a1 (g+8, 2t..)  a2 (g, 2t+8..)  a3 (g+8, 2t+8..)   (low half = lower k)
//   B bf16 (16x8)     b0 (k=2t..2t+1, n=g)   b1 (k=2t+8..2t+9, n=g)
#pragma once
#include <cuda_bf16.h>

#include "query_common.cuh"

namespace pinq {

__device__ __forceinline__ uint32_t tf32_rna(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}
// x = hi + lo + O(2^-22 |x|), both exactly representable in tf32
__device__ __forceinline__ void tf32_split(float x, uint32_t& hi, uint32_t& lo) {
    hi = tf32_rna(x);
    lo = tf32_rna(x - __uint_as_float(hi));
}

// Hot-loop variant: the tensor core ignores the low 13 mantissa bits of a tf32 operand, so x itself serves as
// `hi` (= x truncated) and lo = x - trunc(x) is exact; lo is truncated again by the hardware, leaving
// O(2^-21 |x|).  Two instructions (LOP3 + FADD) instead of the ~9 the emulated cvt.rna pair costs (SASS).
__device__ __forceinline__ void tf32_split_trunc(float x, uint32_t& hi, uint32_t& lo) {
    hi = __float_as_uint(x);
    lo = __float_as_uint(x - __uint_as_float(hi & 0xffffe000u));
}

// x = b1 + b2 + b3 + O(2^-24 |x|)
__device__ __forceinline__ void bf16_split3(float x, __nv_bfloat16& b1, __nv_bfloat16& b2, __nv_bfloat16& b3) {
    b1 = __float2bfloat16_rn(x);
    const float r1 = x - __bfloat162float(b1);
    b2 = __float2bfloat16_rn(r1);
    const float r2 = r1 - __bfloat162float(b2);
    b3 = __float2bfloat16_rn(r2);
}
// Hot-loop variant for a PAIR of values: pieces by truncation (x & 0xffff0000 is a bf16; the remainders are exact), so
// each piece of the pair is one PRMT of the two high halves: a = p[0].lo + p[1].lo + p[2].lo + O(2^-24 |a|), b likewise
// in the high halves.  2 LOP3 + 2 FADD per value + 3 PRMT per pair instead of ~10 instructions per value (F2F based).
__device__ __forceinline__ void bf16_split3_pair(float a, float b, uint32_t (&p)[3]) {
    p[0] = __byte_perm(__float_as_uint(a), __float_as_uint(b), 0x7632);
    const float a1 = a - __uint_as_float(__float_as_uint(a) & 0xffff0000u);
    const float b1 = b - __uint_as_float(__float_as_uint(b) & 0xffff0000u);
    p[1] = __byte_perm(__float_as_uint(a1), __float_as_uint(b1), 0x7632);
    const float a2 = a1 - __uint_as_float(__float_as_uint(a1) & 0xffff0000u);
    const float b2 = b1 - __uint_as_float(__float_as_uint(b1) & 0xffff0000u);
    p[2] = __byte_perm(__float_as_uint(a2), __float_as_uint(b2), 0x7632);
}
__device__ __forceinline__ uint32_t pack_bf16(__nv_bfloat16 lo, __nv_bfloat16 hi) {
    return (uint32_t)__bfloat16_as_ushort(lo) | ((uint32_t)__bfloat16_as_ushort(hi) << 16);
}

__device__ __forceinline__ void mma_tf32(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
    asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
    asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// transpose of an 8x8 b16 matrix held one 32-bit register per lane (row g, columns 2t..2t+1)
__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t a) {
    uint32_t d;
    asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
    return d;
}

constexpr int kRowsPerWarp = 32 * K;  // 192 rows = 12 m-tiles of 16
constexpr int kMTiles = kRowsPerWarp / 16;
constexpr uint32_t kBf16One = 0x3F80u;

// Decoder operands in fragment order, built once per CTA in shared memory.
struct DecFrags {
    // layer 1 B operands of n-tile nt (j = 8 nt + g) as the register PAIRS the five concatenated k-steps consume:
    //   w1a = {hi B[t][j], hi B[t+4][j] | hi B[8+t][j], lo B[t][j]}      k-steps 0/3 and 1
    //   w1b = {lo B[t+4][j], lo B[8+t][j] | hi B[8+t][j], 0}             k-steps 2 and 4     (B[11][j] = b1[j])
    float4 w1a[8][32];
    float4 w1b[8][32];
    float2 w2[32];        // {W2[2i], W2[2i+1]}
    uint2 vf[4][3][32];   // mask x V, columns = feature inputs 0..7:   split s of {V[16p+2t..][g], V[16p+2t+8..][g]}
    uint2 vv[4][3][32];   // mask x V, columns = inputs 8..10 (x - p_k), zero-padded to 8
};

// B[i][j] of layer 1: W1[j][i] for i < 11, b1[j] for i = 11
__device__ __forceinline__ float l1_operand(const pinq_decoder_view& d, int i, int j) {
    return i < IN ? d.W1[j * IN + i] : (i == IN ? d.b1[j] : 0.f);
}

__device__ __forceinline__ void build_dec_frags(const pinq_decoder_view& d, DecFrags* s, bool want_l1, bool want_vf,
                                                bool want_vv) {
    for (int e = threadIdx.x; e < 8 * 32; e += blockDim.x) {
        const int nt = e >> 5, lane = e & 31, g = lane >> 2, t = lane & 3, j = 8 * nt + g;
        if (want_l1) {
            uint32_t h0, l0, h1, l1, h2, l2;
            tf32_split(l1_operand(d, t, j), h0, l0);
            tf32_split(l1_operand(d, t + 4, j), h1, l1);
            tf32_split(l1_operand(d, t + 8, j), h2, l2);
            s->w1a[nt][lane] = make_float4(__uint_as_float(h0), __uint_as_float(h1), __uint_as_float(h2), __uint_as_float(l0));
            s->w1b[nt][lane] = make_float4(__uint_as_float(l1), __uint_as_float(l2), __uint_as_float(h2), 0.f);
        }
    }
    for (int e = threadIdx.x; e < 32; e += blockDim.x) s->w2[e] = make_float2(d.W2[2 * e], d.W2[2 * e + 1]);
    for (int e = threadIdx.x; e < 4 * 32; e += blockDim.x) {
        const int p = e >> 5, lane = e & 31, g = lane >> 2, t = lane & 3;
#pragma unroll
        for (int which = 0; which < 2; ++which) {
            if (which == 0 ? !want_vf : !want_vv) continue;
            const int col = which == 0 ? g : 8 + g;  // decoder input column this lane's n index maps to
            __nv_bfloat16 sp[4][3];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int j = 16 * p + 2 * t + (r & 1) + 8 * (r >> 1);
                const float v = col < IN ? d.W2[j] * d.W1[j * IN + col] : 0.f;
                bf16_split3(v, sp[r][0], sp[r][1], sp[r][2]);
            }
#pragma unroll
            for (int sidx = 0; sidx < 3; ++sidx) {
                const uint2 v = make_uint2(pack_bf16(sp[0][sidx], sp[1][sidx]), pack_bf16(sp[2][sidx], sp[3][sidx]));
                if (which == 0) s->vf[p][sidx][lane] = v;
                else s->vv[p][sidx][lane] = v;
            }
        }
    }
}

// relu'(pre) bits of one (query, neighbour) row = two words (the saved mask is uint2[32-query block][K][query]):
// hidden unit 16 p + 8 h + 2 t + e  <->  bit 8 (p & 1) + 4 h + t + 16 e of word p >> 1, i.e. the two elements of one
// bf16 A-fragment register sit 16 bits apart and shift + and + multiply by bf16(1) expands them.
// Fragments of one 16-row m-tile (4 k-steps of 16 hidden units) for lane (g, t) from the words of rows g and g + 8:
__device__ __forceinline__ void mask_frags_rows(uint2 ra, uint2 rb, int t, uint32_t (&a)[4][4]) {
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const uint32_t wa = (p >> 1) ? ra.y : ra.x, wb = (p >> 1) ? rb.y : rb.x;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int sh = 8 * (p & 1) + 4 * h + t;
            a[p][2 * h] = ((wa >> sh) & 0x00010001u) * kBf16One;      // row g,     hidden 16 p + 8 h + 2 t (+1)
            a[p][2 * h + 1] = ((wb >> sh) & 0x00010001u) * kBf16One;  // row g + 8
        }
    }
}
// position (before the shift by t) of accumulator acc[nt][c] of the mma.sync decode in its row's word nt >> 2
__device__ __forceinline__ constexpr int mask_row_bit0(int nt, int c) { return 4 * (nt & 3) + 16 * (c & 1); }

}  // namespace pinq

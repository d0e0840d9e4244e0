// common.cuh -- shared device/host helpers of libpinq (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "pinq.h"

namespace pinq {

// model/neural_points.py:69
constexpr long long kP1 = 73856093LL;
constexpr long long kP2 = 19349669LL;
constexpr long long kP3 = 83492791LL;

void set_error(const char* fmt, ...);
int check_launch(const char* what);
void ktime_mark(int mark, cudaStream_t st);  // per-kernel timing marks (PINQ_KT_*), no-op unless enabled
void launch_tile_scan(int32_t* tile_sums, int64_t ntiles, int32_t* d_count, cudaStream_t st);
// csrc/sort.cu: perm[] = queries binned by 4x4x4-voxel brick (processing order -> query index)
int64_t bin_scratch_bytes(int64_t n);
int launch_spatial_bin(const float* xyz, int64_t n, float res, bool recip, void* scratch, int32_t* perm, cudaStream_t st);

#define PINQ_REQUIRE(cond, msg)                              \
    do {                                                     \
        if (!(cond)) {                                       \
            pinq::set_error("%s: %s", __func__, msg);        \
            return PINQ_ERR_BAD_ARG;                         \
        }                                                    \
    } while (0)

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

static inline unsigned grid_for(int64_t n, int block) {
    int64_t g = (n + block - 1) / block;
    if (g < 1) g = 1;
    return (unsigned)g;
}

// floor(x / res) with the reference's rounding (hazard H1): true fp32 division on the CPU oracle,
// multiply-by-reciprocal on the reference's CUDA-eager path.
__device__ __forceinline__ int cell_of(float x, float res, float inv_res, bool recip) {
    float q = recip ? __fmul_rn(x, inv_res) : __fdiv_rn(x, res);
    return (int)floorf(q);
}

// slot = (cell . primes) mod B with the reference's signed fmod + negative-index wrap (hazard H2),
// i.e. the mathematical (floored) modulus.  model/neural_points.py:216, 472
__device__ __forceinline__ uint32_t slot_of_cell(int gx, int gy, int gz, uint32_t B) {
    long long s = (long long)gx * kP1 + (long long)gy * kP2 + (long long)gz * kP3;
    long long r = s % (long long)B;
    if (r < 0) r += (long long)B;
    return (uint32_t)r;
}

// ||a - b||^2 with the CPU reference's association and no FMA contraction (hazard H3):
// (dx*dx + dy*dy) + dz*dz, each op rounded to nearest.
__device__ __forceinline__ float dist2_rn(float ax, float ay, float az, float bx, float by, float bz) {
    float dx = __fsub_rn(ax, bx), dy = __fsub_rn(ay, by), dz = __fsub_rn(az, bz);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// L2 prefetch of a contiguous range (16-byte aligned, size a multiple of 16): one instruction, no register or
// shared-memory cost.  The persistent decode / backward CTAs use it for the stage blocks and relu' masks of the block
// AFTER the next one, so that those streaming reads hit L2 instead of DRAM (decode 0.441 -> 0.422 ms, backward
// 0.829 -> 0.768 ms per 4 Mi queries, profiles/r3b_variants.json).
__device__ __forceinline__ void l2_prefetch(const void* p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(p), "r"(bytes) : "memory");
}

// order-preserving float <-> uint32 encoding for atomicMin/Max on floats
__device__ __forceinline__ uint32_t f2ord(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

}  // namespace pinq

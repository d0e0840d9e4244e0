// query_bwd.cu -- hand-derived backward of the fused query w.r.t. neural-point features and decoder
// parameters.  Replaces torch.autograd through
// model/neural_points.py:562-632 + model/decoder.py:66-88 as triggered at utils/mapper.py:572.
//
//   a_k        = g * w_k                    (g = dL/dsdf; the IDW weights carry no parameter gradient)
//   delta_kj   = scale * W2[j] * 1[pre_kj > 0]
//   dL/df_k    = a_k * sum_j delta_kj W1[j][0:8]        -> red.global.add.v4.f32 into grad_features[idx_k]
//   dL/dW1[j]  += a_k delta_kj in_k ; dL/db1[j] += a_k delta_kj
//   dL/dW2[j]  += a_k scale relu(pre_kj) ; dL/db2 += a_k scale
// The hidden layer is recomputed from the saved (knn_idx, knn_w): 48 B/query of saved state instead of
// 1.5 KB/query of activations.
#include "query_common.cuh"

namespace pinq {

struct BwdParams {
    pinq_decoder_view dec;
    const float* xyz;
    int64_t n;
    bool rotate;
    const float* point_xyz;
    const float4* quats;
    const float* features;
    const int32_t* knn_idx;
    const float* knn_w;
    const float* grad_sdf;
    float* grad_features;
    float *grad_W1, *grad_b1, *grad_W2, *grad_b2;
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// per-CTA decoder-gradient accumulator layout: [j*13 + 0..10] dW1[j][:], [j*13+11] db1[j], [j*13+12] dW2[j], [H*13] db2
constexpr int kAccN = H * 13 + 1;

template <bool DEC>
__global__ void __launch_bounds__(kBlock) k_query_bwd(const BwdParams p) {
    __shared__ __align__(16) float4 s_w[H * 4];
    __shared__ float s_acc[DEC ? kAccN : 1];
    load_decoder(p.dec, s_w);
    if (DEC)
        for (int t = threadIdx.x; t < kAccN; t += blockDim.x) s_acc[t] = 0.f;
    __syncthreads();

    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = q < p.n;
    const int lane = threadIdx.x & 31;

    float in[K][IN];
    float a[K];
    int32_t kidx[K];
    const float scale = p.dec.sdf_scale;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        kidx[k] = -1;
        a[k] = 0.f;
#pragma unroll
        for (int i = 0; i < IN; ++i) in[k][i] = 0.f;
    }
    if (live) {
        const float g = p.grad_sdf[q];
        const float x = p.xyz[3 * q], y = p.xyz[3 * q + 1], z = p.xyz[3 * q + 2];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int32_t li = p.knn_idx[q * K + k];
            kidx[k] = li;
            a[k] = g * p.knn_w[q * K + k];
            if (li >= 0) {
                const float4* fr = reinterpret_cast<const float4*>(p.features + (int64_t)li * F);
                const float4 fa = __ldg(fr), fb = __ldg(fr + 1);
                in[k][0] = fa.x; in[k][1] = fa.y; in[k][2] = fa.z; in[k][3] = fa.w;
                in[k][4] = fb.x; in[k][5] = fb.y; in[k][6] = fb.z; in[k][7] = fb.w;
                const float* pp = p.point_xyz + 3 * (int64_t)li;
                float vx = __fsub_rn(x, pp[0]), vy = __fsub_rn(y, pp[1]), vz = __fsub_rn(z, pp[2]);
                if (p.rotate) quat_rotate(__ldg(&p.quats[li]), false, vx, vy, vz);
                in[k][8] = vx; in[k][9] = vy; in[k][10] = vz;
            }
        }
    }

    float gf[K][F];
#pragma unroll
    for (int k = 0; k < K; ++k)
#pragma unroll
        for (int i = 0; i < F; ++i) gf[k][i] = 0.f;
    float db2 = 0.f;

    for (int j = 0; j < H; ++j) {
        const float4 wa = s_w[4 * j], wb = s_w[4 * j + 1], wc = s_w[4 * j + 2], wd = s_w[4 * j + 3];
        float dw[13];
#pragma unroll
        for (int i = 0; i < 13; ++i) dw[i] = 0.f;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const float pre = mlp_pre(in[k], wa, wb, wc);
            const float t = pre > 0.f ? a[k] * scale * wd.x : 0.f;  // a_k * delta_kj
            gf[k][0] = fmaf(t, wa.x, gf[k][0]); gf[k][1] = fmaf(t, wa.y, gf[k][1]);
            gf[k][2] = fmaf(t, wa.z, gf[k][2]); gf[k][3] = fmaf(t, wa.w, gf[k][3]);
            gf[k][4] = fmaf(t, wb.x, gf[k][4]); gf[k][5] = fmaf(t, wb.y, gf[k][5]);
            gf[k][6] = fmaf(t, wb.z, gf[k][6]); gf[k][7] = fmaf(t, wb.w, gf[k][7]);
            if (DEC) {
#pragma unroll
                for (int i = 0; i < IN; ++i) dw[i] = fmaf(t, in[k][i], dw[i]);
                dw[11] += t;
                dw[12] = fmaf(a[k] * scale, fmaxf(pre, 0.f), dw[12]);
            }
        }
        if (DEC) {
#pragma unroll
            for (int i = 0; i < 13; ++i) {
                const float v = warp_sum(dw[i]);
                if (lane == 0) atomicAdd(&s_acc[j * 13 + i], v);
            }
        }
    }
    if (DEC) {
#pragma unroll
        for (int k = 0; k < K; ++k) db2 += a[k] * scale;
        db2 = warp_sum(db2);
        if (lane == 0) atomicAdd(&s_acc[H * 13], db2);
    }

    if (live && p.grad_features) {
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (kidx[k] >= 0) {
                float* dst = p.grad_features + (int64_t)kidx[k] * F;
                red_add_v4(dst, gf[k][0], gf[k][1], gf[k][2], gf[k][3]);
                red_add_v4(dst + 4, gf[k][4], gf[k][5], gf[k][6], gf[k][7]);
            }
        }
    }

    if (DEC) {
        __syncthreads();
        for (int t = threadIdx.x; t < kAccN; t += blockDim.x) {
            const float v = s_acc[t];
            if (t == H * 13) {
                atomicAdd(p.grad_b2, v);
            } else {
                const int j = t / 13, i = t - j * 13;
                if (i < IN) atomicAdd(&p.grad_W1[j * IN + i], v);
                else if (i == 11) atomicAdd(&p.grad_b1[j], v);
                else atomicAdd(&p.grad_W2[j], v);
            }
        }
    }
}

}  // namespace pinq

using namespace pinq;

extern "C" int pinq_query_fused_bwd(const pinq_bwd_args* a, void* stream) {
    PINQ_REQUIRE(a != nullptr, "null args");
    int rc = check_decoder(a->dec, a->k, a->feature_dim);
    if (rc) return rc;
    PINQ_REQUIRE(a->xyz && a->point_xyz && a->features && a->knn_idx && a->knn_w && a->grad_sdf, "null input");
    PINQ_REQUIRE(!(a->flags & PINQ_F_ROTATE) || a->quats, "PINQ_F_ROTATE needs quats");
    const bool dec = a->grad_W1 != nullptr;
    PINQ_REQUIRE(!dec || (a->grad_b1 && a->grad_W2 && a->grad_b2), "decoder gradients: all four or none");
    if (a->n == 0 || (!dec && !a->grad_features)) return PINQ_OK;
    BwdParams p;
    p.dec = a->dec; p.xyz = a->xyz; p.n = a->n; p.rotate = (a->flags & PINQ_F_ROTATE) != 0;
    p.point_xyz = a->point_xyz; p.quats = reinterpret_cast<const float4*>(a->quats);
    p.features = a->features; p.knn_idx = a->knn_idx; p.knn_w = a->knn_w; p.grad_sdf = a->grad_sdf;
    p.grad_features = a->grad_features;
    p.grad_W1 = a->grad_W1; p.grad_b1 = a->grad_b1; p.grad_W2 = a->grad_W2; p.grad_b2 = a->grad_b2;
    cudaStream_t st = as_stream(stream);
    if (dec) k_query_bwd<true><<<grid_for(a->n, kBlock), kBlock, 0, st>>>(p);
    else k_query_bwd<false><<<grid_for(a->n, kBlock), kBlock, 0, st>>>(p);
    return check_launch("pinq_query_fused_bwd");
}

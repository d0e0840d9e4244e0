// gn.cu -- tracker-side kernels: point transform, fused validity + Geman-McClure weights +
// Gauss-Newton normal-equation reduction, and the fused Adam step of the mapper.
// Reference: utils/tools.py:386-399, utils/tracker.py:293-394, 468-491, utils/tools.py:89-116.
#include "common.cuh"

namespace pinq {

__global__ void k_transform(const float* __restrict__ xyz, int64_t n, const float* __restrict__ T12,
                            float* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
#pragma unroll
    for (int r = 0; r < 3; ++r)
        out[3 * i + r] = fmaf(x, T12[4 * r], fmaf(y, T12[4 * r + 1], fmaf(z, T12[4 * r + 2], T12[4 * r + 3])));
}

struct Pose12 { float v[12]; };

// same arithmetic with the pose passed by value: no host->device copy in front of every Gauss-Newton iteration
__global__ void k_transform_pose(const float* __restrict__ xyz, int64_t n, Pose12 T, float* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
#pragma unroll
    for (int r = 0; r < 3; ++r)
        out[3 * i + r] = fmaf(x, T.v[4 * r], fmaf(y, T.v[4 * r + 1], fmaf(z, T.v[4 * r + 2], T.v[4 * r + 3])));
}

constexpr int kGnVals = 30;  // 21 (upper H) + 6 (b) + sum w + sum |r| + count
constexpr int kGnBlock = 256;

// One pass: validity mask, GM weights, J = [p x g, g], fp64 accumulation, warp-shuffle + block reduce,
// one atomicAdd(double) per value per CTA.
__global__ void __launch_bounds__(kGnBlock) k_gn_accumulate(
    const float* __restrict__ xyz, const float* __restrict__ sdf, const float* __restrict__ grad,
    const float* __restrict__ sdf_std, const int32_t* __restrict__ nn_count, const float* __restrict__ label, int64_t n,
    int min_nn, float min_g, float max_g, float max_std, float gm_d, float gm_g, double* out, uint8_t* valid_out) {
    double acc[kGnVals];
#pragma unroll
    for (int i = 0; i < kGnVals; ++i) acc[i] = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float gx = grad[3 * i], gy = grad[3 * i + 1], gz = grad[3 * i + 2];
        const float gn = sqrtf(gx * gx + gy * gy + gz * gz);  // tracker.py:293
        const bool valid = (nn_count[i] >= min_nn) && (gn < max_g) && (gn > min_g) && (sdf_std[i] < max_std);  // :305
        if (valid_out) valid_out[i] = (uint8_t)valid;
        if (!valid) continue;
        const float r = sdf[i] - (label ? label[i] : 0.f);      // :338
        const float an = gn - 1.f;                               // :334
        float wg = gm_g > 0.f ? gm_g / (gm_g * gm_g + an * an) : 1.f;  // :353
        float wr = gm_d > 0.f ? gm_d / (gm_d * gm_d + r * r) : 1.f;    // :354
        wg = gm_g > 0.f ? wg * wg : 1.f;
        wr = gm_d > 0.f ? wr * wr : 1.f;
        const float w = wr * wg;                                 // :392
        const float px = xyz[3 * i], py = xyz[3 * i + 1], pz = xyz[3 * i + 2];
        float J[6];
        J[0] = py * gz - pz * gy;  // torch.cross(points, grad), tracker.py:470
        J[1] = pz * gx - px * gz;
        J[2] = px * gy - py * gx;
        J[3] = gx; J[4] = gy; J[5] = gz;
        int t = 0;
#pragma unroll
        for (int a = 0; a < 6; ++a) {
            const double wa = (double)w * (double)J[a];
#pragma unroll
            for (int b = a; b < 6; ++b) acc[t++] += wa * (double)J[b];
        }
#pragma unroll
        for (int a = 0; a < 6; ++a) acc[21 + a] += (double)w * (double)r * (double)J[a];
        acc[27] += (double)w;
        acc[28] += (double)fabsf(r);
        acc[29] += 1.0;
    }
    __shared__ double s_red[kGnBlock / 32][kGnVals];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < kGnVals; ++i) {
        const double v = warp_sum(acc[i]);
        if (lane == 0) s_red[wid][i] = v;
    }
    __syncthreads();
    if (threadIdx.x < kGnVals) {
        double v = 0.0;
#pragma unroll
        for (int w = 0; w < kGnBlock / 32; ++w) v += s_red[w][threadIdx.x];
        if (v != 0.0) atomicAdd(&out[threadIdx.x], v);
    }
}

__global__ void k_adam(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                       int64_t n, float beta1, float beta2, float eps, float step_size, float bc2_sqrt) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float gi = g[i];
    float mi = m[i], vi = v[i];
    mi = mi + (gi - mi) * (1.f - beta1);              // exp_avg.lerp_(grad, 1 - beta1)
    vi = vi * beta2 + (1.f - beta2) * gi * gi;        // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
    const float denom = sqrtf(vi) / bc2_sqrt + eps;   // (exp_avg_sq.sqrt() / bias_correction2_sqrt).add_(eps)
    p[i] = p[i] - step_size * (mi / denom);           // param.addcdiv_(exp_avg, denom, value=-step_size)
    m[i] = mi;
    v[i] = vi;
}

}  // namespace pinq

using namespace pinq;

extern "C" {

int pinq_transform_points(const float* xyz, int64_t n, const float* T12, float* out, void* stream) {
    PINQ_REQUIRE(xyz && T12 && out, "null pointer");
    if (n == 0) return PINQ_OK;
    k_transform<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(xyz, n, T12, out);
    return check_launch("pinq_transform_points");
}

int pinq_gn_accumulate(const float* xyz, const float* sdf, const float* grad, const float* sdf_std,
                       const int32_t* nn_count, const float* sdf_label, int64_t n, int32_t min_nn, float min_grad_norm,
                       float max_grad_norm, float max_sdf_std, float gm_dist, float gm_grad, double* out32,
                       uint8_t* valid_out, void* stream) {
    PINQ_REQUIRE(xyz && sdf && grad && sdf_std && nn_count && out32, "null pointer");
    if (n == 0) return PINQ_OK;
    unsigned g = grid_for(n, kGnBlock);
    if (g > 148u * 2u) g = 148u * 2u;
    k_gn_accumulate<<<g, kGnBlock, 0, as_stream(stream)>>>(xyz, sdf, grad, sdf_std, nn_count, sdf_label, n, min_nn,
                                                           min_grad_norm, max_grad_norm, max_sdf_std, gm_dist, gm_grad,
                                                           out32, valid_out);
    return check_launch("pinq_gn_accumulate");
}

int pinq_tracking_step(const pinq_tracking_args* a, const float* h_T12, void* stream) {
    PINQ_REQUIRE(a && a->gn_out && a->h_gn_out, "null pointer");
    const pinq_query_args& q = a->query;
    PINQ_REQUIRE(q.xyz && q.sdf && q.grad_x && q.sdf_std && q.nn_count, "query outputs sdf / grad_x / sdf_std / nn_count needed");
    PINQ_REQUIRE(!h_T12 || a->source_xyz, "source points needed with a pose");
    cudaStream_t st = as_stream(stream);
    const int64_t n = q.n;
    int rc;
    if (n > 0 && h_T12) {
        Pose12 T;
        for (int i = 0; i < 12; ++i) T.v[i] = h_T12[i];
        k_transform_pose<<<grid_for(n, 256), 256, 0, st>>>(a->source_xyz, n, T, const_cast<float*>(q.xyz));
        if ((rc = check_launch("k_transform_pose")) != PINQ_OK) return rc;
    }
    if ((rc = pinq_query_fused_fwd(&a->query, stream)) != PINQ_OK) return rc;
    if (cudaMemsetAsync(a->gn_out, 0, 32 * sizeof(double), st) != cudaSuccess) return check_launch("pinq_tracking_step memset");
    rc = pinq_gn_accumulate(q.xyz, q.sdf, q.grad_x, q.sdf_std, q.nn_count, a->sdf_label, n, a->min_nn, a->min_grad_norm,
                            a->max_grad_norm, a->max_sdf_std, a->gm_dist, a->gm_grad, a->gn_out, nullptr, stream);
    if (rc != PINQ_OK) return rc;
    if (cudaMemcpyAsync(a->h_gn_out, a->gn_out, 32 * sizeof(double), cudaMemcpyDeviceToHost, st) != cudaSuccess)
        return check_launch("pinq_tracking_step copy");
    return PINQ_OK;
}

int pinq_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1,
                   float beta2, float eps, int32_t step, void* stream) {
    PINQ_REQUIRE(param && grad && exp_avg && exp_avg_sq && step >= 1, "bad args");
    if (n == 0) return PINQ_OK;
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const double bc2 = 1.0 - pow((double)beta2, (double)step);
    k_adam<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(param, grad, exp_avg, exp_avg_sq, n, beta1, beta2, eps,
                                                            (float)((double)lr / bc1), (float)sqrt(bc2));
    return check_launch("pinq_adam_step");
}

}  // extern "C"

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

// ---- device-side Gauss-Newton loop (Tracker.tracking, utils/tracker.py:91-141) -------------------------------------
// state = 32 doubles: [0..15] pose T (row-major 4x4), [16] last_sdf_residual_cm, [17] sdf_residual_cm, [18..23] spare,
// then 16 int32 in the place of doubles 24..31: [0] done, [1] valid_flag, [2] converged, [3] iterations run,
// [4] valid_point_count, [5] failure reason (1 = residual grew, 2 = too few valid points).
constexpr int kTsLast = 16, kTsRes = 17, kTsInts = 24;
enum { kTiDone = 0, kTiValid = 1, kTiConv = 2, kTiIters = 3, kTiCount = 4, kTiReason = 5 };

__device__ __forceinline__ int32_t* track_ints(double* state) { return reinterpret_cast<int32_t*>(state + kTsInts); }

// cur = T * source with the pose cast to fp32 as the reference does (tracker.py:94, tools.py:386-399)
__global__ void k_track_transform(const float* __restrict__ xyz, int64_t n, const double* __restrict__ state,
                                  float* __restrict__ out) {
    if (reinterpret_cast<const int32_t*>(state + kTsInts)[kTiDone]) return;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
#pragma unroll
    for (int r = 0; r < 3; ++r)
        out[3 * i + r] = fmaf(x, (float)state[4 * r], fmaf(y, (float)state[4 * r + 1], fmaf(z, (float)state[4 * r + 2], (float)state[4 * r + 3])));
}

struct TrackLoop {
    int iter, iter_n, min_valid_points;
    double lm_lambda, term_deg, term_m, min_valid_ratio, max_increment_ratio, source_count;
};

// implicit_reg (tracker.py:468-520) + the loop body's bookkeeping (:108-141), one thread, fp64
__global__ void k_track_solve(const double* __restrict__ g, double* __restrict__ state, TrackLoop L) {
    if (threadIdx.x != 0) return;
    int32_t* fl = track_ints(state);
    if (fl[kTiDone]) return;
    fl[kTiIters] = L.iter + 1;
    const int count = (int)llrint(g[29]);
    fl[kTiCount] = count;
    double dT[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    double residual = 0.0;
    if (count >= 10) {  // tracker.py:310-312 otherwise: identity increment, residual 0
        residual = g[28] / count * 100.0;
        const double norm = 2.0 * g[27] / count;  // w /= 2 mean(w), tracker.py:394
        double H[6][7];
        int k = 0;
        for (int r = 0; r < 6; ++r)
            for (int c = r; c < 6; ++c, ++k) H[r][c] = H[c][r] = g[k] / norm;
        for (int r = 0; r < 6; ++r) {
            H[r][6] = -g[21 + r] / norm;
            H[r][r] += L.lm_lambda * H[r][r];   // H + lambda diag(H), tracker.py:489
        }
        for (int c = 0; c < 6; ++c) {           // t = H^-1 b by elimination with partial pivoting
            int piv = c;
            for (int r = c + 1; r < 6; ++r)
                if (fabs(H[r][c]) > fabs(H[piv][c])) piv = r;
            if (piv != c)
                for (int j = 0; j < 7; ++j) { const double tmp = H[c][j]; H[c][j] = H[piv][j]; H[piv][j] = tmp; }
            for (int r = c + 1; r < 6; ++r) {
                const double f = H[r][c] / H[c][c];
                for (int j = c; j < 7; ++j) H[r][j] -= f * H[c][j];
            }
        }
        double t[6];
        for (int r = 5; r >= 0; --r) {
            double v = H[r][6];
            for (int j = r + 1; j < 6; ++j) v -= H[r][j] * t[j];
            t[r] = v / H[r][r];
        }
        // expmap (tracker.py:580-589): a zero rotation yields NaNs there as here
        const double angle = sqrt(t[0] * t[0] + t[1] * t[1] + t[2] * t[2]);
        const double ax = t[0] / angle, ay = t[1] / angle, az = t[2] / angle;
        const double S[9] = {0, -az, ay, az, 0, -ax, -ay, ax, 0};
        const double sn = sin(angle), cs = 1.0 - cos(angle);
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) {
                double ss = 0.0;
                for (int j = 0; j < 3; ++j) ss += S[3 * r + j] * S[3 * j + c];
                dT[4 * r + c] = (r == c ? 1.0 : 0.0) + S[3 * r + c] * sn + ss * cs;
            }
        dT[3] = t[3]; dT[7] = t[4]; dT[11] = t[5];
    }
    double Tn[16];
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) {
            double v = 0.0;
            for (int j = 0; j < 4; ++j) v += dT[4 * r + j] * state[4 * j + c];
            Tn[4 * r + c] = v;
        }
    for (int i = 0; i < 16; ++i) state[i] = Tn[i];
    state[kTsRes] = residual;
    const double last = state[kTsLast];
    if ((residual - last) / last > L.max_increment_ratio) { fl[kTiValid] = 0; fl[kTiReason] = 1; }
    else state[kTsLast] = residual;
    if (count < L.min_valid_points || (double)count / L.source_count < L.min_valid_ratio) { fl[kTiValid] = 0; fl[kTiReason] = 2; }
    if (!fl[kTiValid] || fl[kTiConv] || L.iter == L.iter_n - 1) { fl[kTiDone] = 1; return; }
    const double rot_deg = acos((dT[0] + dT[5] + dT[10] - 1.0) / 2.0) * (180.0 / 3.14159265358979323846);
    const double tran = sqrt(dT[3] * dT[3] + dT[7] * dT[7] + dT[11] * dT[11]);
    // `and` binds tighter than `or` (tracker.py:135); a NaN angle compares false
    if ((fabs(rot_deg) < L.term_deg && tran < L.term_m) || L.iter == L.iter_n - 2) fl[kTiConv] = 1;
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

int pinq_tracking_run(const pinq_tracking_args* a, const pinq_tracking_loop* L, void* stream) {
    PINQ_REQUIRE(a && L && a->gn_out && L->state && L->h_state, "null pointer");
    const pinq_query_args& q0 = a->query;
    PINQ_REQUIRE(q0.xyz && q0.sdf && q0.grad_x && q0.sdf_std && q0.nn_count && a->source_xyz, "query buffers / source points needed");
    PINQ_REQUIRE(L->iters >= 0 && L->first_iter >= 0 && L->first_iter + L->iters <= L->iter_n, "chunk outside [0, iter_n)");
    cudaStream_t st = as_stream(stream);
    const int64_t n = q0.n;
    if (L->h_state_in &&
        cudaMemcpyAsync(L->state, L->h_state_in, PINQ_TRACK_STATE_DOUBLES * sizeof(double), cudaMemcpyHostToDevice, st) != cudaSuccess)
        return check_launch("pinq_tracking_run upload");
    pinq_query_args q = q0;
    q.skip = reinterpret_cast<const int32_t*>(L->state + kTsInts) + kTiDone;
    TrackLoop tl;
    tl.iter_n = L->iter_n; tl.min_valid_points = L->min_valid_points;
    tl.lm_lambda = L->lm_lambda; tl.term_deg = L->term_thre_deg; tl.term_m = L->term_thre_m;
    tl.min_valid_ratio = L->min_valid_ratio; tl.max_increment_ratio = L->max_increment_ratio; tl.source_count = (double)n;
    int rc;
    for (int j = 0; j < L->iters; ++j) {
        if (n > 0) {
            k_track_transform<<<grid_for(n, 256), 256, 0, st>>>(a->source_xyz, n, L->state, const_cast<float*>(q.xyz));
            if ((rc = check_launch("k_track_transform")) != PINQ_OK) return rc;
        }
        if ((rc = pinq_query_fused_fwd(&q, stream)) != PINQ_OK) return rc;
        if (cudaMemsetAsync(a->gn_out, 0, 32 * sizeof(double), st) != cudaSuccess) return check_launch("pinq_tracking_run memset");
        rc = pinq_gn_accumulate(q.xyz, q.sdf, q.grad_x, q.sdf_std, q.nn_count, a->sdf_label, n, a->min_nn, a->min_grad_norm,
                                a->max_grad_norm, a->max_sdf_std, a->gm_dist, a->gm_grad, a->gn_out, nullptr, stream);
        if (rc != PINQ_OK) return rc;
        tl.iter = L->first_iter + j;
        k_track_solve<<<1, 32, 0, st>>>(a->gn_out, L->state, tl);
        if ((rc = check_launch("k_track_solve")) != PINQ_OK) return rc;
    }
    if (cudaMemcpyAsync(L->h_state, L->state, PINQ_TRACK_STATE_DOUBLES * sizeof(double), cudaMemcpyDeviceToHost, st) != cudaSuccess)
        return check_launch("pinq_tracking_run copy");
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

// frame.cu -- per-frame preprocessing kernels of the data path in front of the hot path:
// motion undistortion ("deskewing") of a sweep.  Reference: utils/tools.py:540-567 (deskewing),
// dataset/slam_dataset.py:233-257 (per-point times), :367-370 and :419-422 (call sites: source points before tracking
// with the previous motion, the full cloud after tracking with the estimated motion).
#include "common.cuh"

namespace pinq {

// per-point time of a scan without time stamps, Velodyne convention (clockwise from the -x axis): 0.5 (-atan2(y, x) / pi + 1)
__global__ void k_velodyne_ts(const float* __restrict__ xyz, int64_t n, int stride, float* __restrict__ ts) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float yaw = -atan2f(xyz[stride * i + 1], xyz[stride * i]);
    ts[i] = 0.5f * (__fdiv_rn(yaw, 3.14159265358979323846f) + 1.0f);
}

// p <- slerp(I, R, tau) p + tau t,  tau = (ts - min) / (max - min) - ts_mid,  [R | t] = T_last<-cur.
// slerp(I, R, tau) = exp(tau log R) (roma.rotmat_slerp: shortest arc, extrapolating for tau < 0): Rodrigues about the
// axis of R by the angle tau * theta.  Axis and angle are derived once per CTA from the 4x4 pose in device memory.
__global__ void k_deskew(float* __restrict__ xyz, int stride, const float* __restrict__ ts, int64_t n,
                         const float* __restrict__ ts_minmax, const float* __restrict__ pose, float ts_mid) {
    __shared__ float s_ax[3], s_theta, s_t[3], s_lo, s_span;
    if (threadIdx.x == 0) {
        const float r00 = pose[0], r01 = pose[1], r02 = pose[2], r10 = pose[4], r11 = pose[5], r12 = pose[6], r20 = pose[8],
                    r21 = pose[9], r22 = pose[10];
        float c = 0.5f * ((r00 + r11 + r22) - 1.0f);
        c = fminf(1.0f, fmaxf(-1.0f, c));
        const float wx = r21 - r12, wy = r02 - r20, wz = r10 - r01;
        const float nw = sqrtf(wx * wx + wy * wy + wz * wz);
        // angle from atan2(|w| / 2, cos): well conditioned for the small rotations between consecutive frames
        const float theta = atan2f(0.5f * nw, c);
        const float inv = nw > 1e-20f ? 1.0f / nw : 0.f;
        s_ax[0] = wx * inv; s_ax[1] = wy * inv; s_ax[2] = wz * inv;
        s_theta = nw > 1e-20f ? theta : 0.f;
        s_t[0] = pose[3]; s_t[1] = pose[7]; s_t[2] = pose[11];
        s_lo = ts_minmax[0];
        s_span = ts_minmax[1] - ts_minmax[0];
    }
    __syncthreads();
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float tau = __fdiv_rn(ts[i] - s_lo, s_span) - ts_mid;
    float sn, cs;
    sincosf(tau * s_theta, &sn, &cs);
    const float x = xyz[stride * i], y = xyz[stride * i + 1], z = xyz[stride * i + 2];
    const float ax = s_ax[0], ay = s_ax[1], az = s_ax[2];
    const float cxp = ay * z - az * y, cyp = az * x - ax * z, czp = ax * y - ay * x;  // a x p
    const float k = (1.0f - cs) * (ax * x + ay * y + az * z);
    xyz[stride * i] = x * cs + cxp * sn + ax * k + tau * s_t[0];
    xyz[stride * i + 1] = y * cs + cyp * sn + ay * k + tau * s_t[1];
    xyz[stride * i + 2] = z * cs + czp * sn + az * k + tau * s_t[2];
}

}  // namespace pinq

using namespace pinq;

extern "C" {

int pinq_velodyne_point_ts(const float* xyz, int64_t n, int32_t stride, float* ts_out, void* stream) {
    PINQ_REQUIRE(stride >= 3 && n >= 0, "stride < 3 or negative size");
    if (n == 0) return PINQ_OK;
    PINQ_REQUIRE(xyz && ts_out, "null pointer");
    k_velodyne_ts<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(xyz, n, stride, ts_out);
    return check_launch("pinq_velodyne_point_ts");
}

int pinq_deskew(float* xyz, int64_t n, int32_t stride, const float* ts, const float* d_ts_minmax2, const float* d_pose16,
                float ts_mid_pose, void* stream) {
    PINQ_REQUIRE(stride >= 3 && n >= 0, "stride < 3 or negative size");
    if (n == 0) return PINQ_OK;
    PINQ_REQUIRE(xyz && ts && d_ts_minmax2 && d_pose16, "null pointer");
    k_deskew<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(xyz, stride, ts, n, d_ts_minmax2, d_pose16, ts_mid_pose);
    return check_launch("pinq_deskew");
}

}  // extern "C"

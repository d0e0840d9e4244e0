// pool.cu -- upkeep of the mapper's replay pool (reference: utils/mapper.py:209-263).
//
// Per frame the reference appends the new samples to five pool tensors with torch.cat, computes the window mask
// ||global_coord - frame_origin||^2 < window_radius^2 over the whole pool, optionally knocks out random survivors when
// the pool exceeds its capacity, and boolean-indexes all five tensors (each index = nonzero + gather + host sync).
// Here: one pass writes the flags of [old pool | new samples], the order-preserving compaction of table.cu turns them
// into the survivor list, and one pass moves every surviving row from the old pool or from the new samples straight into
// the other half of a preallocated ping-pong store -- 12 B read per sample for the mask, 40 B read + 40 B written per survivor.
#include "common.cuh"

namespace pinq {

// flags[i] = 1 if row i of [old | new] lies inside the window (mapper.py:231-233: torch.sum(rel ** 2, -1) < r^2,
// fp32, no contraction: (dx*dx + dy*dy) + dz*dz)
__global__ void k_pool_flags(const float* __restrict__ old_g, int64_t n_old, const float* __restrict__ new_g, int64_t n_new,
                             const float* __restrict__ origin, float r2, uint8_t* __restrict__ flags) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_old + n_new) return;
    const float* p = i < n_old ? old_g + 3 * i : new_g + 3 * (i - n_old);
    const float dx = __fsub_rn(p[0], origin[0]), dy = __fsub_rn(p[1], origin[1]), dz = __fsub_rn(p[2], origin[2]);
    const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
    flags[i] = d2 < r2 ? 1 : 0;
}

// filter_mask[true_indices[discarded]] = False (mapper.py:242-246); duplicates in `discarded` are harmless
__global__ void k_pool_discard(uint8_t* __restrict__ flags, const int32_t* __restrict__ kept_index,
                               const int64_t* __restrict__ discarded, int64_t n_discard) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_discard) flags[kept_index[discarded[i]]] = 0;
}

struct PoolMove {
    pinq_pool_view old_, new_, out;
    int64_t n_old, n_new, new_ts;
    const int32_t* kept_index;
    const int32_t* d_count;
    int32_t* d_kept_new;
};

// survivor j <- row kept_index[j] of [old | new]; the new samples carry the frame id as their timestamp (mapper.py:124)
__global__ void k_pool_move(PoolMove m) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int count = *m.d_count;
    bool from_new = false;
    if (j < count) {
        int64_t i = m.kept_index[j];
        from_new = i >= m.n_old;
        const pinq_pool_view& s = from_new ? m.new_ : m.old_;
        if (from_new) i -= m.n_old;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            m.out.coord[3 * j + c] = s.coord[3 * i + c];
            m.out.global_coord[3 * j + c] = s.global_coord[3 * i + c];
        }
        m.out.sdf_label[j] = s.sdf_label[i];
        m.out.weight[j] = s.weight[i];
        m.out.ts[j] = from_new ? m.new_ts : s.ts[i];
    }
    const unsigned b = __ballot_sync(0xffffffffu, from_new);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(m.d_kept_new, __popc(b));
}

}  // namespace pinq

using namespace pinq;

extern "C" {

int pinq_pool_flags(const float* pool_global_xyz, int64_t n_pool, const float* new_global_xyz, int64_t n_new,
                    const float* d_origin3, float radius2, uint8_t* flags, void* stream) {
    PINQ_REQUIRE(n_pool >= 0 && n_new >= 0 && (n_pool == 0 || pool_global_xyz) && (n_new == 0 || new_global_xyz), "null pool");
    PINQ_REQUIRE(d_origin3 && (flags || n_pool + n_new == 0), "null pointer");
    if (n_pool + n_new == 0) return PINQ_OK;
    k_pool_flags<<<grid_for(n_pool + n_new, 256), 256, 0, as_stream(stream)>>>(pool_global_xyz, n_pool, new_global_xyz, n_new,
                                                                              d_origin3, radius2, flags);
    return check_launch("pinq_pool_flags");
}

int pinq_pool_discard(uint8_t* flags, const int32_t* kept_index, const int64_t* discarded, int64_t n_discard, void* stream) {
    PINQ_REQUIRE(flags && kept_index && (discarded || n_discard == 0) && n_discard >= 0, "null pointer");
    if (n_discard == 0) return PINQ_OK;
    k_pool_discard<<<grid_for(n_discard, 256), 256, 0, as_stream(stream)>>>(flags, kept_index, discarded, n_discard);
    return check_launch("pinq_pool_discard");
}

int pinq_pool_move(const pinq_pool_view* old_pool, int64_t n_pool, const pinq_pool_view* new_samples, int64_t n_new,
                   int64_t new_ts, const int32_t* kept_index, const int32_t* d_count, const pinq_pool_view* out,
                   int32_t* d_kept_new, void* stream) {
    PINQ_REQUIRE(old_pool && new_samples && out && kept_index && d_count && d_kept_new, "null pointer");
    PINQ_REQUIRE(n_pool >= 0 && n_new >= 0, "negative size");
    PINQ_REQUIRE(n_pool == 0 || (old_pool->coord && old_pool->global_coord && old_pool->sdf_label && old_pool->weight && old_pool->ts),
                 "null pool array");
    PINQ_REQUIRE(n_new == 0 || (new_samples->coord && new_samples->global_coord && new_samples->sdf_label && new_samples->weight),
                 "null sample array");
    if (n_pool + n_new == 0) return PINQ_OK;
    PINQ_REQUIRE(out->coord && out->global_coord && out->sdf_label && out->weight && out->ts, "null output array");
    PoolMove m;
    m.old_ = *old_pool; m.new_ = *new_samples; m.out = *out;
    m.n_old = n_pool; m.n_new = n_new; m.new_ts = new_ts;
    m.kept_index = kept_index; m.d_count = d_count; m.d_kept_new = d_kept_new;
    k_pool_move<<<grid_for(n_pool + n_new, 256), 256, 0, as_stream(stream)>>>(m);
    return check_launch("pinq_pool_move");
}

}  // extern "C"

// ---- DataSampler.sample (reference: utils/data_sampler.py:20-192) as one kernel ---------------------------------------
namespace pinq {

struct RaySampleParams {
    const float* xyz;       // [n,3] sensor-frame points
    const float* r_surf;    // [n_surf * n] N(0,1)   (element rep * n + ray, the order the reference's repeat() pairs them)
    const float* r_front;   // [n_front * n] U(0,1)
    const float* r_behind;  // [n_behind * n] U(0,1)
    int64_t n;
    int n_surf, n_front, n_behind;
    float range, begin_ratio, end_dist, sigma_ratio, max_range, dist_weight_scale;
    int dist_weight_on, behind_dropoff_on;
    float* out_xyz;         // [n * n_all, 3] ray-wise: sample s of ray i at row i * n_all + s
    float* out_label;
    float* out_weight;
};

__global__ void k_ray_samples(RaySampleParams p) {
    const int n_all = 1 + p.n_surf + p.n_front + p.n_behind;
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= p.n * n_all) return;
    const int64_t i = j / n_all;
    const int s = (int)(j - i * n_all);
    const float x = p.xyz[3 * i], y = p.xyz[3 * i + 1], z = p.xyz[3 * i + 2];
    const float d = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z)));
    float disp = 0.f, ratio = 1.f;
    if (s >= 1 && s <= p.n_surf) {                      // gaussian around the surface (data_sampler.py:46-55)
        disp = __fmul_rn(p.r_surf[(int64_t)(s - 1) * p.n + i], p.range);
        ratio = __fadd_rn(__fdiv_rn(disp, d), 1.f);
    } else if (s > p.n_surf && s <= p.n_surf + p.n_front) {  // free space in front (:76-83)
        const float fmax = __fsub_rn(1.f, __fdiv_rn(__fmul_rn(p.sigma_ratio, p.range), d));
        const float r = p.r_front[(int64_t)(s - 1 - p.n_surf) * p.n + i];
        ratio = __fadd_rn(__fmul_rn(r, __fsub_rn(fmax, p.begin_ratio)), p.begin_ratio);
        disp = __fmul_rn(__fsub_rn(ratio, 1.f), d);
    } else if (s > p.n_surf + p.n_front) {              // free space behind (:90-97)
        const float bmax = __fadd_rn(__fdiv_rn(p.end_dist, d), 1.f);
        const float bmin = __fadd_rn(1.f, __fdiv_rn(__fmul_rn(p.sigma_ratio, p.range), d));
        const float r = p.r_behind[(int64_t)(s - 1 - p.n_surf - p.n_front) * p.n + i];
        ratio = __fadd_rn(__fmul_rn(r, __fsub_rn(bmax, bmin)), bmin);
        disp = __fmul_rn(__fsub_rn(ratio, 1.f), d);
    }
    p.out_xyz[3 * j] = __fmul_rn(x, ratio);
    p.out_xyz[3 * j + 1] = __fmul_rn(y, ratio);
    p.out_xyz[3 * j + 2] = __fmul_rn(z, ratio);
    float w = 1.f;
    if (p.dist_weight_on && s <= p.n_surf)              // :127-128
        w = __fsub_rn(__fadd_rn(1.f, __fmul_rn(p.dist_weight_scale, 0.5f)), __fmul_rn(__fdiv_rn(d, p.max_range), p.dist_weight_scale));
    if (p.behind_dropoff_on) {                          // :131-135
        const float dmin = __fmul_rn(0.2f, p.end_dist), dmax = p.end_dist;
        const float c = fminf(fmaxf(__fdiv_rn(__fsub_rn(dmax, disp), __fsub_rn(dmax, dmin)), 0.f), 1.f);
        w = __fmul_rn(w, __fadd_rn(__fmul_rn(c, 0.8f), 0.2f));
    }
    if (s > p.n_surf) w = -w;                           // the sign flags free-space samples
    p.out_label[j] = -disp;
    p.out_weight[j] = w;
}

}  // namespace pinq

extern "C" int pinq_ray_samples(const float* xyz, int64_t n, const float* r_surf, const float* r_front, const float* r_behind,
                                int32_t n_surf, int32_t n_front, int32_t n_behind, float surface_range, float free_begin_ratio,
                                float free_end_dist, float max_range, float dist_weight_scale, int32_t dist_weight_on,
                                int32_t behind_dropoff_on, float* out_xyz, float* out_label, float* out_weight, void* stream) {
    PINQ_REQUIRE(n >= 0 && n_surf >= 0 && n_front >= 0 && n_behind >= 0, "negative size");
    if (n == 0) return PINQ_OK;
    PINQ_REQUIRE(xyz && out_xyz && out_label && out_weight, "null pointer");
    PINQ_REQUIRE((n_surf == 0 || r_surf) && (n_front == 0 || r_front) && (n_behind == 0 || r_behind), "null random numbers");
    pinq::RaySampleParams p;
    p.xyz = xyz; p.r_surf = r_surf; p.r_front = r_front; p.r_behind = r_behind; p.n = n;
    p.n_surf = n_surf; p.n_front = n_front; p.n_behind = n_behind;
    p.range = surface_range; p.begin_ratio = free_begin_ratio; p.end_dist = free_end_dist; p.sigma_ratio = 2.0f;
    p.max_range = max_range; p.dist_weight_scale = dist_weight_scale;
    p.dist_weight_on = dist_weight_on; p.behind_dropoff_on = behind_dropoff_on;
    p.out_xyz = out_xyz; p.out_label = out_label; p.out_weight = out_weight;
    const int64_t total = n * (1 + n_surf + n_front + n_behind);
    pinq::k_ray_samples<<<pinq::grid_for(total, 256), 256, 0, pinq::as_stream(stream)>>>(p);
    return pinq::check_launch("pinq_ray_samples");
}

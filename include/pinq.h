/*
 * pinq.h -- C-ABI of libpinq: the B200 (sm_100a) neural-point SDF query path of PIN-SLAM.
 *
 * The reference (kelly7707/PIN_SLAM) is pure Python/PyTorch and has no FFI layer; the seam it
 * offers is its Python class API.  Every entry point below names the reference function
 * (path:line relative to the reference root) whose arithmetic it replaces.  The Python mirror
 * in pin_slam_b200/{model,utils} binds these with ctypes.
 *
 * Conventions
 *   - all pointers are DEVICE pointers unless the parameter name starts with h_;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*), never
 *     synchronises the device and never allocates (callers own all buffers);
 *   - return value: 0 = success, negative = PINQ_ERR_*; pinq_last_error() gives the text;
 *   - index type on the device is int32 (the reference uses int64: model/neural_points.py:47);
 *     -1 means "empty / invalid" exactly as in the reference.
 */
#ifndef PINQ_H_
#define PINQ_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PINQ_OK 0
#define PINQ_ERR_BAD_ARG (-1)
#define PINQ_ERR_CUDA (-2)
#define PINQ_ERR_UNSUPPORTED (-3)

#define PINQ_MAX_STENCIL 343 /* (2*3+1)^3: num_nei_cells <= 3 */

/* flags for pinq_query_args.flags */
#define PINQ_F_LOCAL 1u        /* query_locally=True: candidate filter + local index from cand[].w    */
#define PINQ_F_TRAINING 2u     /* training_mode=True: certainty scatter-add (+ ts amax when query_ts)  */
#define PINQ_F_ROTATE 4u       /* after_pgo=True: rotate x-p_k into the point frame (tools.py:316)     */
#define PINQ_F_RECIP_DIV 8u    /* floor(x*(1/res)) instead of floor(x/res): CUDA-eager oracle (H1)     */
#define PINQ_F_SKIP_EMPTY 16u  /* mesher: sdf = 0 where nn_count == 0 (utils/mesher.py:96-103)         */
#define PINQ_F_EXACT_KEYS 64u  /* brick search: every query through the 64-bit-key pass (the fast pass falls back to it on truncated-key ties; tests) */
#define PINQ_F_STENCIL_IN_RANGE 128u /* brick search: the caller has verified that max over the stencil of sum_i ((|dx_i| + 1) res)^2 < max_valid_dist2, i.e. no candidate of a stencil cell can fail the distance test (model/neural_points.py:499): the test is skipped */
#define PINQ_F_MMA_SYNC 32u    /* decode / backward: the mma.sync kernels instead of the tcgen05 ones (A/B tests; also pinq_bwd_args.flags) */

const char* pinq_last_error(void);
int pinq_version(void);

/* Measurement aid (no reference counterpart; the role of the reference's tic()/toc() in utils/tools.py around the
 * query calls): when enabled, the query entry points record CUDA events on the launching stream between their
 * kernels.  pinq_kernel_times() waits for them and returns, per phase i, the milliseconds between mark i and the
 * next recorded mark (-1 where a phase did not run).  Phases of one fused forward + backward: */
#define PINQ_KT_BIN 0      /* spatial binning (memset + 5 kernels)          */
#define PINQ_KT_SEARCH 1   /* stencil masks + K-nearest search (+ resolve)  */
#define PINQ_KT_RESOLVE 2  /* k_resolve (slot-probing search only)          */
#define PINQ_KT_DECODE 3   /* k_decode_tc / k_decode_mma                    */
#define PINQ_KT_SCATTER 4  /* second-pass certainty scatter (training)      */
#define PINQ_KT_FWD_END 5
#define PINQ_KT_BWD 6      /* k_bwd_tc / k_bwd_mma                          */
#define PINQ_KT_BWD_END 7
#define PINQ_KT_MARKS 8
void pinq_kernel_timing(int enable);
int pinq_kernel_times(float* ms_out, int cap);

/* ---- voxel-hash index -------------------------------------------------------------------
 * The table is the reference's direct-mapped `buffer_pt_index` (model/neural_points.py:73)
 * held as dense int32[buffer_size]; slot(cell) = (cell . primes) mod buffer_size with the
 * reference's signed-fmod + negative-index wrap (neural_points.py:216, 472). */

/* model/neural_points.py:73, 377 -- table[:] = -1 */
int pinq_table_clear(int32_t* table, uint32_t buffer_size, void* stream);

/* model/neural_points.py:214-216 -- slots[i] = slot(floor(xyz[i]/res)) */
int pinq_hash_points(const float* xyz, int64_t n, float res, uint32_t buffer_size, uint32_t flags,
                     int32_t* slots, void* stream);

/* model/neural_points.py:218, 242 -- out[i] = table[slots[i]] */
int pinq_table_gather(const int32_t* table, const int32_t* slots, int64_t n, int32_t* out, void* stream);

/* model/neural_points.py:247, 400, 422 -- table[slots[i]] = values[i] with the CPU reference's
 * sequential semantics: among duplicate slots the highest i wins.  `scratch` = int32[n]. */
int pinq_table_scatter_last_wins(int32_t* table, const int32_t* slots, const int32_t* values, int64_t n,
                                 void* stream);

/* model/neural_points.py:221-231 -- update_mask[i] = empty | moved-away | stale (uint8 0/1).
 * hash_idx = table[slots]; map arrays may be empty (m == 0 -> all ones). */
int pinq_update_mask(const float* sample_xyz, const int32_t* hash_idx, int64_t n, const float* map_xyz,
                     const int64_t* map_ts_update, int64_t m, const float* travel_dist, int64_t cur_ts,
                     float res, float diff_travel_dist_local, uint8_t* update_mask, void* stream);

/* Order-preserving stream compaction helper: out_index[j] = i for the j-th set flag; *d_count = #set.
 * `scratch` must hold pinq_compact_scratch_bytes(n) bytes.  (replaces boolean-mask indexing,
 * model/neural_points.py:239, 293-296, 302) */
int64_t pinq_compact_scratch_bytes(int64_t n);
int pinq_compact_flags(const uint8_t* flags, int64_t n, int32_t* out_index, int32_t* d_count, void* scratch,
                       void* stream);

/* Allocation-free variant of pinq_compact_flags.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same inputs, outputs and results as pinq_compact_flags, but the scan's tile sums live in a static device
 * array inside the library (64 KB per GPU, held from library load until exit) instead of caller scratch; the call issues
 * no cudaMalloc / cudaFree / allocating library routine.  Limit: n <= PINQ_COMPACT_STATIC_MAX_N flags; larger n returns
 * PINQ_ERR_BAD_ARG with a message naming the limit.  The array is shared by every call on the device and keeps the last
 * call's contents (nothing needs clearing: each call overwrites what it reads): calls on one stream are ordered, calls
 * that overlap on the GPU (several streams or host threads) RACE -- serialise them or use pinq_compact_flags.  The
 * first call on a device resolves the array's address (cudaGetSymbolAddress): make it once outside graph capture. */
#define PINQ_COMPACT_STATIC_MAX_N (16ll * 1024 * 1024)
int pinq_compact_flags_static(const uint8_t* flags, int64_t n, int32_t* out_index, int32_t* d_count, void* stream);

/* utils/tools.py:409-442 (mode 0, closest to voxel centre) and :444-477 (mode 1, min `value`).
 * Three steps so that the caller owns all memory:
 *   1. pinq_vds_prepare : per-point quirky linear key (scalar V, hazard H5) + quantised score;
 *                         needs the two global reductions -> d_minmax float[8] scratch.
 *   2. caller sorts nothing: pinq_vds_select groups by key with an open-addressing scratch table
 *      (uint64 keys[cap], uint64 best[cap], cap = power of two >= 2n) and atomically keeps the
 *      minimal packed (score, index) per key;
 *   3. pinq_vds_emit compacts the winners; the caller orders them by key (ascending) -- the
 *      reference's output order is the sorted unique key order. */
int pinq_vds_prepare(const float* xyz, int64_t n, float voxel, const float* value /*nullable*/, int mode,
                     float* d_scratch8, int64_t* keys, int32_t* qscore, void* stream);
int pinq_vds_select(const int64_t* keys, const int32_t* qscore, int64_t n, uint64_t* tab_keys,
                    uint64_t* tab_best, uint64_t cap, void* stream);
int pinq_vds_emit(const uint64_t* tab_keys, const uint64_t* tab_best, uint64_t cap, int64_t n,
                  int64_t* out_keys, int32_t* out_index, int32_t* d_count, void* stream);

/* model/neural_points.py:272-305 -- local-map window: flags[i] = ||p_i - sensor||^2 < r^2 and
 * |travel_dist[cur_ts] - travel_dist[ts_used[i]]| < diff (uint8). */
int pinq_local_mask(const float* map_xyz, const int64_t* ts_create, const int64_t* ts_update, int64_t m,
                    const float* sensor_xyz3, const float* travel_dist, int64_t cur_ts, float radius2,
                    float diff_travel_dist_local, int use_mid_ts, uint8_t* flags, void* stream);

/* ---- occupancy index: the L2-resident, bit-identical read view of the slot table ----------
 * occ[w]  = { bits: occupancy of slots 32w..32w+31, prefix: #occupied slots below 32w }  (uint32[nwords][2])
 * rec[r]  = { x, y, z, bits(index) } of the r-th occupied slot; lookup(slot) == table[slot] by construction.
 * Built from the dense table by two streaming passes whenever the map changes (csrc/index.cu). */
int64_t pinq_index_words(uint32_t buffer_size);
int64_t pinq_index_scratch_bytes(uint32_t buffer_size);
int pinq_index_build(const int32_t* table, uint32_t buffer_size, uint32_t* occ2, int32_t* d_count, void* scratch,
                     void* stream);
/* rec_global.w = table value (global point index); rec_local.w = global2local[] of it after the
 * travel-distance filter of model/neural_points.py:480-488 (-1 = rejected).  Either may be NULL.
 * Hazard H13: the reference remaps points OUTSIDE the local map to local index 1, not -1
 * (neural_points.py:301, full_like on a bool mask); such candidates are stored as -2 and the decode
 * kernel resolves them to index 1 with the position of local point 1, exactly as the reference does.
 * local_mask = uint8[m+1] flags of reset_local_map. */
int pinq_index_pack(const int32_t* table, uint32_t buffer_size, const uint32_t* occ2, const float* map_xyz,
                    const int64_t* ts_create, const int64_t* global2local, const uint8_t* local_mask,
                    const float* travel_dist,
                    int64_t cur_ts, float diff_travel_dist_local, float* rec_global4, float* rec_local4,
                    void* stream);

/* ---- brick index: spatially coherent read view (csrc/brick.cuh) ---------------------------------
 * A brick = 4x4x4 voxels; open-addressing table keyed by brick coordinates -> {64-bit voxel occupancy,
 * rank of the brick's first record}; records of a brick are contiguous.  Holds exactly the points that
 * are findable through the slot table (table[slot(cell(p))] == p), so a lookup by voxel returns what
 * the reference's probe of that voxel returns, minus far-collision aliases that the reference rejects
 * by distance anyway (caller checks the no-near-collision property of buffer_size first).
 * rec_*4 = slot-rank-ordered records of pinq_index_pack, d_count = *pinq_index_build count,
 * tab = pinq_brick_capacity(m) * 32 bytes, d_counters2 = int32[2] {records placed, invariant violations}. */
int64_t pinq_brick_capacity(int64_t m);
int pinq_brick_build(const float* rec_global4, const float* rec_local4, const int32_t* d_count, int64_t m, float res,
                     uint32_t flags, void* tab, uint32_t cap, int32_t* d_counters2, float* brick_rec_global4,
                     float* brick_rec_local4, void* stream);

/* ---- neighbour search -------------------------------------------------------------------- */
typedef struct pinq_map_view {
    const uint32_t* occ;        /* [nwords][2] occupancy index (pinq_index_build)             */
    const float* rec4;          /* [n_occupied][4] rank-ordered records: local flavour for
                                   PINQ_F_LOCAL queries, global flavour otherwise             */
    uint32_t buffer_size;
    float resolution;           /* voxel_size_m                                               */
    float max_valid_dist2;      /* model/neural_points.py:457                                 */
    int32_t kn;                 /* stencil size (81 for run_kitti)                            */
    const int32_t* stencil_dx;  /* [kn,3] int32 neighbour cell offsets (neural_points.py:439) */
    int64_t m;                  /* number of neural points (global)                           */
    /* optional brick index (pinq_brick_build); NULL = probe the slot view.  Used by the fused query
       and the mesher; the API-form search / query_certainty always read the slot view. */
    const void* brick_tab;      /* [brick_cap] 32-byte entries                                */
    const float* brick_rec4;    /* [n_occupied][4] records in brick order (same flavour as rec4) */
    uint32_t brick_cap;         /* power of two                                               */
    int32_t stencil_radius;     /* max |dx| over the stencil (<= 2 for the brick path)        */
    const uint64_t* stencil_masks; /* [512] from pinq_stencil_masks_build for this stencil, or NULL =
                                   rebuilt inside every query launch (costs one small kernel)    */
} pinq_map_view;

/* Brick-voxel membership masks of a stencil (64 box phases x 8 bricks, one 64-bit word each): depends on
 * stencil_dx only, so set_search_neighborhood() (neural_points.py:430-457) can build it once. */
int pinq_stencil_masks_build(const int32_t* stencil_dx, int32_t kn, int32_t stencil_radius, uint64_t* out512,
                             void* stream);

/* model/neural_points.py:459-509 -- standalone API form: writes dist2[n,kn] f32 and idx[n,kn] i64
 * (global indices).  time_filtering reads ts_create / travel_dist exactly as the reference does. */
int pinq_neighbor_search(const pinq_map_view* map, const float* xyz, int64_t n, uint32_t flags,
                         const float* map_xyz, const int64_t* ts_create, const float* travel_dist,
                         int64_t cur_ts, float diff_travel_dist_local, int time_filtering, float* out_dist2,
                         int64_t* out_idx, void* stream);

/* model/neural_points.py:511-525 -- max certainty over the stencil (global certainties). */
int pinq_query_certainty(const pinq_map_view* map, const float* xyz, int64_t n, uint32_t flags,
                         const float* map_xyz, const float* certainties, float* out, void* stream);

/* ---- fused query -------------------------------------------------------------------------- */
typedef struct pinq_decoder_view {
    const float* W1; /* layers.0.weight [H, F+3] row-major (model/decoder.py:45) */
    const float* b1; /* layers.0.bias   [H]                                      */
    const float* W2; /* lout.weight     [1, H]                                   */
    const float* b2; /* lout.bias       [1]                                      */
    int32_t hidden;  /* H = 64                                                   */
    int32_t in_dim;  /* F + 3 = 11                                               */
    float sdf_scale; /* model/decoder.py:51-52, hazard H12                       */
} pinq_decoder_view;

typedef struct pinq_query_args {
    pinq_map_view map;
    pinq_decoder_view dec;
    const float* xyz;          /* [n,3] query points                                              */
    int64_t n;
    uint32_t flags;            /* PINQ_F_*                                                        */
    int32_t k;                 /* query_nn_k = 6                                                  */
    int32_t feature_dim;       /* 8                                                               */
    const float* features;     /* [Mx+1, F]: local_geo_features (LOCAL) or geo_features           */
    const float* point_xyz;    /* [Mx,3] positions indexed like `features` (hazard H13 only)      */
    float* certainties;        /* [Mx]: read for queried certainty; += w in TRAINING mode         */
    const float* quats;        /* [Mx,4] or NULL (needed with PINQ_F_ROTATE)                      */
    int64_t* ts_update;        /* [Mx] or NULL: amax with query_ts in TRAINING mode               */
    const int64_t* query_ts;   /* [n] or NULL                                                     */
    /* outputs (any may be NULL) */
    float* sdf;                /* [n]   sum_k w_k s_k                 (tracker.py:245, mapper.py:468) */
    float* grad_x;             /* [n,3] d sdf / d x                   (tools.py:174-184)              */
    float* sdf_std;            /* [n]                                 (tracker.py:246-247)            */
    int32_t* nn_count;         /* [n]   valid candidates over the whole stencil (neural_points.py:557) */
    float* certainty;          /* [n]   sum_k w_k c_k                 (neural_points.py:651-656)      */
    int32_t* knn_idx;          /* [n,k] top-k index (local or global), -1 invalid   } written by              */
    float* knn_w;              /* [n,k] normalised IDW weights                       } pinq_query_feature only */
    float* knn_d2;             /* [n,k] squared distances (9e3 invalid)                              */
    void* stage;               /* pinq_stage_bytes(n) or NULL: the search -> decode hand-off of the fused query
                                  (top-k index, IDW weight, x - p_k per neighbour; 124 B/query, opaque,
                                  processing order).  Give a buffer and keep it to call pinq_query_fused_bwd;
                                  NULL = scratch inside `workspace`.                                  */
    uint32_t* relu_mask;       /* pinq_relu_mask_words(n) words or NULL: relu'(pre) of the hidden layer,
                                  one bit per (query, neighbour, hidden unit) in tensor-core fragment
                                  order; saved for pinq_query_fused_bwd (48 B/query, opaque)          */
    int32_t* perm;             /* [n] or NULL.  When given, the fused forward bins the queries by 4x4x4-voxel
                                  brick and processes them in that order (perm[i] = query at position i):
                                  gathers become cache hits, results are unchanged.  knn_idx / knn_w /
                                  knn_d2 / relu_mask are then stored in PROCESSING order; pass the same
                                  perm to pinq_query_fused_bwd.  Ignored by pinq_query_feature and the mesher. */
    void* workspace;           /* pinq_query_workspace_bytes(n) bytes: search -> decode hand-off      */
    const int32_t* skip;       /* NULL, or a device flag: when non-zero at launch time the search / resolve /
                                  decode kernels of this call return at once (pinq_tracking_run)         */
    void* decode_wait_event;   /* NULL, or a cudaEvent_t the DECODE waits for (cudaStreamWaitEvent on `stream`):
                                  binning and the neighbour search read only query points and map geometry, so a
                                  data-parallel caller lets them run under the previous step's gradient exchange /
                                  optimizer (which the decode, reading features and decoder, must follow)   */
    void* zero_fill;           /* NULL, or a device buffer of zero_fill_bytes (a multiple of 16, 16-byte aligned) that the
                                  fused forward sets to zero on its way: the feature-gradient block the backward of this
                                  query accumulates into (replaces the 53 MB memset of `torch.zeros_like(features)`
                                  between forward and backward: one streaming 16-byte store per thread of the
                                  neighbour search, which has the HBM bandwidth to spare).  Done by a plain memset
                                  when the batch has fewer than zero_fill_bytes / 16 queries or takes the slot-probing
                                  search.  Not touched when `skip` is set.  pinq_query_fused_fwd only.           */
    int64_t zero_fill_bytes;
} pinq_query_args;

int64_t pinq_query_workspace_bytes(int64_t n);
int64_t pinq_relu_mask_words(int64_t n);
int64_t pinq_stage_bytes(int64_t n);

/* neural_points.py:528-674 + decoder.py:66-88 + the weighted combine, in one pass.
 * Supported specialisation: k=6, feature_dim=8, hidden=64, one hidden level, weighted_first=False. */
int pinq_query_fused_fwd(const pinq_query_args* args, void* stream);

/* Un-fused API form of query_feature (neural_points.py:574-632): geo_features_vector [n,k,F+3],
 * weight [n,k], nn_count, certainty; same side effects as the fused form. */
int pinq_query_feature(const pinq_query_args* args, float* feat_out, void* stream);

typedef struct pinq_bwd_args {
    pinq_decoder_view dec;
    int64_t n;
    uint32_t flags;            /* unused (the rotated x - p_k comes from the stage)        */
    int32_t k, feature_dim;
    const float* features;     /* [Mx+1,F]                                                 */
    const void* stage;         /* pinq_stage_bytes(n), saved by the forward                 */
    const uint32_t* relu_mask; /* pinq_relu_mask_words(n), saved by the forward            */
    const int32_t* perm;       /* [n] processing order of the forward, or NULL             */
    const float* grad_sdf;     /* [n] dL/dsdf                                              */
    float* grad_features;      /* [Mx+1,F] accumulated (+=), may be NULL                   */
    float* grad_W1;            /* [H,F+3] accumulated, NULL when the decoder is frozen (H11) */
    float* grad_b1;            /* [H]                                                      */
    float* grad_W2;            /* [1,H]                                                    */
    float* grad_b2;            /* [1]                                                      */
} pinq_bwd_args;

/* hand-derived backward of the fused query w.r.t. features and decoder (replaces autograd through
 * neural_points.py:562-632 and decoder.py:66-88 as triggered at utils/mapper.py:572). */
int pinq_query_fused_bwd(const pinq_bwd_args* args, void* stream);

/* ---- Gauss-Newton accumulation (utils/tracker.py:293-394, 468-491) ------------------------
 * One pass over the source points: validity mask (nn >= min_nn, min_g < |g| < max_g, std < max_std),
 * Geman-McClure weights (un-normalised), and the reduction of
 *   out[0..20]  upper triangle of sum w J^T J (row-major, J = [p x g, g])
 *   out[21..26] sum w r J          out[27] sum w      out[28] sum |r|      out[29] valid count
 * in fp64.  `out` (double[32]) must be zeroed by the caller.  valid_out (uint8[n]) optional. */
int pinq_gn_accumulate(const float* xyz, const float* sdf, const float* grad, const float* sdf_std,
                       const int32_t* nn_count, const float* sdf_label /*nullable = 0*/, int64_t n,
                       int32_t min_nn, float min_grad_norm, float max_grad_norm, float max_sdf_std,
                       float gm_dist, float gm_grad, double* out32, uint8_t* valid_out, void* stream);

/* utils/tools.py:386-399 -- out = (T[:3,:3] x + T[:3,3]) with T given as 12 floats (row-major 3x4). */
int pinq_transform_points(const float* xyz, int64_t n, const float* T12, float* out, void* stream);

/* ---- mesher dense grid (utils/mesher.py:41-180) -------------------------------------------
 * Grid coordinates are generated in-kernel: p = float(i)*voxel + origin, z fastest
 * (mesher.py:168-178); linear range [first, first+count) of the nx*ny*nz grid.
 * sdf_out [count] f32, mask_out [count] u8 = nn_count >= min_nn. */
int pinq_mesher_grid_query(const pinq_query_args* args, const float* origin3_host, const int32_t* dims3_host,
                           float voxel, int64_t first, int64_t count, int32_t min_nn, float* sdf_out,
                           uint8_t* mask_out, void* stream);

/* One Gauss-Newton iteration of Tracker.tracking (utils/tracker.py:91-107 + registration_step :277-452) as a
 * single call: cur = T * source (utils/tools.py:386-399; skipped when h_T12 is NULL: query.xyz is read as is),
 * fused query with gradient and std, the reduction above, and the copy of its 32 doubles to pinned host
 * memory.  The caller synchronises the stream, then solves the 6x6 system on the host (tracker.py:468-520). */
typedef struct pinq_tracking_args {
    pinq_query_args query;     /* xyz = [n,3] buffer for the transformed points; sdf, grad_x, sdf_std, nn_count set */
    const float* source_xyz;   /* [n,3] points in the sensor frame                                  */
    const float* sdf_label;    /* [n] or NULL                                                       */
    int32_t min_nn;
    float min_grad_norm, max_grad_norm, max_sdf_std, gm_dist, gm_grad;
    double* gn_out;            /* [32] device                                                       */
    double* h_gn_out;          /* [32] pinned host                                                  */
} pinq_tracking_args;
int pinq_tracking_step(const pinq_tracking_args* args, const float* h_T12, void* stream);

/* A chunk of Gauss-Newton iterations of Tracker.tracking (utils/tracker.py:91-141) WITHOUT returning to the host in
 * between: per iteration  cur = T * source  ->  fused query  ->  normal-equation reduction  ->  one-thread solve
 * (implicit_reg, tracker.py:468-520: LM damping, 6x6 solve, expmap, T = dT @ T in fp64) + the loop's bookkeeping
 * (residual growth, valid-point ratio, convergence thresholds, `converged` one iteration ahead of the break).
 * Once the loop's break condition holds, the remaining launches of the chunk return at once (pinq_query_args.skip).
 * The caller enqueues chunks until state says done (one stream synchronise per chunk instead of per iteration).
 * state: PINQ_TRACK_STATE_DOUBLES doubles -- [0..15] T row-major, [16] last_sdf_residual_cm, [17] sdf_residual_cm;
 * doubles 24..31 hold int32 flags: [0] done, [1] valid_flag, [2] converged, [3] iterations run, [4] valid_point_count,
 * [5] failure reason (1 residual grew beyond max_increment_ratio, 2 too few valid points). */
#define PINQ_TRACK_STATE_DOUBLES 32
typedef struct pinq_tracking_loop {
    double* state;             /* device                                                              */
    const double* h_state_in;  /* pinned host; non-NULL: uploaded before the chunk (first chunk)       */
    double* h_state;           /* pinned host; written after the chunk                                 */
    int32_t first_iter, iters, iter_n;
    int32_t min_valid_points;
    float lm_lambda, term_thre_deg, term_thre_m, min_valid_ratio, max_increment_ratio;
} pinq_tracking_loop;
int pinq_tracking_run(const pinq_tracking_args* args, const pinq_tracking_loop* loop, void* stream);

/* ---- fused Adam (utils/tools.py:89-116; torch.optim.Adam single-tensor semantics) -------- */
int pinq_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                   float beta1, float beta2, float eps, int32_t step, void* stream);

/* ---- fused mapping iterations (utils/mapper.py:425-593) -----------------------------------
 * One call runs `iters` iterations of Mapper.mapping() on the device without returning to the host:
 *   get_batch (mapper.py:323-361)  -> k_map_gather: pool rows of the pre-drawn indices, plus the
 *                                     6 x ceil(bs/decimation) numerical-gradient points (mapper.py:683-733)
 *                                     appended to the same query batch (query_ts = -1: no ts amax, H9)
 *   sdf prediction (mapper.py:461-468)  -> pinq_query_fused_fwd over the bs + 6*ng points
 *   bce + eikonal loss (mapper.py:515-547, loss.py:40-47) -> k_map_loss_grad: dL/dsdf in closed form
 *   cur_loss.backward() (mapper.py:572) -> pinq_query_fused_bwd
 *   opt.step() (mapper.py:573, tools.py:89-116) -> one multi-tensor Adam launch (also zeroes the grads)
 * The optimizer is fresh per call (mapper.py:441): iteration `it` is Adam step it+1 and step 1 ignores
 * what exp_avg / exp_avg_sq held before. */
typedef struct pinq_adam_tensor {
    float* param;
    float* grad;               /* zero on entry; left zero on return                        */
    float* exp_avg;            /* scratch, same size (contents ignored at step 1)           */
    float* exp_avg_sq;
    int64_t n;
} pinq_adam_tensor;

#define PINQ_EIK_ALL 0
#define PINQ_EIK_SURFACE 1
#define PINQ_EIK_FREESPACE 2

typedef struct pinq_mapping_args {
    pinq_query_args query;     /* xyz = batch_xyz, n = bs + 6*ng, query_ts = batch_ts, sdf, nn_count,
                                  stage, relu_mask, workspace set; TRAINING | LOCAL flags          */
    pinq_bwd_args bwd;         /* n as above, grad_sdf = grad_sdf below, grad_* = the adam grads   */
    /* replay pool (mapper.py:59-66) */
    const float* pool_xyz;     /* [P,3] global_coord_pool (or coord_pool after BA)                 */
    const float* pool_label;   /* [P]                                                              */
    const float* pool_weight;  /* [P]                                                              */
    const int64_t* pool_ts;    /* [P]                                                              */
    /* pre-drawn uniform indices, one row per iteration, in the order get_batch draws them */
    const int64_t* draws_hist; /* [iters][bs - bs_new] in [0, P)                                   */
    const int64_t* draws_new;  /* [iters][bs_new] in [0, new_count) or NULL when bs_new == 0       */
    const int64_t* new_idx;    /* [new_count] pool rows of the newly observed samples              */
    int32_t bs, bs_new;
    int32_t decimation;        /* gradient_decimation; ng = ceil(bs / decimation); 0 = no eikonal  */
    int32_t eik_mode;          /* PINQ_EIK_*  (ekional_add_to)                                     */
    int32_t weighted;          /* loss_weight_on                                                   */
    float eps_step;            /* voxel_size_m * num_grad_step_ratio                               */
    float sigma;               /* logistic_gaussian_ratio * sigma_sigmoid_m                        */
    float weight_e;
    float surface_range;       /* surface_sample_range_m (eik_mode != ALL)                         */
    /* per-iteration batch buffers */
    float* batch_xyz;          /* [n,3]                                                            */
    int64_t* batch_ts;         /* [n]                                                              */
    float* batch_label;        /* [bs]                                                             */
    float* batch_weight;       /* [bs]  |weight|                                                   */
    float* grad_sdf;           /* [n]                                                              */
    float* loss;               /* [iters][2] (bce, eikonal) or NULL; zero on entry                 */
    int32_t* eik_count;        /* [iters] selected eikonal samples; zero on entry                  */
    /* optimizer */
    pinq_adam_tensor tensors[8];
    int32_t tensor_count;
    float lr, beta1, beta2, adam_eps, weight_decay;
    /* data-parallel form (all zero / NULL = one device).  The batch above is this rank's contiguous slice
     * [shard_first, shard_first + bs) of a GLOBAL batch of global_bs samples drawn with one shared seed: the losses are
     * normalised by the global counts and the eikonal subset is the global batch's x[::decimation], so the sum of the
     * ranks' gradients is the single-device gradient.  With `peer` (a HOST pointer, see pinq_peer_args below) the local
     * Adam launch is replaced by pinq_peer_exchange with the fused optimizer: tensors[].param / .grad must then be the
     * param / sum regions of this rank's peer block in one common order, epoch = peer->epoch + iteration, the
     * caller zeroes peer->exp_avg / exp_avg_sq, and lr / betas / eps are taken from this struct. */
    int32_t global_bs;
    int32_t shard_first;
    const struct pinq_peer_args* peer;
} pinq_mapping_args;

int pinq_mapping_run(const pinq_mapping_args* args, int32_t iters, void* stream);

/* ---- data-parallel exchange over NVLink peer memory ------------------------------------------
 * Replaces what a data-parallel run has to do between `cur_loss.backward()` and `opt.step()` of Mapper.mapping
 * (utils/mapper.py:572-573; optimizer: utils/tools.py:89-116): the sum of the feature-table and decoder gradients
 * (and of the certainty increments, model/neural_points.py:637-642) and the max of local_point_ts_update (:643-645)
 * over the ranks.  Every rank owns one "peer block" with the same layout -- [signal pad | payload] -- allocated by
 * pinq_peer_alloc (cudaMalloc, so it can be exported), exported as a 64-byte cudaIpc handle, and mapped by the other
 * ranks of the node (one process per GPU).  pinq_peer_exchange launches ONE kernel per rank:
 *   barrier (flags in peer memory) -> this rank's 1/world slice: sum over all ranks' blocks (P2P loads)
 *   [-> Adam on the slice, optimizer state held by the slice owner only] -> result stored into every rank's block
 *   (P2P stores: summed gradients, or updated parameters + zeroed gradients) -> barrier.
 * Ranks must call it in lock-step with the same epoch (1, 2, 3, ...).  The flag barrier pairs CTA b of every rank, so
 * all CTAs of the grid must be resident at once: set `cooperative` (launch-time guarantee).  A peer that does not arrive
 * within timeout_ms sets the error word (pinq_peer_error_offset() bytes into the block, non-zero = failed) instead of hanging.
 * These are the only entry points that allocate or that touch other devices. */
#define PINQ_PEER_MAX 8
#define PINQ_PEER_HANDLE_BYTES 64
#define PINQ_PEER_MAX_GRID 256
#define PINQ_PEER_PAD_BYTES 16384 /* flags [PINQ_PEER_MAX_GRID][8] uint32, error word, padding */
int pinq_peer_alloc(int64_t bytes, void** d_out);          /* zero-filled; bytes >= PINQ_PEER_PAD_BYTES   */
int pinq_peer_free(void* d_block);
int pinq_peer_export(void* d_block, void* h_handle);       /* h_handle: PINQ_PEER_HANDLE_BYTES            */
int pinq_peer_open(const void* h_handle, void** d_out);    /* in another process of the same node         */
int pinq_peer_close(void* d_mapped);
int64_t pinq_peer_error_offset(void);
typedef struct pinq_peer_args {
    int32_t rank, world;       /* world in {1, 2, 4, 8}                                                  */
    void* block[PINQ_PEER_MAX]; /* block[r] = rank r's peer block as mapped in this process               */
    int64_t sum_off, n_sum;    /* fp32 sum region: byte offset (16-byte aligned), element count (% 4)    */
    int64_t max_off, n_max;    /* int64 max region: byte offset (16-byte aligned), element count (% 2)   */
    uint32_t epoch;            /* call counter, identical on every rank, starting at 1                   */
    int32_t grid;              /* CTAs (<= PINQ_PEER_MAX_GRID, identical on every rank); 0 = 148         */
    float timeout_ms;          /* 0 = 5000                                                               */
    /* fused optimizer (n_adam > 0): the first n_adam elements of the sum region are the gradients of the
     * n_adam parameters at param_off of the same block; exp_avg / exp_avg_sq are LOCAL arrays of n_adam
     * floats of which only this rank's slice is used; gradients come back zeroed. */
    int64_t param_off, n_adam;
    float* exp_avg;
    float* exp_avg_sq;
    float lr, beta1, beta2, adam_eps;
    int32_t adam_step;         /* 1-based                                                                */
    int32_t cooperative;       /* non-zero: cudaLaunchCooperativeKernel -- all `grid` CTAs are guaranteed co-resident
                                * (needed as soon as anything else may share the device, e.g. the next step's search
                                * on another stream); 0: plain launch, the caller vouches for co-residency (several
                                * simulated ranks in ONE process must use 0: their kernels wait for each other)     */
    int32_t threads;           /* threads per CTA, identical on every rank: 0 / 512 = stand-alone form; 128 = the form
                                * for an exchange that runs UNDER another kernel (plain launch on a side stream): CTAs of
                                * 128 threads x <= 56 registers fit into what ONE retiring CTA of the neighbour search
                                * frees, a 512-thread CTA needs a completely drained SM                               */
    void* mc_block;            /* NULL, or the NVLS multicast address of the peer blocks (one address that reaches the
                                * block of EVERY rank through NVSwitch; the blocks must then be allocations bound to one
                                * multicast object, e.g. torch.distributed._symmetric_memory): the fp32 sum region is
                                * reduced with multimem.ld_reduce and broadcast with multimem.st -- 2 memory
                                * instructions per 16 bytes of the owned slice instead of 2 x world; the sum is then the
                                * switch's order, not rank order.  Flags and the int64 max region keep using block[] */
} pinq_peer_args;
int pinq_peer_exchange(const pinq_peer_args* args, void* stream);

/* ---- replay pool upkeep (utils/mapper.py:209-263) ------------------------------------------
 * Per frame: append the new samples (mapper.py:209-222), keep the rows inside the window around the sensor
 * (mapper.py:229-233), knock out random survivors above pool_capacity (:237-246), compact all five pool tensors
 * (:248-263).  Three calls around pinq_compact_flags so that the caller owns all memory and draws the random
 * numbers (the reference's torch.randint stream):
 *   pinq_pool_flags    flags[i] = ||g_i - origin||^2 < radius2 over [old pool | new samples]
 *   pinq_compact_flags survivors' row numbers, in order (boolean-mask indexing keeps order)
 *   [pinq_pool_discard flags[kept_index[discarded[k]]] = 0, then compact again]
 *   pinq_pool_move     survivor j <- its row of the old pool or of the new samples (timestamp = new_ts), written
 *                      into `out` (must not alias the inputs); *d_kept_new += survivors among the new samples
 *                      (= filter_mask[-cur_sample_count:].sum(), mapper.py:262).                                  */
typedef struct pinq_pool_view {
    float* coord;          /* [n,3] sensor-frame sample (coord_pool)          */
    float* global_coord;   /* [n,3] world-frame sample (global_coord_pool)    */
    float* sdf_label;      /* [n]                                              */
    float* weight;         /* [n]                                              */
    int64_t* ts;           /* [n] frame id (time_pool); unused for new samples */
} pinq_pool_view;
int pinq_pool_flags(const float* pool_global_xyz, int64_t n_pool, const float* new_global_xyz, int64_t n_new,
                    const float* d_origin3, float radius2, uint8_t* flags, void* stream);
int pinq_pool_discard(uint8_t* flags, const int32_t* kept_index, const int64_t* discarded, int64_t n_discard, void* stream);
int pinq_pool_move(const pinq_pool_view* old_pool, int64_t n_pool, const pinq_pool_view* new_samples, int64_t n_new,
                   int64_t new_ts, const int32_t* kept_index, const int32_t* d_count, const pinq_pool_view* out,
                   int32_t* d_kept_new, void* stream);

/* ---- iso-surface extraction over the mesher's dense grid -----------------------------------
 * Replaces utils/mesher.py:310-337 `mc_mesh` (skimage.measure.marching_cubes(sdf, level, allow_degenerate=False,
 * mask=mask) on the host).  Grid = what pinq_mesher_grid_query writes: x-major, z fastest, dims = {nx, ny, nz}.
 * A cell is processed when all eight of its corners are masked in (mask may be NULL = everything).  Two calls so that
 * the caller owns all memory:
 *   pinq_mc_count  classifies cells and owned edges, scans -> d_counts2 = {vertices, triangles} (device int32[2])
 *   pinq_mc_emit   verts_out float32 [V,3] = origin + (grid position) * voxel, faces_out int32 [T,3] (+ vertex_base:
 *                  lets several slabs share one vertex array).  `dims3_host` describes the slab handed in, `x_first` the
 *                  index of its first x layer in the whole grid and `origin3_host` the WHOLE grid's origin: coordinates
 *                  are computed from global grid indices, so the vertices on the layer two slabs (or two ranks) share
 *                  are bit-identical on both sides.  Deterministic order: vertices by owner grid point then
 *                  axis, triangles by cell; every vertex is emitted once (indexed mesh, no duplicate removal needed
 *                  inside a slab).  Zero-area triangles are NOT removed here (`allow_degenerate=False` is a filter
 *                  on the output, done by the caller).
 * `scratch`: pinq_mc_scratch_bytes(nx * ny * nz) bytes, shared by the two calls.  pinq_mc_tables copies the
 * 256-case table the kernels use (edge ids, 16 per case, -1 padded; triangles per case) for inspection / tests. */
int64_t pinq_mc_scratch_bytes(int64_t n_points);
int pinq_mc_tables(int8_t* tri_out_256x16, uint8_t* ntri_out_256);
int pinq_mc_count(const float* sdf, const uint8_t* mask, const int32_t* dims3_host, float level, void* scratch,
                  int32_t* d_counts2, void* stream);
int pinq_mc_emit(const float* sdf, const int32_t* dims3_host, int32_t x_first, float level, const float* origin3_host,
                 float voxel, const void* scratch, int32_t vertex_base, float* verts_out, int32_t* faces_out, void* stream);

/* Motion undistortion of a sweep (row f1): utils/tools.py:540-567 `deskewing` with the per-point times of
 * dataset/slam_dataset.py:233-257.  pinq_velodyne_point_ts: ts = 0.5 (-atan2(y, x) / pi + 1) for scans without time stamps
 * (KITTI .bin); pinq_deskew, in place: tau = (ts - min ts) / (max ts - min ts) - ts_mid_pose,
 * p <- slerp(I, R, tau) p + tau t with [R | t] = T_last<-cur (row-major 4x4 float32 in device memory; the slerp is
 * roma.rotmat_slerp's: rotation about the axis of R by tau times its angle, extrapolating for tau < 0).
 * d_ts_minmax2 = {min ts, max ts} in device memory (torch.aminmax), `stride` = floats per point (>= 3). */
int pinq_velodyne_point_ts(const float* xyz, int64_t n, int32_t stride, float* ts_out, void* stream);
int pinq_deskew(float* xyz, int64_t n, int32_t stride, const float* ts, const float* d_ts_minmax2, const float* d_pose16,
                float ts_mid_pose, void* stream);

/* utils/data_sampler.py:20-192 -- DataSampler.sample: per ray the measured point, n_surf gaussian samples around it,
 * n_front uniform samples in the free space in front and n_behind behind, with projective SDF labels (-displacement)
 * and weights (distance weighting :127-128, behind drop-off :131-135, negative sign = free space), written in ray-wise
 * order (:169-175).  The random numbers are the caller's (torch.randn / torch.rand, the reference's RNG stream):
 * r_surf[rep * n + ray] ~ N(0,1), r_front / r_behind ~ U(0,1). */
int pinq_ray_samples(const float* xyz, int64_t n, const float* r_surf, const float* r_front, const float* r_behind,
                     int32_t n_surf, int32_t n_front, int32_t n_behind, float surface_range, float free_begin_ratio,
                     float free_end_dist, float max_range, float dist_weight_scale, int32_t dist_weight_on,
                     int32_t behind_dropoff_on, float* out_xyz, float* out_label, float* out_weight, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PINQ_H_ */

"""ctypes binding of libpinq.so (include/pinq.h).  No CPU fallback: a missing library or a non-CUDA
tensor raises immediately."""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libpinq.so")
if os.environ.get("PINQ_LIB"):  # kernel experiments only (scripts/variants.py): another build of the same sources
    _LIB_PATH = os.environ["PINQ_LIB"]

F_LOCAL, F_TRAINING, F_ROTATE, F_RECIP_DIV, F_SKIP_EMPTY, F_MMA_SYNC, F_EXACT_KEYS, F_STENCIL_IN_RANGE = 1, 2, 4, 8, 16, 32, 64, 128

_vp, _i64, _i32, _u32, _u64, _f32 = C.c_void_p, C.c_int64, C.c_int32, C.c_uint32, C.c_uint64, C.c_float


class MapView(C.Structure):
    _fields_ = [("occ", _vp), ("rec4", _vp), ("buffer_size", _u32), ("resolution", _f32), ("max_valid_dist2", _f32),
                ("kn", _i32), ("stencil_dx", _vp), ("m", _i64), ("brick_tab", _vp), ("brick_rec4", _vp), ("brick_cap", _u32),
                ("stencil_radius", _i32), ("stencil_masks", _vp)]


class DecoderView(C.Structure):
    _fields_ = [("W1", _vp), ("b1", _vp), ("W2", _vp), ("b2", _vp), ("hidden", _i32), ("in_dim", _i32),
                ("sdf_scale", _f32)]


class QueryArgs(C.Structure):
    _fields_ = [("map", MapView), ("dec", DecoderView), ("xyz", _vp), ("n", _i64), ("flags", _u32), ("k", _i32),
                ("feature_dim", _i32), ("features", _vp), ("point_xyz", _vp), ("certainties", _vp), ("quats", _vp), ("ts_update", _vp),
                ("query_ts", _vp), ("sdf", _vp), ("grad_x", _vp), ("sdf_std", _vp), ("nn_count", _vp),
                ("certainty", _vp), ("knn_idx", _vp), ("knn_w", _vp), ("knn_d2", _vp), ("stage", _vp), ("relu_mask", _vp), ("perm", _vp), ("workspace", _vp), ("skip", _vp), ("decode_wait_event", _vp),
                ("zero_fill", _vp), ("zero_fill_bytes", _i64)]


class BwdArgs(C.Structure):
    _fields_ = [("dec", DecoderView), ("n", _i64), ("flags", _u32), ("k", _i32), ("feature_dim", _i32),
                ("features", _vp), ("stage", _vp), ("relu_mask", _vp), ("perm", _vp), ("grad_sdf", _vp), ("grad_features", _vp), ("grad_W1", _vp), ("grad_b1", _vp), ("grad_W2", _vp),
                ("grad_b2", _vp)]


class TrackingArgs(C.Structure):
    _fields_ = [("query", QueryArgs), ("source_xyz", _vp), ("sdf_label", _vp), ("min_nn", _i32), ("min_grad_norm", _f32),
                ("max_grad_norm", _f32), ("max_sdf_std", _f32), ("gm_dist", _f32), ("gm_grad", _f32), ("gn_out", _vp),
                ("h_gn_out", _vp)]


class TrackingLoop(C.Structure):
    _fields_ = [("state", _vp), ("h_state_in", _vp), ("h_state", _vp), ("first_iter", _i32), ("iters", _i32), ("iter_n", _i32),
                ("min_valid_points", _i32), ("lm_lambda", _f32), ("term_thre_deg", _f32), ("term_thre_m", _f32),
                ("min_valid_ratio", _f32), ("max_increment_ratio", _f32)]


TRACK_STATE_DOUBLES = 32


class AdamTensor(C.Structure):
    _fields_ = [("param", _vp), ("grad", _vp), ("exp_avg", _vp), ("exp_avg_sq", _vp), ("n", _i64)]


EIK_ALL, EIK_SURFACE, EIK_FREESPACE = 0, 1, 2

COMPACT_STATIC_MAX_N = 16 * 1024 * 1024  # PINQ_COMPACT_STATIC_MAX_N

PEER_MAX, PEER_HANDLE_BYTES, PEER_PAD_BYTES, PEER_MAX_GRID = 8, 64, 16384, 256


class PeerArgs(C.Structure):
    _fields_ = [("rank", _i32), ("world", _i32), ("block", _vp * PEER_MAX), ("sum_off", _i64), ("n_sum", _i64),
                ("max_off", _i64), ("n_max", _i64), ("epoch", _u32), ("grid", _i32), ("timeout_ms", _f32),
                ("param_off", _i64), ("n_adam", _i64), ("exp_avg", _vp), ("exp_avg_sq", _vp), ("lr", _f32), ("beta1", _f32),
                ("beta2", _f32), ("adam_eps", _f32), ("adam_step", _i32), ("cooperative", _i32), ("threads", _i32), ("mc_block", _vp)]


class MappingArgs(C.Structure):
    _fields_ = [("query", QueryArgs), ("bwd", BwdArgs), ("pool_xyz", _vp), ("pool_label", _vp), ("pool_weight", _vp),
                ("pool_ts", _vp), ("draws_hist", _vp), ("draws_new", _vp), ("new_idx", _vp), ("bs", _i32), ("bs_new", _i32),
                ("decimation", _i32), ("eik_mode", _i32), ("weighted", _i32), ("eps_step", _f32), ("sigma", _f32),
                ("weight_e", _f32), ("surface_range", _f32), ("batch_xyz", _vp), ("batch_ts", _vp), ("batch_label", _vp),
                ("batch_weight", _vp), ("grad_sdf", _vp), ("loss", _vp), ("eik_count", _vp), ("tensors", AdamTensor * 8),
                ("tensor_count", _i32), ("lr", _f32), ("beta1", _f32), ("beta2", _f32), ("adam_eps", _f32),
                ("weight_decay", _f32), ("global_bs", _i32), ("shard_first", _i32), ("peer", C.POINTER(PeerArgs))]


class PoolView(C.Structure):
    _fields_ = [("coord", _vp), ("global_coord", _vp), ("sdf_label", _vp), ("weight", _vp), ("ts", _vp)]


# name -> (restype, argtypes); mirrors include/pinq.h one to one
SIGNATURES = {
    "pinq_last_error": (C.c_char_p, []),
    "pinq_version": (C.c_int, []),
    "pinq_table_clear": (C.c_int, [_vp, _u32, _vp]),
    "pinq_hash_points": (C.c_int, [_vp, _i64, _f32, _u32, _u32, _vp, _vp]),
    "pinq_table_gather": (C.c_int, [_vp, _vp, _i64, _vp, _vp]),
    "pinq_table_scatter_last_wins": (C.c_int, [_vp, _vp, _vp, _i64, _vp]),
    "pinq_update_mask": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _i64, _vp, _i64, _f32, _f32, _vp, _vp]),
    "pinq_compact_scratch_bytes": (_i64, [_i64]),
    "pinq_compact_flags": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _vp]),
    "pinq_compact_flags_static": (C.c_int, [_vp, _i64, _vp, _vp, _vp]),
    "pinq_vds_prepare": (C.c_int, [_vp, _i64, _f32, _vp, C.c_int, _vp, _vp, _vp, _vp]),
    "pinq_vds_select": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _u64, _vp]),
    "pinq_vds_emit": (C.c_int, [_vp, _vp, _u64, _i64, _vp, _vp, _vp, _vp]),
    "pinq_local_mask": (C.c_int, [_vp, _vp, _vp, _i64, _vp, _vp, _i64, _f32, _f32, C.c_int, _vp, _vp]),
    "pinq_index_words": (_i64, [_u32]),
    "pinq_index_scratch_bytes": (_i64, [_u32]),
    "pinq_index_build": (C.c_int, [_vp, _u32, _vp, _vp, _vp, _vp]),
    "pinq_index_pack": (C.c_int, [_vp, _u32, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _f32, _vp, _vp, _vp]),
    "pinq_brick_capacity": (_i64, [_i64]),
    "pinq_brick_build": (C.c_int, [_vp, _vp, _vp, _i64, _f32, _u32, _vp, _u32, _vp, _vp, _vp, _vp]),
    "pinq_kernel_timing": (None, [C.c_int]),
    "pinq_kernel_times": (C.c_int, [C.POINTER(C.c_float), C.c_int]),
    "pinq_query_workspace_bytes": (_i64, [_i64]),
    "pinq_relu_mask_words": (_i64, [_i64]),
    "pinq_stage_bytes": (_i64, [_i64]),
    "pinq_neighbor_search": (C.c_int, [C.POINTER(MapView), _vp, _i64, _u32, _vp, _vp, _vp, _i64, _f32, C.c_int, _vp,
                                       _vp, _vp]),
    "pinq_query_certainty": (C.c_int, [C.POINTER(MapView), _vp, _i64, _u32, _vp, _vp, _vp, _vp]),
    "pinq_query_fused_fwd": (C.c_int, [C.POINTER(QueryArgs), _vp]),
    "pinq_query_feature": (C.c_int, [C.POINTER(QueryArgs), _vp, _vp]),
    "pinq_query_fused_bwd": (C.c_int, [C.POINTER(BwdArgs), _vp]),
    "pinq_gn_accumulate": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _f32, _f32, _f32, _f32, _f32, _vp,
                                     _vp, _vp]),
    "pinq_transform_points": (C.c_int, [_vp, _i64, _vp, _vp, _vp]),
    "pinq_mesher_grid_query": (C.c_int, [C.POINTER(QueryArgs), C.POINTER(_f32), C.POINTER(_i32), _f32, _i64, _i64,
                                         _i32, _vp, _vp, _vp]),
    "pinq_adam_step": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _f32, _f32, _f32, _f32, _i32, _vp]),
    "pinq_stencil_masks_build": (C.c_int, [_vp, _i32, _i32, _vp, _vp]),
    "pinq_tracking_step": (C.c_int, [C.POINTER(TrackingArgs), C.POINTER(_f32), _vp]),
    "pinq_mapping_run": (C.c_int, [C.POINTER(MappingArgs), _i32, _vp]),
    "pinq_tracking_run": (C.c_int, [C.POINTER(TrackingArgs), C.POINTER(TrackingLoop), _vp]),
    "pinq_peer_alloc": (C.c_int, [_i64, C.POINTER(_vp)]),
    "pinq_peer_free": (C.c_int, [_vp]),
    "pinq_peer_export": (C.c_int, [_vp, _vp]),
    "pinq_peer_open": (C.c_int, [_vp, C.POINTER(_vp)]),
    "pinq_peer_close": (C.c_int, [_vp]),
    "pinq_peer_error_offset": (_i64, []),
    "pinq_peer_exchange": (C.c_int, [C.POINTER(PeerArgs), _vp]),
    "pinq_ray_samples": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _i32, _i32, _i32, _f32, _f32, _f32, _f32, _f32, _i32, _i32, _vp, _vp, _vp, _vp]),
    "pinq_mc_scratch_bytes": (_i64, [_i64]),
    "pinq_mc_tables": (C.c_int, [_vp, _vp]),
    "pinq_mc_count": (C.c_int, [_vp, _vp, C.POINTER(_i32), _f32, _vp, _vp, _vp]),
    "pinq_mc_emit": (C.c_int, [_vp, C.POINTER(_i32), _i32, _f32, C.POINTER(_f32), _f32, _vp, _i32, _vp, _vp, _vp]),
    "pinq_velodyne_point_ts": (C.c_int, [_vp, _i64, _i32, _vp, _vp]),
    "pinq_deskew": (C.c_int, [_vp, _i64, _i32, _vp, _vp, _vp, _f32, _vp]),
    "pinq_pool_flags": (C.c_int, [_vp, _i64, _vp, _i64, _vp, _f32, _vp, _vp]),
    "pinq_pool_discard": (C.c_int, [_vp, _vp, _vp, _i64, _vp]),
    "pinq_pool_move": (C.c_int, [C.POINTER(PoolView), _i64, C.POINTER(PoolView), _i64, _i64, _vp, _vp, C.POINTER(PoolView), _vp, _vp]),
}

_lib = None


def lib_path() -> str:
    return _LIB_PATH


def load():
    """dlopen libpinq.so and declare every prototype.  Raises if the library is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise RuntimeError(
            "libpinq.so is missing (%s): build it with `python -m pin_slam_b200.build` "
            "(there is no CPU or PyTorch fallback for the neural-point query path)" % _LIB_PATH)
    # a library built from other sources than the ones in the tree must not be used silently (the digest of the .cu /
    # .cuh / pinq.h files and the nvcc flags is written next to the .so by build.py; both are git-ignored build artefacts)
    if os.environ.get("PINQ_SKIP_STAMP_CHECK") != "1" and not os.environ.get("PINQ_LIB"):
        from . import build as _build
        if not _build.is_current():
            try:
                _build.build()
            except Exception as e:
                raise RuntimeError("libpinq.so is stale (sources changed since it was built) and rebuilding failed: %s" % e)
    lib = C.CDLL(_LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class PinqError(RuntimeError):
    pass


def check(rc: int):
    if rc != 0:
        raise PinqError("libpinq error %d: %s" % (rc, load().pinq_last_error().decode()))


def ptr(t):
    """Device pointer of a CUDA tensor (None -> NULL).  Refuses CPU tensors: no host path exists."""
    if t is None:
        return None
    if not t.is_cuda:
        raise PinqError("libpinq needs CUDA tensors (got a %s tensor); there is no CPU fallback" % t.device)
    if not t.is_contiguous():
        raise PinqError("libpinq needs contiguous tensors")
    return t.data_ptr()


def stream():
    return torch.cuda.current_stream().cuda_stream

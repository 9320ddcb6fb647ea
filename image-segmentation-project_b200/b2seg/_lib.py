"""ctypes binding of libb2seg.so (C ABI declared in include/b2seg.h).

There is deliberately no fallback: if the shared library is missing or a call
fails, a B2SegError is raised.  Nothing under ``oracle/`` is ever imported here.
"""
from __future__ import annotations

import ctypes as C
import os

ABI_VERSION = 3
_HERE = os.path.dirname(os.path.abspath(__file__))
# B2SEG_LIB_PATH: development A/B runs against another build of the same ABI (never a fallback)
LIB_PATH = os.environ.get("B2SEG_LIB_PATH") or os.path.join(_HERE, "libb2seg.so")


class B2SegError(RuntimeError):
    pass


# ---------------------------------------------------------------- structs ----
class LossSpec(C.Structure):
    _fields_ = [
        ("w_dice", C.c_float), ("w_logcosh_dice", C.c_float), ("dice_smooth", C.c_float),
        ("w_tversky", C.c_float), ("tv_alpha", C.c_float), ("tv_beta", C.c_float),
        ("tv_smooth", C.c_float), ("tv_gamma", C.c_float),
        ("w_sens_spec", C.c_float), ("ss_w", C.c_float), ("ss_eps", C.c_float),
        ("w_bce", C.c_float), ("bce_mode", C.c_int32), ("bce_beta", C.c_float),
        ("w_focal", C.c_float), ("focal_alpha", C.c_float), ("focal_gamma", C.c_float),
        ("w_explog_dice", C.c_float), ("w_explog_cross", C.c_float),
        ("el_gamma_dice", C.c_float), ("el_gamma_cross", C.c_float),
    ]


STATS_LEN = 16
COEF_LEN = 8
S_P, S_T, S_PT, S_BCE, S_BCE_T, S_FOCAL, S_PB, S_PB_T, S_N, S_LOSS, S_TP, S_FP, S_FN, S_TN = range(14)

AUG_HFLIP, AUG_VFLIP, AUG_ROT90, AUG_TRANSPOSE = 0, 1, 2, 3
INTERP_NEAREST, INTERP_LINEAR = 0, 1
BORDER_CONSTANT, BORDER_REFLECT, BORDER_REFLECT101 = 0, 1, 2

MAX_TAPS = 9
MAX_SUB = 4


class Tensor4(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("n", C.c_int32), ("h", C.c_int32), ("w", C.c_int32),
                ("c", C.c_int32), ("sn", C.c_int64), ("sh", C.c_int64), ("sw", C.c_int64)]


class Tap(C.Structure):
    _fields_ = [("src", C.c_int8), ("dh", C.c_int8), ("dw", C.c_int8), ("wtap", C.c_int8)]


class TapSub(C.Structure):
    _fields_ = [("ntaps", C.c_int32), ("taps", Tap * MAX_TAPS), ("out_offset", C.c_int64)]


class TapGemmDesc(C.Structure):
    _fields_ = [
        ("src", Tensor4 * 4), ("nsrc", C.c_int32), ("nsub", C.c_int32),
        ("sub", TapSub * MAX_SUB),
        ("weights", C.c_void_p), ("wtaps", C.c_int32), ("ngemm", C.c_int32), ("cin", C.c_int32),
        ("on", C.c_int32), ("oh", C.c_int32), ("ow", C.c_int32),
        ("out", C.c_void_p), ("osn", C.c_int64), ("osh", C.c_int64), ("osw", C.c_int64),
        ("cq", C.c_int32), ("oqh", C.c_int64), ("oqw", C.c_int64),
        ("bias", C.c_void_p), ("stats", C.c_void_p), ("accumulate", C.c_int32),
        ("stats_fixed", C.c_int32),
    ]


class WgradTap(C.Structure):
    _fields_ = [("src", C.c_int8), ("dh", C.c_int8), ("dw", C.c_int8), ("pad", C.c_int8),
                ("out_base", C.c_int64)]


class WgradDesc(C.Structure):
    _fields_ = [
        ("a_src", Tensor4 * 4), ("na_src", C.c_int32), ("b_src", Tensor4),
        ("ntaps", C.c_int32), ("taps", WgradTap * MAX_TAPS),
        ("pn", C.c_int32), ("ph", C.c_int32), ("pw", C.c_int32),
        ("grad", C.c_void_p), ("row_stride", C.c_int64), ("col_stride", C.c_int64),
        ("accumulate", C.c_int32), ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
    ]


class EwDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("p", C.c_void_p * 18), ("i", C.c_int64 * 12), ("f", C.c_float * 4)]


class WarpItem(C.Structure):
    _fields_ = [("minv", C.c_double * 6), ("interp", C.c_int32 * 4), ("cval", C.c_int32 * 4),
                ("border", C.c_int32), ("pad", C.c_int32)]


class PackJob(C.Structure):
    _fields_ = [("inp", C.c_void_p), ("out", C.c_void_p), ("T", C.c_int32), ("A", C.c_int32),
                ("B", C.c_int32), ("pad", C.c_int32), ("sa", C.c_int64), ("sb", C.c_int64),
                ("st", C.c_int64), ("start", C.c_int64)]


EW_MAX_PARTIAL_ROWS = 592
(EW_BN_FINALIZE, EW_BN_APPLY, EW_BN_BWD_REDUCE, EW_BN_BWD_FINALIZE, EW_BN_BWD_APPLY, EW_MAXPOOL_FWD,
 EW_MAXPOOL_BWD, EW_STEM_FWD, EW_STEM_WGRAD, EW_DOT1_FWD, EW_DOT1_BWD, EW_COLSUM_FINALIZE, EW_COLSUM,
 EW_PACK, EW_GATE_APPLY, EW_GATE_BWD, EW_SE_POOL, EW_SE_MLP, EW_SE_SCALE, EW_SE_BWD_REDUCE,
 EW_SE_MLP_BWD, EW_SE_BWD_APPLY, EW_FILL, EW_BN_APPLY_FUSED, EW_BN_BWD_FUSED) = range(1, 26)


# ------------------------------------------------------------- signatures ----
_vp, _i, _i64, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_size_t
_SIGS = {
    "b2seg_abi_version": (C.c_int, []),
    "b2seg_last_error": (C.c_char_p, []),
    "b2seg_device_info": (C.c_int, [C.c_char_p, _i, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "b2seg_seg_reduce_workspace_bytes": (_sz, []),
    "b2seg_seg_loss_fwd": (C.c_int, [_vp, _vp, _vp, _i64, C.POINTER(LossSpec), _vp, _vp, _vp, _vp, _vp]),
    "b2seg_seg_loss_finish": (C.c_int, [C.POINTER(LossSpec), _i64, _vp, _vp, _vp, _vp]),
    "b2seg_seg_loss_bwd": (C.c_int, [_vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp]),
    "b2seg_seg_metrics_from_probs": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _vp]),
    "b2seg_seg_metrics_from_probs_noalloc": (C.c_int, [_vp, _vp, _i64, _vp, _vp]),
    "b2seg_loss_boundary_weight_map": (C.c_int, [_vp, _vp, _i, _i, _i, _i, _i, C.POINTER(C.c_float), C.c_float, _vp]),
    "b2seg_loss_edt_workspace_bytes": (C.c_size_t, [_i, _i, _i]),
    "b2seg_loss_edt_f32": (C.c_int, [_vp, _vp, _i, _i, _i, _i, C.c_float, _vp, _vp]),
    "b2seg_aug_permute_u8": (C.c_int, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "b2seg_aug_warp_affine_u8": (C.c_int, [_vp, _vp, _i, _i, _i, C.POINTER(C.c_double), _i, _i, _i, _vp]),
    "b2seg_aug_remap_u8": (C.c_int, [_vp, _vp, _i, _i, _i, _vp, _vp, _i, _i, _i, _vp]),
    "b2seg_aug_warp_affine_batch_u8": (C.c_int, [_vp, _vp, _i, _i, _i, _i, _vp, _vp]),
    "b2seg_aug_remap_batch_u8": (C.c_int, [_vp, _vp, _i, _i, _i, _i, _vp, _vp, _i64, _vp, _vp]),
    "b2seg_aug_lut_u8": (C.c_int, [_vp, _vp, _i64, _vp, _vp]),
    "b2seg_aug_hist_u8": (C.c_int, [_vp, _i64, _i, _vp, _vp]),
    "b2seg_aug_affine_lut_u8": (C.c_int, [_vp, _vp, _i64, _vp, C.c_double, C.c_double, _vp]),
    "b2seg_aug_fill_rects_u8": (C.c_int, [_vp, _i, _i, _vp, _i, _i, _vp]),
    "b2seg_aug_gauss_noise_u8": (C.c_int, [_vp, _vp, _i64, C.c_float, C.c_float, C.c_uint64, _vp]),
    "b2seg_aug_filter2d_u8": (C.c_int, [_vp, _vp, _i, _i, _i, _vp, _i, _vp]),
    "b2seg_aug_pad_u8": (C.c_int, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "b2seg_aug_uniform_f32": (C.c_int, [_vp, _i64, C.c_uint64, _vp]),
    "b2seg_aug_blur1d_f32": (C.c_int, [_vp, _vp, _i, _i, _i, _vp, _i, _i, C.c_float, _i, _vp]),
    "b2seg_resize_nearest_u8": (C.c_int, [_vp, _i, _i, _i, _vp, _vp, _i, _i, _vp, _vp, _vp]),
    "b2seg_resize_linear_u8": (C.c_int, [_vp, _i, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "b2seg_resize_areafast_u8": (C.c_int, [_vp, _i, _i, _i, _vp, _i, _i, _vp]),
    "b2seg_resize_area_u8": (C.c_int, [_vp, _i, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "b2seg_pre_percentile_lut": (C.c_int, [_vp, _i, _i64, C.c_double, C.c_double, _vp, _vp, _vp]),
    "b2seg_pre_clahe_workspace_bytes": (_sz, [_i, _i, _i]),
    "b2seg_pre_clahe_normalize": (C.c_int, [_vp, _i, _i, _i, _vp, C.c_double, _i, _i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "b2seg_pre_scale255_f32": (C.c_int, [_vp, _i, _i64, _vp, _vp]),
    "b2seg_adam_step": (C.c_int, [_vp, _i, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float,
                                  C.c_double, _vp]),
    "b2seg_op_run": (C.c_int, [_vp, _vp]),
    "b2seg_op_destroy": (C.c_int, [_vp]),
    "b2seg_op_launches": (C.c_int, [_vp]),
    "b2seg_seq_create": (C.c_int, [C.POINTER(_vp)]),
    "b2seg_seq_append": (C.c_int, [_vp, _vp]),
    "b2seg_seq_append_lane": (C.c_int, [_vp, _vp, C.c_int]),
    "b2seg_seq_join": (C.c_int, [_vp]),
    "b2seg_seq_run": (C.c_int, [_vp, _vp]),
    "b2seg_seq_launches": (C.c_int, [_vp]),
    "b2seg_seq_is_graph": (C.c_int, [_vp]),
    "b2seg_seq_run_timed": (C.c_int, [_vp, _vp, C.POINTER(C.c_float), _i]),
    "b2seg_seq_destroy": (C.c_int, [_vp]),
    "b2seg_tapgemm_create": (C.c_int, [C.POINTER(TapGemmDesc), C.POINTER(_vp)]),
    "b2seg_tapgemm_stat_rows": (C.c_int, [_i, _i, _i]),
    "b2seg_wgrad_workspace_bytes": (_sz, [_i, _i, _i, _i, _i, _i]),
    "b2seg_wgrad_create": (C.c_int, [C.POINTER(WgradDesc), C.POINTER(_vp)]),
    "b2seg_ew_create": (C.c_int, [C.POINTER(EwDesc), C.POINTER(_vp)]),
    "b2seg_ew_partial_rows": (C.c_int, [_i, _i64, _i]),
    "b2seg_stem_stat_rows": (C.c_int, [_i, _i, _i]),
    "b2seg_stem_wgrad_rows": (C.c_int, [_i, _i, _i]),
}

_lib = None


def exported_symbols():
    """Names include/b2seg.h declares (used by the CPU-side ABI test)."""
    return sorted(_SIGS)


def register(name, restype, argtypes):
    _SIGS[name] = (restype, argtypes)
    if _lib is not None:
        fn = getattr(_lib, name)
        fn.restype, fn.argtypes = restype, argtypes


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B2SegError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` (nvcc, sm_100a).  b2seg has no CPU or PyTorch fallback.")
        handle = C.CDLL(LIB_PATH)
        for name, (restype, argtypes) in _SIGS.items():
            fn = getattr(handle, name)  # AttributeError here = ABI mismatch, fail loudly
            fn.restype, fn.argtypes = restype, argtypes
        ver = handle.b2seg_abi_version()
        if ver != ABI_VERSION:
            raise B2SegError(f"libb2seg ABI {ver} != expected {ABI_VERSION}")
        _lib = handle
    return _lib


def check(rc: int):
    if rc != 0:
        msg = lib().b2seg_last_error().decode("utf-8", "replace")
        raise B2SegError(f"libb2seg error {rc}: {msg}")


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise B2SegError("b2seg kernels run on a Blackwell GPU only; got a CPU tensor "
                             "(there is no CPU fallback)")

"""ctypes binding of the fod_b200 C ABI (include/fod_b200.h) — the stub INTEGRATION.md describes.

The product path has no fallback: if `libfod_b200.so` is missing or a call returns a non-zero status the
caller gets an exception (`FodError`).  Build the library with `python -c "import __graft_entry__ as g; g.build()"`
or `make -C future-object-detection_b200/csrc`.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfod_b200.so")
ABI_VERSION = 13

FOD_OK, FOD_E_ARG, FOD_E_TMAP, FOD_E_LAUNCH, FOD_E_NUMERIC = 0, 1, 2, 3, 4


class FodError(RuntimeError):
    pass


class Dropout(C.Structure):
    """fod_dropout (include/fod_b200.h)."""
    _fields_ = [("rng_state", C.c_void_p), ("p", C.c_float), ("call", C.c_int32)]


class LinearArgs(C.Structure):
    _fields_ = [
        ("x", C.c_void_p), ("w", C.c_void_p), ("bias", C.c_void_p), ("residual", C.c_void_p),
        ("ln_gamma", C.c_void_p), ("ln_beta", C.c_void_p), ("out_bf16", C.c_void_p), ("out_f32", C.c_void_p),
        ("pre_ln_bf16", C.c_void_p), ("ln_mean", C.c_void_p), ("ln_rstd", C.c_void_p),
        ("ldx", C.c_int64), ("ld_res", C.c_int64), ("ld_out", C.c_int64),
        ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32), ("res_row_mod", C.c_int32), ("relu", C.c_int32),
        ("ln_eps", C.c_float),
        ("relu_mask", C.c_void_p), ("col_scale", C.c_void_p), ("ldw", C.c_int64),
        ("a_mn", C.c_int32), ("b_mn", C.c_int32), ("split_k", C.c_int32), ("atomic_out", C.c_int32),
        ("colsum_out", C.c_void_p), ("drop", Dropout), ("mask_scale", C.c_float), ("pad1", C.c_int32),
    ]


class AttnArgs(C.Structure):
    _fields_ = [
        ("q1", C.c_void_p), ("k1", C.c_void_p), ("q2", C.c_void_p), ("k2", C.c_void_p), ("v", C.c_void_p),
        ("out", C.c_void_p), ("lse", C.c_void_p),
        ("ldq1", C.c_int64), ("ldk1", C.c_int64), ("ldq2", C.c_int64), ("ldk2", C.c_int64), ("ldv", C.c_int64),
        ("ldo", C.c_int64),
        ("bsq1", C.c_int64), ("bsk1", C.c_int64), ("bsq2", C.c_int64), ("bsk2", C.c_int64), ("bsv", C.c_int64),
        ("bso", C.c_int64),
        ("B", C.c_int32), ("H", C.c_int32), ("Lq", C.c_int32), ("Lk", C.c_int32), ("scale", C.c_float),
        ("pad0", C.c_int32), ("drop", Dropout), ("keep_bits", C.c_void_p), ("keep_pitch", C.c_int64),
    ]


class AttnBwdArgs(C.Structure):
    _fields_ = [
        ("fwd", AttnArgs), ("dout", C.c_void_p), ("lddo", C.c_int64), ("bsdo", C.c_int64), ("delta", C.c_void_p),
        ("dq1", C.c_void_p), ("dq2", C.c_void_p), ("dk1", C.c_void_p), ("dk2", C.c_void_p), ("dv", C.c_void_p),
        ("ld_dq", C.c_int64), ("ld_dk", C.c_int64), ("mode", C.c_int32), ("pad0", C.c_int32),
        ("ws", C.c_void_p), ("ws_floats", C.c_int64),
    ]


class MatchArgs(C.Structure):
    _fields_ = [
        ("logits", C.c_void_p), ("boxes", C.c_void_p), ("tgt_boxes", C.c_void_p), ("tgt_labels", C.c_void_p),
        ("tgt_count", C.c_void_p), ("match_tgt", C.c_void_p), ("pred_idx", C.c_void_p), ("tgt_idx", C.c_void_p),
        ("status", C.c_void_p), ("cost_out", C.c_void_p), ("cost_ws", C.c_void_p),
        ("L", C.c_int32), ("B", C.c_int32), ("M", C.c_int32), ("C", C.c_int32), ("NT", C.c_int32),
        ("w_class", C.c_float), ("w_bbox", C.c_float), ("w_giou", C.c_float), ("alpha", C.c_float),
    ]


class SetLossArgs(C.Structure):
    _fields_ = [
        ("logits", C.c_void_p), ("boxes", C.c_void_p), ("match_tgt", C.c_void_p), ("tgt_boxes", C.c_void_p),
        ("tgt_labels", C.c_void_p), ("tgt_count", C.c_void_p), ("num_boxes", C.c_void_p), ("dlogits", C.c_void_p),
        ("dboxes", C.c_void_p), ("partial", C.c_void_p), ("out", C.c_void_p),
        ("L", C.c_int32), ("B", C.c_int32), ("M", C.c_int32), ("C", C.c_int32), ("NT", C.c_int32),
        ("main_level", C.c_int32),
        ("w_ce", C.c_float), ("w_bbox", C.c_float), ("w_giou", C.c_float), ("alpha", C.c_float),
    ]


class FoldDesc(C.Structure):
    """fod_fold_desc (include/fod_b200.h)."""
    _fields_ = [("w", C.c_void_p), ("scale", C.c_void_p), ("out", C.c_void_p), ("O", C.c_int32), ("I", C.c_int32),
                ("HW", C.c_int32), ("pad", C.c_int32)]


class ApArgs(C.Structure):
    _fields_ = [
        ("pred_boxes", C.c_void_p), ("scores", C.c_void_p), ("anno_boxes", C.c_void_p), ("anno_classes", C.c_void_p),
        ("anno_active", C.c_void_p), ("thresholds", C.c_void_p), ("confs", C.c_void_p), ("is_positive", C.c_void_p),
        ("size_cat", C.c_void_p), ("num_annos", C.c_void_p),
        ("B", C.c_int32), ("M", C.c_int32), ("C", C.c_int32), ("N", C.c_int32), ("K", C.c_int32), ("T", C.c_int32),
        ("size_lo", C.c_float), ("size_hi", C.c_float),
    ]


class AdamWArgs(C.Structure):
    _fields_ = [
        ("params", C.c_void_p), ("grads", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
        ("shadow_bf16", C.c_void_p), ("state", C.c_void_p), ("partial", C.c_void_p),
        ("n", C.c_int64), ("group_begin", C.c_int64 * 4), ("group_lr", C.c_float * 4), ("group_wd", C.c_float * 4),
        ("num_groups", C.c_int32), ("zero_grad", C.c_int32),
        ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float), ("max_norm", C.c_float),
    ]


# every symbol include/fod_b200.h declares: name -> (restype, argtypes)
_VP, _I32, _I64, _F = C.c_void_p, C.c_int32, C.c_int64, C.c_float
SYMBOLS = {
    "fod_version": (C.c_int, []),
    "fod_last_error": (C.c_char_p, []),
    "fod_linear_fwd": (C.c_int, [C.POINTER(LinearArgs), _VP]),
    "fod_attn_fwd": (C.c_int, [C.POINTER(AttnArgs), _VP]),
    "fod_attn_bwd": (C.c_int, [C.POINTER(AttnBwdArgs), _VP]),
    "fod_add_ln_fwd": (C.c_int, [_VP, _VP, _I32, _VP, _VP, _VP, _VP, _VP, _I64, _F, _VP]),
    "fod_add_ln_fwd_drop": (C.c_int, [_VP, _VP, _VP, _I32, _VP, _VP, _VP, _VP, _VP, _VP, _I64, _F, C.POINTER(Dropout), _VP]),
    "fod_ln_bwd": (C.c_int, [_VP, _VP, _VP, _VP, _I32, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _I64, _VP]),
    "fod_colsum": (C.c_int, [_VP, _I64, _VP, _I64, _I32, _I32, _VP]),
    "fod_ln_bwd_drop": (C.c_int, [_VP, _VP, _VP, _VP, _I32, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _I64, C.POINTER(Dropout), _VP]),
    "fod_rng_advance": (C.c_int, [_VP, _VP]),
    "fod_maxpool3x3s2_nhwc": (C.c_int, [_VP, _VP, _I32, _I32, _I32, _I32, _VP]),
    "fod_batch_sum": (C.c_int, [_VP, _VP, _I64, _I64, _VP]),
    "fod_frames_to_nhwc_bf16": (C.c_int, [_VP, _I64, _I64, _I32, _I32, _I32, _I32, _VP, _VP]),
    "fod_relu_bwd_add": (C.c_int, [_VP, _VP, _VP, _VP, _I64, _VP]),
    "fod_stem_conv7x7s2": (C.c_int, [_VP, _VP, _VP, _VP, _I32, _I32, _I32, _VP]),
    "fod_stem_conv7x7s2_p4": (C.c_int, [_VP, _VP, _VP, _VP, _I32, _I32, _I32, _VP]),
    "fod_attn_keep_bits": (C.c_int, [_VP, _I32, C.c_float, _I64, _I32, _VP, _I64, _VP]),
    "fod_stem_conv_pool_p4": (C.c_int, [_VP, _VP, _VP, _VP, _I32, _I32, _I32, _VP]),
    "fod_frames_to_nhwc4p_bf16": (C.c_int, [_VP, _I64, _I64, _I32, _I32, _I32, _I32, _VP, _VP]),
    "fod_fold_conv_weights": (C.c_int, [_VP, _I32, _I64, _VP]),
    "fod_conv_wgrad_unfold": (C.c_int, [_VP, _VP, _VP, _I32, _I32, _I32, _VP]),
    "fod_targets_compact": (C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP, _VP, _I32, _I32, _F, _F, _VP]),
    "fod_match": (C.c_int, [C.POINTER(MatchArgs), _VP]),
    "fod_set_loss": (C.c_int, [C.POINTER(SetLossArgs), _VP]),
    "fod_set_loss_static": (C.c_int, [C.POINTER(SetLossArgs), _VP]),
    "fod_ap_match": (C.c_int, [C.POINTER(ApArgs), _VP]),
    "fod_flat_adamw_step": (C.c_int, [C.POINTER(AdamWArgs), _VP]),
    "fod_optim_workspace_floats": (C.c_int, []),
    "fod_cast_bf16": (C.c_int, [_VP, _VP, _I64, _VP]),
}

_lib = None
launches = 0  # number of C-ABI kernel-launching calls made (bench.py reports kernels via ops.LAUNCH_COUNTS)


def load():
    """dlopen the library once; raises FodError when it is missing (no CPU / torch fallback exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FodError(f"{LIB_PATH} not found: build it first (python -c 'import __graft_entry__ as g; g.build()')")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    got = lib.fod_version()
    if got != ABI_VERSION:
        raise FodError(f"libfod_b200.so ABI {got} != binding ABI {ABI_VERSION}: rebuild")
    _lib = lib
    return lib


def check(rc, what):
    if rc != FOD_OK:
        msg = load().fod_last_error()
        raise FodError(f"{what} failed with status {rc}: {msg.decode() if msg else ''}")

"""ctypes binding of libinr_b200.so -- the only way the Python host code reaches the kernels.

Every entry point of include/inr_b200.h is declared here with its argument types; `check(rc)` turns the C
status codes into Python exceptions carrying inr_last_error().  The library must have been built
(`python -m inr_for_audio_b200.build`); nothing here falls back to PyTorch or the CPU.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libinr_b200.so")

INR_OK, INR_E_BADARG, INR_E_UNSUPPORTED, INR_E_CUDA, INR_E_WORKSPACE = 0, -1, -2, -3, -4
PREC_FP32, PREC_BF16, PREC_FP16, PREC_FP16S = 0, 1, 2, 3
LOSS_MSE, LOSS_MAE = 0, 1
ABI_VERSION = 102                          # INR_ABI_VERSION of include/inr_b200.h
STATIC_SCRATCH_BYTES = 32 * 1024 * 1024   # INR_STATIC_SCRATCH_BYTES
PRECISIONS = {"fp32": PREC_FP32, "bf16": PREC_BF16, "fp16": PREC_FP16, "fp16s": PREC_FP16S}


class SirenDesc(C.Structure):
    _fields_ = [("in_features", C.c_int32), ("hidden_features", C.c_int32), ("hidden_layers", C.c_int32),
                ("out_features", C.c_int32), ("outermost_linear", C.c_int32), ("first_omega_0", C.c_float),
                ("hidden_omega_0", C.c_float), ("precision", C.c_int32), ("num_snake", C.c_int32),
                ("num_tanh", C.c_int32), ("first_linear", C.c_int32)]


KAN_MAX_LAYERS = 8


class KanDesc(C.Structure):
    _fields_ = [("num_layers", C.c_int32), ("widths", C.c_int32 * (KAN_MAX_LAYERS + 1)), ("grid_size", C.c_int32),
                ("spline_order", C.c_int32), ("precision", C.c_int32)]


class OptState(C.Structure):
    _fields_ = [("lr", C.c_float), ("min_lr", C.c_float), ("factor", C.c_float), ("threshold", C.c_float),
                ("sched_eps", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("adam_eps", C.c_float),
                ("patience", C.c_int32), ("num_bad", C.c_int32), ("step", C.c_int32), ("hist_cap", C.c_int32),
                ("best", C.c_float), ("last_loss", C.c_float), ("step_size", C.c_float), ("bc2_sqrt", C.c_float),
                ("gmax_acc", C.c_float), ("gmax_prev", C.c_float), ("omb1", C.c_float), ("omb2", C.c_float),
                ("hist_base", C.c_int32), ("reserved0", C.c_int32),
                ("loss_acc", C.c_double * 4)]


OPT_STATE_BYTES = C.sizeof(OptState)
OPT_LOSS_ACC_OFFSET = OptState.loss_acc.offset

_p = C.c_void_p
_i64 = C.c_int64
_f = C.c_float
_i = C.c_int

# name -> (restype, argtypes); mirrors include/inr_b200.h (+ the inr_dbg_* test hooks of inr_b200_debug.h) one to one
SIGNATURES = {
    "inr_last_error": (C.c_char_p, []),
    "inr_version": (_i, []),
    "inr_launch_count": (_i64, []),
    "inr_set_sm_limit": (_i, [_i]),
    "inr_profile_begin": (_i, []),
    "inr_profile_end": (_i64, [C.c_char_p, _i64]),
    "inr_siren_num_params": (_i64, [C.POINTER(SirenDesc)]),
    "inr_siren_workspace_bytes": (_i64, [C.POINTER(SirenDesc), _i64, _i]),
    "inr_coords_linspace": (_i, [_p, _i64, _i64, _i64, _f, _p]),
    "inr_coords_grid2d": (_i, [_p, _i64, _i64, _i64, _i64, _p]),
    "inr_encode_gaussian": (_i, [_p, _i, _i64, _i64, _i64, _f, _p, _i, _p, _p]),
    "inr_encode_nerf": (_i, [_p, _i, _i64, _i64, _i64, _f, _i, _f, _p, _p]),
    "inr_siren_forward": (_i, [C.POINTER(SirenDesc), _p, _p, _i64, _i64, _i64, _f, _p, _p, _i64, _p]),
    "inr_siren_forward_static": (_i, [C.POINTER(SirenDesc), _p, _p, _i64, _i64, _i64, _f, _p, _p]),
    "inr_siren_loss_backward": (_i, [C.POINTER(SirenDesc), _p, _p, _p, _i64, _i64, _i64, _f, _i, _f, _p, _p, _p, _p,
                                     _p, _i64, _p]),
    "inr_siren_train_forward": (_i, [C.POINTER(SirenDesc), _p, _p, _i64, _i64, _i64, _f, _p, _p, _i64, _p]),
    "inr_siren_train_backward": (_i, [C.POINTER(SirenDesc), _p, _p, _p, _i64, _i64, _i64, _f, _i, _f, _p, _p, _p, _p,
                                      _i64, _p]),
    "inr_opt_loss_to_f32": (_i, [_p, _p, _p]),
    "inr_opt_loss_from_f32": (_i, [_p, _p, _f, _p]),
    "inr_stmdct_num_frames": (_i64, [_i64, _i]),
    "inr_stmdct_forward": (_i, [_p, _i64, _i, _p, _p, _p]),
    "inr_istmdct": (_i, [_p, _i64, _i, _p, _p, _p]),
    "inr_mdct_loss": (_i, [_p, _p, _i64, _i, _p, _i, _f, _p, _p, _p]),
    "inr_stft_num_frames": (_i64, [_i64, _i]),
    "inr_stft_mag": (_i, [_p, _i64, _i, _i, _p, _f, _p, _p, _p]),
    "inr_stft_loss": (_i, [_p, _p, C.c_double, _i64, _i, _i, _p, _f, _f, _p, _p, _p]),
    "inr_snr_loss": (_i, [_p, _p, _i64, _f, _f, _p, _p, _p, _p]),
    "inr_opt_init": (_i, [_p, _f, _f, C.c_int32, _p]),
    "inr_adam_step": (_i, [_p, _p, _p, _p, _i64, _p, _f, _p, _p]),
    "inr_siren_fit_step": (_i, [C.POINTER(SirenDesc), _p, _p, _p, _p, _p, _p, _i64, _i64, _i64, _f, _i, _f, _p, _p,
                                _p, _i64, _p]),
    "inr_kan_num_params": (_i64, [C.POINTER(KanDesc)]),
    "inr_kan_num_grid_floats": (_i64, [C.POINTER(KanDesc)]),
    "inr_kan_workspace_bytes": (_i64, [C.POINTER(KanDesc), _i64, _i]),
    "inr_kan_b_splines": (_i, [_p, _p, _i64, _i, _i, _i, _p, _p]),
    "inr_kan_forward": (_i, [C.POINTER(KanDesc), _p, _p, _p, _i64, _i64, _i64, _f, _p, _p, _i64, _p]),
    "inr_kan_loss_backward": (_i, [C.POINTER(KanDesc), _p, _p, _p, _p, _i64, _i64, _i64, _f, _i, _f, _p, _p, _p, _p, _p,
                                   _i64, _p]),
    "inr_dbg_fused_trace_enable": (_i, [_i]),
    "inr_dbg_fused_trace_read": (_i, [_p, _i]),
    "inr_dbg_tc_scratch_bytes": (_i64, [_i, _i64]),
    "inr_dbg_tc_layer": (_i, [_i, _i, _p, _p, _p, _f, _p, _p, _p, _i64, _i, _p, _i64, _p]),
    "inr_dbg_tc_dw": (_i, [_i, _p, _p, _p, _i64, _i, _p, _i64, _p]),
}

_lib = None


class InrLibraryError(RuntimeError):
    pass


def lib():
    """The loaded library (built on demand is NOT attempted here: a missing .so is a loud error)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise InrLibraryError(
                "libinr_b200.so is not built (%s). Run `python -m inr_for_audio_b200.build`; there is no CPU or "
                "PyTorch fallback for the INR hot path." % LIB_PATH)
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        got = l.inr_version()
        if got != ABI_VERSION:
            raise InrLibraryError("libinr_b200.so reports ABI version %d, this binding was written against %d: rebuild "
                                  "with `python -m inr_for_audio_b200.build`" % (got, ABI_VERSION))
        _lib = l
    return _lib


def check(rc):
    if rc is not None and rc < 0:
        msg = lib().inr_last_error()
        raise InrLibraryError("libinr_b200 error %d: %s" % (rc, msg.decode() if msg else "?"))
    return rc


def ptr(t):
    """Raw device pointer of a torch tensor (or None)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def profile_begin():
    check(lib().inr_profile_begin())


def profile_end():
    """-> {kernel tag: (launches, total_ms)} for the launches since profile_begin()."""
    buf = C.create_string_buffer(1 << 16)
    check(lib().inr_profile_end(buf, len(buf)))
    out = {}
    for line in buf.value.decode().splitlines():
        tag, n, ms = line.rsplit(",", 2)
        out[tag] = (int(n), float(ms))
    return out

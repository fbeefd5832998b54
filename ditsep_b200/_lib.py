"""ctypes binding of libditsep_b200.so (the C ABI declared in include/ditsep_b200.h).

There is deliberately no fallback: if the shared library is missing or a call fails,
an exception is raised.  The library is built in-tree by ``ditsep_b200.build``.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libditsep_b200.so")

c_void_p, c_int, c_int64, c_uint64, c_float = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_float


class SdeParams(C.Structure):
    """ditsep_sde_params (include/ditsep_b200.h)."""

    _fields_ = [("sde", C.c_int32), ("n_src", C.c_int32), ("p0", C.c_float), ("sigma_min", C.c_float),
                ("sigma_max", C.c_float), ("N", C.c_int32), ("snr", C.c_float), ("elem_scale", C.c_void_p)]


class GnJob(C.Structure):
    """ditsep_gn_job (include/ditsep_b200.h): one GroupNorm consumer of up to 4 fixed-point statistics sources."""

    _fields_ = [("src_sums", C.c_void_p * 4), ("src_C", C.c_int32 * 4), ("n_src", C.c_int32), ("C", C.c_int32),
                ("G", C.c_int32), ("eps", C.c_float), ("HW", C.c_int64), ("gamma", C.c_void_p), ("beta", C.c_void_p),
                ("scale_shift", C.c_void_p)]


# name -> (restype, argtypes); must list EVERY symbol of include/ditsep_b200.h
SIGNATURES = {
    "ditsep_abi_version": (c_int, []),
    "ditsep_last_error": (C.c_char_p, []),
    "ditsep_launch_count": (c_int64, []),
    "ditsep_note_launches": (None, [c_int64]),
    "ditsep_pc_corrector": (c_int, [C.POINTER(SdeParams), c_void_p, c_void_p, c_void_p, c_uint64, c_uint64, c_void_p,
                                    c_void_p, c_int, c_int64, c_void_p, c_void_p]),
    "ditsep_pc_predictor": (c_int, [C.POINTER(SdeParams), c_void_p, c_void_p, c_void_p, c_uint64, c_uint64, c_void_p,
                                    c_void_p, c_void_p, c_int, c_int64, c_void_p, c_void_p]),
    "ditsep_pc_prior": (c_int, [C.POINTER(SdeParams), c_void_p, c_void_p, c_uint64, c_uint64, c_void_p, c_int, c_int64,
                                c_void_p, c_void_p]),
    "ditsep_pc_langevin": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_void_p,
                                   c_void_p]),
    "ditsep_energy_envelope": (c_int, [c_void_p, c_void_p, c_int, c_int64, c_int, c_void_p]),
    "ditsep_pc_perturb": (c_int, [C.POINTER(SdeParams), c_void_p, c_void_p, c_uint64, c_uint64, c_void_p, c_void_p, c_void_p,
                                  c_int, c_int64, c_void_p, c_void_p]),
    "ditsep_score_loss": (c_int, [C.POINTER(SdeParams), c_void_p, c_void_p, c_uint64, c_uint64, c_void_p, c_void_p, c_void_p,
                                  c_int, c_int64, c_void_p]),
    "ditsep_score_loss_sigma": (c_int, [c_void_p, c_void_p, c_uint64, c_uint64, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                        c_int64, c_void_p]),
    "ditsep_sb_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_uint64, c_uint64, c_void_p, c_void_p,
                               c_void_p, c_void_p, c_void_p, c_int, c_int, c_int64, c_void_p, c_void_p]),
    "ditsep_conv1d_tc": (c_int, [c_void_p, c_int, c_int, c_int, c_int64, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                                 c_void_p, c_void_p, c_int, c_int64, c_void_p, c_void_p, c_int, c_int, c_int64, c_void_p]),
    "ditsep_conv1d_in": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "ditsep_conv1d_out": (c_int, [c_void_p, c_int, c_int64, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "ditsep_vae_sample": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "ditsep_philox_normal": (c_int, [c_void_p, c_int64, c_uint64, c_uint64, c_void_p]),
    "ditsep_box_muller": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "ditsep_stft_compress": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_int,
                                     c_float, c_float, c_int, c_void_p]),
    "ditsep_stft_compress_cached": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_int,
                                     c_float, c_float, c_int, c_void_p, c_int, c_void_p]),
    "ditsep_expand_istft": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int64, c_int, c_float, c_float,
                                    c_void_p]),
    "ditsep_gn_sums": (c_int, [c_void_p, c_int, c_int, c_int64, c_int, c_int, c_void_p, c_void_p]),
    "ditsep_gn_finalize": (c_int, [C.POINTER(GnJob), c_int, c_void_p]),
    "ditsep_gn_stats_static": (c_int, [c_void_p, c_int, c_int, c_int64, c_int, c_int, c_int, c_void_p, c_void_p, c_float,
                                       c_void_p, c_void_p]),
    "ditsep_gn_small": (c_int, [c_void_p, c_int, c_int, c_int64, c_int, c_int, c_int, c_void_p, c_void_p, c_float, c_int,
                                c_void_p, c_int, c_void_p]),
    "ditsep_gn_apply_job": (c_int, [c_void_p, c_int, c_int, c_int64, C.POINTER(GnJob), c_int, c_int, c_void_p, c_int, c_void_p]),
    "ditsep_gn_apply": (c_int, [c_void_p, c_int, c_int, c_int64, c_int, c_int, c_void_p, c_int, c_void_p, c_int,
                                c_void_p]),
    "ditsep_fir_resample": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                    c_int, c_void_p, c_int, c_void_p]),
    "ditsep_conv_direct": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                   c_int, c_void_p, c_int, c_float, c_int, c_int, c_void_p, c_int, c_int, c_void_p]),
    "ditsep_split_hilo": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_int, c_void_p]),
    "ditsep_conv_tc_plan_create": (c_int, [C.POINTER(c_void_p), c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p,
                                           c_int, c_int]),
    "ditsep_conv_tc_plan_create_sc": (c_int, [C.POINTER(c_void_p), c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p,
                                              c_int, c_int, c_void_p, c_int, c_int]),
    "ditsep_debug_conv_trace": (None, [c_void_p]),
    "ditsep_conv_tc_plan_destroy": (None, [c_void_p]),
    "ditsep_conv_tc_has_stats": (c_int, [c_void_p]),
    "ditsep_conv_tc_run": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_float, c_void_p,
                                   c_int, c_int, c_void_p, c_void_p]),
    "ditsep_attention": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_int,
                                 c_void_p]),
    "ditsep_temb_mlp": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                c_void_p]),
    "ditsep_temb_dense": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "ditsep_output_layer": (c_int, [c_void_p, c_int, c_int, c_int64, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p,
                                    c_int, c_void_p, c_void_p]),
}

_lib = None


class DitsepError(RuntimeError):
    pass


def load():
    """Load the shared library (once) and declare every signature.  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise DitsepError(f"{LIB_PATH} not found: build it with `python -m ditsep_b200.build` "
                          "(there is no CPU or PyTorch fallback for this path)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if a declared symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str):
    if rc != 0:
        msg = load().ditsep_last_error().decode(errors="replace")
        raise DitsepError(f"{what} failed (rc={rc}): {msg}")

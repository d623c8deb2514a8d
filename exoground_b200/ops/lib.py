"""ctypes binding of libexg.so (include/exg.h).  No torch types cross this boundary.

The library is built in-tree by exoground_b200/build.py.  Loading fails loudly: there is no CPU or
PyTorch fallback for any op in this package.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(HERE, "libexg.so")

F32, BF16 = 0, 1
EPI_NONE, EPI_QGELU, EPI_RESID, EPI_QGELU_BWD, EPI_RELU, EPI_RELU_BWD = 0, 1, 2, 3, 4, 5
GEMM_AUTO, GEMM_SIMT, GEMM_TC, GEMM_TC2 = 0, 1, 2, 3

vp = C.c_void_p


class RowMap(C.Structure):
    _fields_ = [("group", C.c_long), ("group_stride", C.c_long), ("offset", C.c_long)]


class GemmParams(C.Structure):
    _fields_ = [("A", vp), ("B", vp), ("C", vp),
                ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
                ("lda", C.c_long), ("ldb", C.c_long), ("ldc", C.c_long),
                ("transA", C.c_int), ("transB", C.c_int),
                ("batch", C.c_int), ("strideA", C.c_long), ("strideB", C.c_long), ("strideC", C.c_long),
                ("in_dtype", C.c_int), ("out_dtype", C.c_int), ("epilogue", C.c_int),
                ("bias", vp), ("aux", vp), ("ldaux", C.c_long), ("strideAux", C.c_long),
                ("aux_out", vp), ("alpha", C.c_float), ("impl", C.c_int)]


_SIGS = {
    "exg_version": (C.c_int, []),
    "exg_last_error": (C.c_char_p, []),
    "exg_launch_count": (C.c_longlong, []),
    "exg_reserve_sms": (C.c_int, [C.c_int]),
    "exg_profile_classes": (C.c_int, []),
    "exg_profile_class_name": (C.c_char_p, [C.c_int]),
    "exg_profile_begin": (C.c_int, []),
    "exg_profile_end": (C.c_int, [C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_longlong)]),
    "exg_gemm": (C.c_int, [C.POINTER(GemmParams), vp]),
    "exg_gemm_tc_supported": (C.c_int, [C.POINTER(GemmParams)]),
    "exg_gemm_trace": (C.c_int, [vp, C.c_int, C.c_int]),
    "exg_gemm_add_ln_supported": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_long, C.c_long]),
    "exg_gemm_add_ln": (C.c_int, [vp, vp, vp, vp, vp, vp, C.c_float, vp, vp, C.c_int, C.c_int, C.c_long, C.c_long, vp]),
    "exg_align_finish": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp]),
    "exg_layernorm_fwd": (C.c_int, [vp, C.c_int, vp, vp, vp, C.c_int, RowMap, vp, C.c_int, vp, C.c_long,
                                    vp, vp, C.c_long, C.c_int, C.c_float, vp]),
    "exg_add_layernorm_fwd": (C.c_int, [vp, C.c_int, vp, vp, vp, vp, vp, C.c_int, RowMap, C.c_long, C.c_int,
                                        C.c_float, vp]),
    "exg_layernorm_bwd": (C.c_int, [vp, C.c_int, RowMap, vp, C.c_int, RowMap, vp, C.c_int, vp, vp, vp, vp,
                                    vp, vp, C.c_int, vp, vp, vp, C.c_long, C.c_int, vp]),
    "exg_add_rows": (C.c_int, [vp, RowMap, vp, RowMap, vp, C.c_long, vp, C.c_int, RowMap, vp, C.c_int,
                               C.c_long, C.c_int, vp]),
    "exg_l2norm_fwd": (C.c_int, [vp, C.c_int, RowMap, vp, C.c_int, vp, C.c_long, C.c_int, vp]),
    "exg_l2norm_bwd": (C.c_int, [vp, C.c_int, vp, C.c_int, vp, vp, C.c_int, RowMap, C.c_int, C.c_long,
                                 C.c_int, vp]),
    "exg_colsum": (C.c_int, [vp, C.c_int, C.c_long, C.c_long, C.c_long, vp, C.c_int, vp]),
    "exg_wcolsum": (C.c_int, [vp, C.c_int, vp, C.c_long, C.c_long, C.c_long, vp, C.c_int, vp]),
    "exg_rowdot": (C.c_int, [vp, C.c_int, vp, vp, vp, C.c_long, C.c_int, vp]),
    "exg_outer_rows": (C.c_int, [vp, vp, vp, C.c_int, C.c_long, C.c_int, vp]),
    "exg_cast": (C.c_int, [vp, C.c_int, vp, C.c_int, C.c_long, vp]),
    "exg_stage_mean": (C.c_int, [vp, C.c_int, C.c_long, C.c_int, vp, vp, C.c_int, C.c_long, vp]),
    "exg_bcast_rows": (C.c_int, [vp, C.c_float, vp, C.c_int, C.c_long, C.c_int, C.c_long, vp]),
    "exg_xent_diag_fwd": (C.c_int, [vp, C.c_int, vp, vp, C.c_long, C.c_int, C.c_int, vp]),
    "exg_xent_diag_bwd": (C.c_int, [vp, C.c_int, vp, C.c_float, vp, C.c_int, C.c_long, C.c_int, C.c_int, vp]),
    "exg_xview_nce_fwd": (C.c_int, [vp, C.c_int, vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.c_float, vp, vp, vp, vp, vp, vp]),
    "exg_xview_nce_bwd": (C.c_int, [vp, C.c_int, vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.c_float, vp, vp, vp, C.c_int, vp]),
    "exg_ground_loss": (C.c_int, [vp, vp, vp, vp, C.c_long, C.c_int, C.c_float, C.POINTER(C.c_float), C.c_int,
                                  vp, vp, vp, vp]),
    "exg_attn_fwd": (C.c_int, [vp, vp, vp, C.c_int, C.c_long, C.c_long, C.c_long, vp, vp, C.c_long, vp,
                               C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, vp]),
    "exg_attn_bwd": (C.c_int, [vp, vp, vp, vp, vp, C.c_int, C.c_long, C.c_long, C.c_long, C.c_long, vp, vp,
                               vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, vp]),
    "exg_attn_bwd_workspace": (C.c_long, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "exg_nce_pack_targets": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_int, vp]),
    "exg_nce_fwd": (C.c_int, [vp, C.c_int, vp, vp, vp, vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int,
                              C.c_int, C.c_int, C.c_float, vp]),
    "exg_nce_bwd": (C.c_int, [vp, C.c_int, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int,
                              C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, vp]),
    "exg_own_max": (C.c_int, [vp, C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp]),
    "exg_nce_logdiff": (C.c_int, [vp, vp, C.c_int, vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp]),
    "exg_tan_loss_tail_workspace": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "exg_tan_loss_tail": (C.c_int, [vp] * 10 + [C.c_int] * 5 + [C.c_float, C.c_float, C.c_int, C.c_float] +
                          [vp] * 7 + [C.c_size_t, vp]),
    "exg_tan_loss_tail_static_max_bytes": (C.c_size_t, []),
    "exg_tan_loss_tail_static": (C.c_int, [vp] * 10 + [C.c_int] * 5 + [C.c_float, C.c_float, C.c_int, C.c_float] +
                                 [vp] * 7),
    "exg_adamw": (C.c_int, [vp, vp, vp, vp, vp, C.c_long, C.c_float, C.c_float, C.c_float, C.c_float,
                            C.c_float, C.c_int, C.c_float, vp]),
    "exg_adamw_dev": (C.c_int, [vp, vp, vp, vp, vp, C.c_long, vp, C.c_float, C.c_float, C.c_float, C.c_float,
                                C.c_float, vp]),
    "exg_grad_sumsq": (C.c_int, [vp, vp, C.c_long, vp, C.c_int, vp]),
    "exg_grad_clip_scale": (C.c_int, [vp, vp, C.c_long, vp, C.c_float, C.c_float, vp]),
    "exg_adamw_clip": (C.c_int, [vp, vp, vp, vp, vp, C.c_long, vp, C.c_float, C.c_float, C.c_float, C.c_float,
                                 C.c_float, C.c_int, C.c_float, vp, vp, C.c_float, vp]),
    "exg_ema": (C.c_int, [vp, vp, C.c_long, C.c_float, vp]),
}

_lib = None


class ExgError(RuntimeError):
    pass


def exported_symbols():
    """Names include/exg.h declares (used by the CPU test that checks the library exports them all)."""
    return sorted(_SIGS)


def load():
    """dlopen libexg.so and attach signatures.  Raises if the library is missing (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ExgError(f"{LIB_PATH} not found: build it with `python -m exoground_b200.build` "
                       "(exoground_b200 has no CPU / PyTorch fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _SIGS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().exg_last_error()
        raise ExgError(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")


def launch_count():
    return int(load().exg_launch_count())


def profile_begin():
    check(load().exg_profile_begin(), "exg_profile_begin")


def profile_end():
    """{class name: {"ms", "work", "launches"}} for the launches since profile_begin (synchronises)."""
    lib = load()
    n = lib.exg_profile_classes()
    ms, work, cnt = (C.c_double * n)(), (C.c_double * n)(), (C.c_longlong * n)()
    check(lib.exg_profile_end(ms, work, cnt), "exg_profile_end")
    return {lib.exg_profile_class_name(i).decode(): {"ms": ms[i], "work": work[i], "launches": int(cnt[i])}
            for i in range(n) if cnt[i]}

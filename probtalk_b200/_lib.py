"""ctypes binding of ``libprobtalk_b200.so`` (the C ABI declared in ``include/probtalk_b200.h``).

The product path has NO fallback: if the CUDA library is missing or fails to load, importing a
compute op raises immediately.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# PROBTALK_B200_LIB: developer override (A/B measurements of two builds inside one GPU call)
LIB_PATH = os.environ.get("PROBTALK_B200_LIB") or os.path.join(_HERE, "lib", "libprobtalk_b200.so")

_lib = None


class PtError(RuntimeError):
    pass


class GemmDesc(C.Structure):
    """Mirror of ``pt_gemm_desc`` (include/probtalk_b200.h)."""
    _fields_ = [
        ("a", C.c_void_p), ("a_rows", C.c_longlong), ("a_pitch", C.c_longlong), ("w", C.c_void_p),
        ("M", C.c_int), ("N", C.c_int), ("C", C.c_int), ("taps", C.c_int), ("s", C.c_int),
        ("base_row", C.c_int), ("rows_per_sample", C.c_int), ("t_valid", C.c_int), ("n_valid", C.c_int),
        ("out", C.c_void_p), ("oss", C.c_longlong), ("ors", C.c_longlong), ("ocs", C.c_longlong),
        ("out_f32", C.c_int), ("bias", C.c_void_p), ("residual", C.c_void_p),
        ("rss", C.c_longlong), ("rrs", C.c_longlong), ("res_f32", C.c_int), ("act", C.c_int),
        ("slope", C.c_float),
        ("row_norm", C.c_int), ("norm_a", C.c_void_p), ("norm_b", C.c_void_p), ("labels", C.c_void_p),
        ("label_rows", C.c_int),
        ("norm_classes", C.c_int),
        ("split", C.c_int), ("a_lo", C.c_void_p), ("out_lo", C.c_void_p), ("residual_lo", C.c_void_p),
        ("a_col_table", C.c_void_p),
    ]


# symbol -> (restype, argtypes); tests/test_abi.py checks this table against include/probtalk_b200.h
SIGNATURES = {
    "pt_version": (C.c_int, []),
    "pt_last_error": (C.c_char_p, []),
    "pt_pq_search": (C.c_int, [C.c_void_p, C.c_longlong, C.c_longlong, C.c_longlong, C.c_int, C.c_int, C.c_int,
                               C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                               C.c_void_p]),
    "pt_pq_gather": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_longlong, C.c_longlong,
                               C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "pt_gemm_bf16": (C.c_int, [C.POINTER(GemmDesc), C.c_void_p]),
    "pt_attention": (C.c_int, [C.c_void_p, C.c_longlong, C.c_void_p, C.c_longlong, C.c_void_p, C.c_longlong,
                               C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float,
                               C.c_void_p]),
    "pt_pq_ema_workspace_bytes": (C.c_longlong, [C.c_int, C.c_int, C.c_int]),
    "pt_pq_ema_update": (C.c_int, [C.c_void_p, C.c_longlong, C.c_longlong, C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_int,
                                   C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_int,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p]),
    "pt_pq_gather_conv1x1": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong,
                                       C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "pt_attention_f32": (C.c_int, [C.c_void_p, C.c_longlong, C.c_void_p, C.c_longlong, C.c_void_p, C.c_longlong,
                                   C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                   C.c_float, C.c_void_p]),
    "pt_pack_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                               C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "pt_pack_mask": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                               C.c_void_p]),
    "pt_feature_stats_workspace_bytes": (C.c_longlong, [C.c_int]),
    "pt_feature_stats": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong,
                                  C.c_void_p]),
    "pt_feature_stats_static": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "pt_lvd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "pt_diversity": (C.c_int, [C.c_void_p, C.c_int, C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p]),
    "pt_pose_postprocess": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                     C.c_int, C.c_void_p]),
    "pt_interp_blend": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                  C.c_int, C.c_void_p]),
    "pt_layernorm512": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "pt_ln_mod_silu512": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                    C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "pt_embed_tokens512": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "pt_maskgit_step": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                  C.c_int, C.c_int, C.c_int, C.c_float, C.c_void_p]),
    "pt_w2v_frontend_workspace_bytes": (C.c_longlong, [C.c_int]),
    "pt_w2v_frontend": (C.c_int, [C.c_void_p, C.c_int, C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_longlong, C.c_void_p, C.c_longlong, C.c_void_p]),
    "pt_w2v_interp": (C.c_int, [C.c_void_p, C.c_int, C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                C.c_void_p]),
    "pt_layernorm_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_int,
                                    C.c_int, C.c_void_p]),
}


def lib():
    """Load (once) and return the CUDA library; raises PtError if it is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PtError(
                "probtalk_b200 CUDA library not found at %s -- run `python -m probtalk_b200.build` "
                "(there is no CPU fallback)" % LIB_PATH)
        try:
            handle = C.CDLL(LIB_PATH)
        except OSError as e:   # pragma: no cover
            raise PtError("cannot load %s: %s" % (LIB_PATH, e))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = lib().pt_last_error()
        raise PtError("%s failed: %s" % (what or "probtalk_b200 call", msg.decode() if msg else "unknown error"))

"""ctypes binding of libmercury_b200.so (the C ABI declared in include/mercury_b200.h).

There is no CPU fallback: if the library is missing or fails to load, importing the ops raises.
"""
from __future__ import annotations

import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "libmercury_b200.so")

ABI_VERSION = 3

OK, ERR_ARG, ERR_CUDA, ERR_WORKSPACE = 0, -1, -2, -3
F32, BF16 = 0, 1
CDF_EXACT_SERIAL_F32, CDF_F64 = 0, 1
ST_BAD_LABEL, ST_NEGATIVE, ST_NONFINITE, ST_ZERO_SUM, ST_TOPK_TIES, ST_PEER_TIMEOUT, ST_INTERNAL, ST_GROUP_COUNT = 1, 2, 4, 8, 16, 32, 64, 128
PEER_BOX_BYTES, PEER_HANDLE_BYTES = 4096, 64
STAT_MEAN, STAT_EMA, STAT_SHIFT, STAT_TOTAL, STAT_CHAIN_TOTAL, STAT_COUNT = 0, 1, 2, 3, 4, 8
STATIC_MAX_N = 1 << 22

c_i64, c_f64, c_int, c_vp, c_sz = ctypes.c_int64, ctypes.c_double, ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t


class UpdateArgs(ctypes.Structure):
    """struct mercury_update_args (include/mercury_b200.h)."""

    _fields_ = [
        ("logits", c_vp), ("dtype", c_int), ("labels", c_vp), ("n", c_i64), ("c", c_i64), ("row_stride", c_i64),
        ("ignore_index", c_i64),
        ("scores_in", c_vp),
        ("h_bounds", ctypes.POINTER(c_i64)), ("n_bounds", c_int),
        ("ema_state", c_vp), ("ema_first", c_int), ("ema_alpha", c_f64), ("smooth_alpha", c_f64),
        ("uniforms", c_vp), ("b", c_i64),
        ("data", c_vp), ("row_bytes", c_i64), ("index", c_vp),
        ("scores", c_vp), ("stats", c_vp), ("idx", c_vp), ("nprob", c_vp), ("seed", c_vp),
        ("out_data", c_vp), ("out_labels", c_vp), ("out_index", c_vp),
        ("status", c_vp),
        ("global_sums", c_vp), ("n_global_sums", c_int), ("n_global", c_i64), ("index_base", c_i64), ("draw_only", c_int),
        ("peer_boxes", c_vp), ("n_shards", c_int), ("shard_rank", c_int), ("sums_all", c_vp), ("totals_all", c_vp),
        ("local_uniforms", c_vp),
        ("cdf_mode", c_int),
    ]


# name -> (restype, argtypes); every symbol include/mercury_b200.h declares
SIGNATURES = {
    "mercury_abi_version": (c_int, []),
    "mercury_last_error": (ctypes.c_char_p, []),
    "mercury_score": (c_int, [c_vp, c_int, c_vp, c_i64, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp]),
    "mercury_pool_stats_workspace_bytes": (c_sz, [c_i64]),
    "mercury_pool_stats": (c_int, [c_vp, c_i64, ctypes.POINTER(c_i64), c_int, c_vp, c_int, c_f64, c_f64, c_vp, c_vp, c_sz, c_vp]),
    "mercury_normalize": (c_int, [c_vp, c_i64, c_vp, c_vp, c_vp]),
    "mercury_cdf_workspace_bytes": (c_sz, [c_i64]),
    "mercury_cdf": (c_int, [c_vp, c_i64, c_int, c_vp, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "mercury_cdf_static": (c_int, [c_vp, c_i64, c_int, c_vp, c_vp, c_vp, c_vp]),
    "mercury_sample_with_replacement": (c_int, [c_vp, c_i64, c_vp, c_i64, c_vp, c_vp]),
    "mercury_sample_topk_workspace_bytes": (c_sz, [c_i64, c_i64]),
    "mercury_sample_topk": (c_int, [c_vp, c_i64, c_vp, c_i64, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "mercury_groupwise_draw_workspace_bytes": (c_sz, [c_i64, c_i64]),
    "mercury_groupwise_draw": (c_int, [c_vp, c_vp, ctypes.c_int32, c_i64, c_i64, c_vp, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "mercury_gather_reweight": (c_int, [c_vp, c_vp, c_i64, c_i64, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "mercury_gather_owned": (c_int, [c_vp, c_i64, c_i64, c_i64, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "mercury_test_set_guess_bias": (None, [c_f64]),
    "mercury_test_set_fused": (None, [c_int]),
    "mercury_debug_step_times": (c_int, [c_vp, c_i64, ctypes.POINTER(ctypes.c_ulonglong)]),
    "mercury_sum_scores": (c_int, [c_vp, c_i64, c_vp, c_vp, c_sz, c_vp]),
    "mercury_shard_select": (c_int, [c_vp, c_int, c_int, c_vp, c_i64, c_vp, c_vp]),
    "mercury_score_head": (c_int, [c_vp, c_int, c_i64, c_i64, c_i64, c_vp, c_vp, c_i64, c_vp, c_i64, c_vp, c_vp, c_vp]),
    "mercury_score_head_wide": (c_int, [c_vp, c_i64, c_i64, c_i64, c_vp, c_i64, c_vp, c_i64, c_vp, c_i64, c_vp, c_vp, c_vp]),
    "mercury_reweighted_mean": (c_int, [c_vp, c_vp, c_i64, c_vp, c_vp]),
    "mercury_peer_create": (c_int, [c_vp, c_vp]),
    "mercury_peer_open": (c_int, [c_vp, c_vp]),
    "mercury_peer_close": (c_int, [c_vp]),
    "mercury_peer_destroy": (c_int, [c_vp]),
    "mercury_peer_set_wait": (c_int, [c_vp, c_i64, c_vp]),
    "mercury_peer_reset": (c_int, [c_vp, c_vp]),
    "mercury_sum_scores_exchange": (c_int, [c_vp, c_i64, c_vp, c_int, c_int, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "mercury_shard_select_exchange": (c_int, [c_vp, c_vp, c_int, c_int, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp]),
    "mercury_update_samples_workspace_bytes": (c_sz, [c_i64, c_i64]),
    "mercury_update_samples": (c_int, [ctypes.POINTER(UpdateArgs), c_vp, c_sz, c_vp]),
}

_lib = None


def load() -> ctypes.CDLL:
    """Load the CUDA library.  Raises (never falls back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise ImportError(
            f"{SO_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  mercury_b200 has no CPU or PyTorch fallback.")
    lib = ctypes.CDLL(SO_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    got = lib.mercury_abi_version()
    if got != ABI_VERSION:
        raise ImportError(f"libmercury_b200.so has ABI {got}, binding expects {ABI_VERSION}")
    _lib = lib
    return lib


def last_error() -> str:
    return load().mercury_last_error().decode("utf-8", "replace")


def check(rc: int) -> None:
    if rc != OK:
        msg = last_error()
        if rc == ERR_ARG and ("2^24" in msg or "n_sample" in msg):
            raise RuntimeError(msg)  # the texts torch.multinomial raises with
        raise RuntimeError(f"mercury_b200 error {rc}: {msg}")

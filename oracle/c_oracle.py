"""ctypes view of oracle/liboracle_c.so (plain-C oracle).  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle_c.so")
_lib = None

STATUS = {0: "ok", 1: "negative", 2: "nonfinite", 3: "zero_sum", 4: "bad_label"}


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "mercury_oracle_c.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "liboracle_c.so"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
    return _lib


def _p(a, t):
    return a.ctypes.data_as(ctypes.POINTER(t))


def score(logits: np.ndarray, labels: np.ndarray, ignore_index: int = -100, bf16: bool = False):
    n, c = logits.shape
    out = np.empty(n, dtype=np.float32)
    labels = np.ascontiguousarray(labels, dtype=np.int64)
    if bf16:
        x = np.ascontiguousarray(logits).view(np.uint16)
        st = lib().orc_score_bf16(_p(x, ctypes.c_uint16), _p(labels, ctypes.c_int64), ctypes.c_int64(n),
                                  ctypes.c_int64(c), ctypes.c_int64(c), ctypes.c_int64(ignore_index),
                                  _p(out, ctypes.c_float))
    else:
        x = np.ascontiguousarray(logits, dtype=np.float32)
        st = lib().orc_score_f32(_p(x, ctypes.c_float), _p(labels, ctypes.c_int64), ctypes.c_int64(n),
                                 ctypes.c_int64(c), ctypes.c_int64(c), ctypes.c_int64(ignore_index),
                                 _p(out, ctypes.c_float))
    return st, out


def serial_cdf(p: np.ndarray):
    p = np.ascontiguousarray(p, dtype=np.float32)
    chain = np.empty_like(p)
    cdf = np.empty_like(p)
    st = lib().orc_serial_cdf_f32(_p(p, ctypes.c_float), ctypes.c_int64(p.size), _p(chain, ctypes.c_float),
                                  _p(cdf, ctypes.c_float))
    return st, chain, cdf


def lower_bound(cdf: np.ndarray, u: np.ndarray):
    cdf = np.ascontiguousarray(cdf, dtype=np.float32)
    u = np.ascontiguousarray(u, dtype=np.float64)
    idx = np.empty(u.size, dtype=np.int64)
    lib().orc_lower_bound(_p(cdf, ctypes.c_float), ctypes.c_int64(cdf.size), _p(u, ctypes.c_double),
                          ctypes.c_int64(u.size), _p(idx, ctypes.c_int64))
    return idx


def exp_race_topk(p: np.ndarray, u: np.ndarray, k: int):
    p = np.ascontiguousarray(p, dtype=np.float32)
    u = np.ascontiguousarray(u, dtype=np.float64)
    keys = np.empty_like(p)
    lib().orc_exp_race_keys(_p(p, ctypes.c_float), _p(u, ctypes.c_double), ctypes.c_int64(p.size),
                            _p(keys, ctypes.c_float))
    idx = np.empty(k, dtype=np.int64)
    lib().orc_topk_desc(_p(keys, ctypes.c_float), ctypes.c_int64(p.size), ctypes.c_int64(k), _p(idx, ctypes.c_int64))
    return keys, idx


def gather_weights(p: np.ndarray, idx: np.ndarray):
    p = np.ascontiguousarray(p, dtype=np.float32)
    idx = np.ascontiguousarray(idx, dtype=np.int64)
    nprob = np.empty(idx.size, dtype=np.float32)
    seed = np.empty(idx.size, dtype=np.float32)
    lib().orc_gather_weights(_p(p, ctypes.c_float), ctypes.c_int64(p.size), _p(idx, ctypes.c_int64),
                             ctypes.c_int64(idx.size), _p(nprob, ctypes.c_float), _p(seed, ctypes.c_float))
    return nprob, seed


def score_head(features: np.ndarray, weight: np.ndarray, bias, labels: np.ndarray, ignore_index: int = -100):
    n, d = features.shape
    c = weight.shape[0]
    f = np.ascontiguousarray(features, dtype=np.float32)
    w = np.ascontiguousarray(weight, dtype=np.float32)
    bb = None if bias is None else np.ascontiguousarray(bias, dtype=np.float32)
    labels = np.ascontiguousarray(labels, dtype=np.int64)
    out = np.empty(n, dtype=np.float32)
    st = lib().orc_score_head_f32(_p(f, ctypes.c_float), _p(w, ctypes.c_float), None if bb is None else _p(bb, ctypes.c_float),
                                  _p(labels, ctypes.c_int64), ctypes.c_int64(n), ctypes.c_int64(d), ctypes.c_int64(c),
                                  ctypes.c_int64(ignore_index), _p(out, ctypes.c_float))
    return st, out


def two_level_draw(p_shards, u: np.ndarray):
    p = np.ascontiguousarray(np.concatenate(p_shards), dtype=np.float32)
    offsets = np.ascontiguousarray(np.concatenate([[0], np.cumsum([len(s) for s in p_shards])]), dtype=np.int64)
    u = np.ascontiguousarray(u, dtype=np.float64)
    idx = np.empty(len(u), dtype=np.int64)
    st = lib().orc_two_level_draw(_p(p, ctypes.c_float), _p(offsets, ctypes.c_int64), ctypes.c_int64(len(p_shards)),
                                  _p(u, ctypes.c_double), ctypes.c_int64(len(u)), _p(idx, ctypes.c_int64))
    return st, idx

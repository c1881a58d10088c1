"""ctypes wrapper for oracle/phylo_oracle.c — test infrastructure / CPU baseline only."""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libphylo_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    if force or not os.path.exists(_LIB_PATH) or any(
            os.path.getmtime(_LIB_PATH) < os.path.getmtime(os.path.join(_HERE, src))
            for src in ("phylo_oracle.c", "phylo_oracle_dense.c")):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B"])
    return _LIB_PATH


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        lib = ctypes.CDLL(_LIB_PATH)
        dp = ctypes.POINTER(ctypes.c_double)
        lib.oracle_loglik_grad.restype = ctypes.c_int
        lib.oracle_loglik_grad.argtypes = [
            ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int32),
            ctypes.POINTER(ctypes.c_uint8), dp, dp, dp, dp, dp, dp, dp, dp, dp, dp, ctypes.c_int]
        lib.oracle_max_threads.restype = ctypes.c_int
        lib.oracle_dense_loglik_grad.restype = ctypes.c_int
        lib.oracle_dense_loglik_grad.argtypes = [
            ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int32),
            ctypes.POINTER(ctypes.c_uint8), dp, dp, dp, dp, dp, dp, dp, dp, dp, dp, ctypes.c_int]
        _lib = lib
    return _lib


def max_threads() -> int:
    return int(_load().oracle_max_threads())


def _p(a, ct):
    return a.ctypes.data_as(ctypes.POINTER(ct))


def loglik_and_gradient(triples, masks, weights, bl, evec, ivec, evals, freqs, rates,
                        cat_weights, n_threads: int = 0, want_grad: bool = True):
    """Same contract as oracle.loglik_and_gradient but takes uint8 masks [N, P]."""
    lib = _load()
    triples = np.ascontiguousarray(triples, dtype=np.int32)
    masks = np.ascontiguousarray(masks, dtype=np.uint8)
    f = lambda a: np.ascontiguousarray(a, dtype=np.float64)
    weights, bl, evec, ivec, evals, freqs, rates, cat_weights = map(
        f, (weights, bl, evec, ivec, evals, freqs, rates, cat_weights))
    n, p = masks.shape
    logl = ctypes.c_double(0.0)
    grad = np.zeros(2 * n - 2) if want_grad else None
    d = ctypes.c_double
    rc = lib.oracle_loglik_grad(
        n, p, len(rates), _p(triples, ctypes.c_int32), _p(masks, ctypes.c_uint8),
        _p(weights, d), _p(bl, d), _p(evec, d), _p(ivec, d), _p(evals, d), _p(freqs, d),
        _p(rates, d), _p(cat_weights, d), ctypes.byref(logl),
        _p(grad, d) if want_grad else None, int(n_threads))
    if rc != 0:
        raise RuntimeError(f"oracle_loglik_grad failed: {rc}")
    return float(logl.value), grad


def dense_loglik_and_gradient(triples, codes, weights, bl, evec, ivec, evals, freqs, rates,
                              cat_weights, n_threads: int = 0, want_grad: bool = True):
    """Any number of states (oracle/phylo_oracle_dense.c): ``codes`` uint8 [N, P] are state codes
    (>= n_states: missing), the eigen system is [S, S] / [S]."""
    lib = _load()
    triples = np.ascontiguousarray(triples, dtype=np.int32)
    codes = np.ascontiguousarray(codes, dtype=np.uint8)
    f = lambda a: np.ascontiguousarray(a, dtype=np.float64)
    weights, bl, evec, ivec, evals, freqs, rates, cat_weights = map(
        f, (weights, bl, evec, ivec, evals, freqs, rates, cat_weights))
    n, p = codes.shape
    s = len(evals)
    assert evec.shape == (s, s) and ivec.shape == (s, s) and freqs.shape == (s,)
    logl = ctypes.c_double(0.0)
    grad = np.zeros(2 * n - 2) if want_grad else None
    d = ctypes.c_double
    rc = lib.oracle_dense_loglik_grad(
        s, n, p, len(rates), _p(triples, ctypes.c_int32), _p(codes, ctypes.c_uint8),
        _p(weights, d), _p(bl, d), _p(evec, d), _p(ivec, d), _p(evals, d), _p(freqs, d),
        _p(rates, d), _p(cat_weights, d), ctypes.byref(logl),
        _p(grad, d) if want_grad else None, int(n_threads))
    if rc != 0:
        raise RuntimeError(f"oracle_dense_loglik_grad failed: {rc}")
    return float(logl.value), grad

"""ctypes binding of ``libb200phylo.so`` (the C ABI in ``include/b200phylo.h``).

This is the only compute path of the package: there is no CPU fallback.  If the shared
library is missing or no B200 is present, construction raises ``EngineError``.

The role this object plays in the reference is the one ``bito``'s instance plays
under torchtree-bito's ``TreeLikelihoodModel`` (selected with ``--engine bitorch``,
/root/reference/modules/macro_flu.nf:119-122): set tree / tip data / model once, then
ask for log-likelihood and branch gradients for new branch lengths.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional, Tuple

import numpy as np

__all__ = ["Engine", "EngineError", "library_path", "load_library", "EXPORTED_SYMBOLS"]

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_NAME = "libb200phylo.so"

EXPORTED_SYMBOLS = (
    "b2p_create", "b2p_destroy", "b2p_last_error", "b2p_version", "b2p_get_info",
    "b2p_set_topology", "b2p_set_tip_states", "b2p_set_pattern_weights", "b2p_set_eigen",
    "b2p_set_categories", "b2p_loglik", "b2p_loglik_grad", "b2p_loglik_grad_host",
    "b2p_set_profiling", "b2p_last_kernel_ms", "b2p_kernel_ms_stats", "b2p_get_branch_partials",
    "b2p_peer_export", "b2p_peer_attach", "b2p_peer_status", "b2p_set_gradient_operator",
    "b2p_branch_outer_products", "b2p_model_gradients",
)


class EngineError(RuntimeError):
    pass


FLAG_OUTER_PRODUCTS = 1  # B2P_FLAG_OUTER_PRODUCTS (include/b200phylo.h)


class _Config(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in (
        "n_tips", "n_patterns", "n_categories", "n_states", "precision", "device",
        "warps_per_block", "flags")]


class _Info(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in (
        "n_tips", "n_patterns", "n_categories", "n_states", "precision", "warps_per_block",
        "cols_per_tile", "n_tiles", "grid_blocks", "blocks_per_sm", "n_sms", "post_rounds",
        "tree_depth", "regs_per_thread", "tasks_per_round", "pre_rounds", "pool_blocks",
        "smem_bytes", "copy_depth", "reserved0")] + [(n, ctypes.c_int64) for n in (
            "device_bytes", "alg_bytes_loglik", "alg_bytes_grad", "resident_reads", "block_loads",
            "block_stores")]


def library_path() -> str:
    # B2P_LIB: a differently built copy of the same library (debug builds: -DB2P_TRACE_CLOCKS)
    return os.environ.get("B2P_LIB") or os.path.join(_HERE, _LIB_NAME)


_lib = None
_host_raw_fn = None


def _host_raw(lib):
    """``b2p_loglik_grad_host`` through a prototype that takes plain integer addresses (no
    ``ctypes.cast`` objects on the per-evaluation path)."""
    global _host_raw_fn
    if _host_raw_fn is None:
        vp = ctypes.c_void_p
        _host_raw_fn = ctypes.CFUNCTYPE(ctypes.c_int, vp, vp, vp, vp)(("b2p_loglik_grad_host", lib))
    return _host_raw_fn


def load_library():
    """Load the C-ABI library and declare prototypes.  Fails loudly when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    path = library_path()
    if not os.path.exists(path):
        raise EngineError(
            f"{path} not found: build it with `python -m b200phylo.build` "
            "(nvcc, sm_100a).  There is no CPU fallback.")
    lib = ctypes.CDLL(path)
    vp, dp = ctypes.c_void_p, ctypes.POINTER(ctypes.c_double)
    lib.b2p_create.argtypes = [ctypes.POINTER(_Config), ctypes.POINTER(vp)]
    lib.b2p_destroy.argtypes = [vp]
    lib.b2p_last_error.argtypes = [vp]
    lib.b2p_last_error.restype = ctypes.c_char_p
    lib.b2p_version.argtypes = []
    lib.b2p_get_info.argtypes = [vp, ctypes.POINTER(_Info)]
    lib.b2p_set_topology.argtypes = [vp, ctypes.POINTER(ctypes.c_int32)]
    lib.b2p_set_tip_states.argtypes = [vp, ctypes.POINTER(ctypes.c_uint8)]
    lib.b2p_set_pattern_weights.argtypes = [vp, dp]
    lib.b2p_set_eigen.argtypes = [vp, dp, dp, dp, dp]
    lib.b2p_set_categories.argtypes = [vp, dp, dp]
    lib.b2p_loglik.argtypes = [vp, vp, vp, vp]
    lib.b2p_loglik_grad.argtypes = [vp, vp, vp, vp, vp]
    lib.b2p_loglik_grad_host.argtypes = [vp, dp, dp, dp]
    lib.b2p_set_profiling.argtypes = [vp, ctypes.c_int]
    lib.b2p_last_kernel_ms.argtypes = [vp, ctypes.POINTER(ctypes.c_float)]
    lib.b2p_kernel_ms_stats.argtypes = [vp, dp, ctypes.POINTER(ctypes.c_int32)]
    lib.b2p_get_branch_partials.argtypes = [vp, ctypes.c_int32, dp]
    lib.b2p_set_gradient_operator.argtypes = [vp, dp, dp]
    lib.b2p_branch_outer_products.argtypes = [vp, vp, vp, vp, vp, vp]
    lib.b2p_model_gradients.argtypes = [vp, vp, vp, ctypes.c_int32, vp, vp]
    lib.b2p_peer_export.argtypes = [vp, ctypes.c_int32, ctypes.c_char_p]
    lib.b2p_peer_attach.argtypes = [vp, ctypes.c_int32, ctypes.c_int32, ctypes.c_char_p]
    lib.b2p_peer_status.argtypes = [vp, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32)]
    for name in EXPORTED_SYMBOLS:
        if name != "b2p_last_error":
            getattr(lib, name).restype = ctypes.c_int
    _lib = lib
    return lib


def _dptr(a: np.ndarray):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


class Engine:
    """One likelihood instance on one B200.

    Parameters mirror ``b2p_config``; arrays follow ``include/b200phylo.h``:
    ``triples`` int32 [N-1, 3] post-order, ``masks`` uint8 [N, P], ``weights`` [P].
    ``n_states`` 4 selects the nucleotide walk kernel; 5..64 (e.g. 61 codons) the
    dense-state path, where the tip bytes are state codes (>= n_states: missing).
    """

    def __init__(self, n_tips: int, n_patterns: int, n_categories: int = 1,
                 precision: str = "float64", device: int = 0, warps_per_block: int = 0, n_states: int = 4,
                 outer_products: bool = False):
        self._h = ctypes.c_void_p()
        self._lib = load_library()
        self._setup = {}      # what the set_* calls were given: an outer-products twin is set up from it
        self._twin = None
        self._twin_version = -1
        self._version = 0
        prec = {"float64": 64, "float32": 32, 64: 64, 32: 32}.get(precision)
        if prec is None:
            raise ValueError(f"precision must be float64 or float32, got {precision!r}")
        cfg = _Config(int(n_tips), int(n_patterns), int(n_categories), int(n_states), prec, int(device),
                      int(warps_per_block), FLAG_OUTER_PRODUCTS if outer_products else 0)
        self.outer_products = bool(outer_products)
        self.n_states = int(n_states)
        rc = self._lib.b2p_create(ctypes.byref(cfg), ctypes.byref(self._h))
        if rc != 0:
            msg = self._lib.b2p_last_error(None).decode()
            self._h = ctypes.c_void_p()
            raise EngineError(f"b2p_create failed ({rc}): {msg}")
        self.n_tips, self.n_patterns, self.n_categories = int(n_tips), int(n_patterns), int(n_categories)
        self.n_branches = 2 * self.n_tips - 2
        self.precision = "float32" if prec == 32 else "float64"
        self.device = int(device)

    # ------------------------------------------------------------------ plumbing
    def _check(self, rc: int, what: str) -> None:
        if rc != 0:
            raise EngineError(f"{what} failed ({rc}): {self._lib.b2p_last_error(self._h).decode()}")

    def close(self) -> None:
        if getattr(self, "_twin", None) is not None:
            self._twin.close()
            self._twin = None
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.b2p_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self) -> dict:
        out = _Info()
        self._check(self._lib.b2p_get_info(self._h, ctypes.byref(out)), "b2p_get_info")
        return {n: int(getattr(out, n)) for n, _ in _Info._fields_}

    # --------------------------------------------------------------------- setup
    def set_topology(self, triples) -> None:
        t = np.ascontiguousarray(triples, dtype=np.int32)
        if t.shape != (self.n_tips - 1, 3):
            raise ValueError(f"triples must have shape ({self.n_tips - 1}, 3), got {t.shape}")
        self._check(self._lib.b2p_set_topology(
            self._h, t.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))), "b2p_set_topology")
        self._remember("topology", t)

    def set_tip_states(self, masks) -> None:
        m = np.ascontiguousarray(masks, dtype=np.uint8)
        if m.shape != (self.n_tips, self.n_patterns):
            raise ValueError(f"masks must have shape ({self.n_tips}, {self.n_patterns}), got {m.shape}")
        self._check(self._lib.b2p_set_tip_states(
            self._h, m.ctypes.data_as(ctypes.POINTER(ctypes.c_uint8))), "b2p_set_tip_states")
        self._remember("tips", m)

    def set_pattern_weights(self, weights) -> None:
        w = np.ascontiguousarray(weights, dtype=np.float64)
        if w.shape != (self.n_patterns,):
            raise ValueError(f"weights must have shape ({self.n_patterns},), got {w.shape}")
        self._check(self._lib.b2p_set_pattern_weights(self._h, _dptr(w)), "b2p_set_pattern_weights")
        self._remember("weights", w)

    def set_eigen(self, evec, ivec, evals, freqs) -> None:
        a = [np.ascontiguousarray(x, dtype=np.float64) for x in (evec, ivec, evals, freqs)]
        n = self.n_states
        if a[0].shape != (n, n) or a[1].shape != (n, n) or a[2].shape != (n,) or a[3].shape != (n,):
            raise ValueError(f"eigen system must be {n}x{n}, {n}x{n}, {n}, {n}")
        self._check(self._lib.b2p_set_eigen(self._h, *map(_dptr, a)), "b2p_set_eigen")
        self._remember("eigen", a)

    def set_categories(self, rates, weights) -> None:
        r = np.ascontiguousarray(rates, dtype=np.float64)
        w = np.ascontiguousarray(weights, dtype=np.float64)
        if r.shape != (self.n_categories,) or w.shape != (self.n_categories,):
            raise ValueError(f"rates/weights must have shape ({self.n_categories},)")
        self._check(self._lib.b2p_set_categories(self._h, _dptr(r), _dptr(w)), "b2p_set_categories")
        self._remember("categories", (r, w))

    def set_model(self, subst_model, site_model) -> None:
        e = subst_model.eigen()
        self.set_eigen(e.evec, e.ivec, e.evals, e.freqs)
        self.set_categories(site_model.rates, site_model.weights)

    # ---------------------------------------------------------------- evaluation
    def loglik_grad_device(self, d_bl: int, d_logl: int, d_grad: Optional[int], stream: int = 0) -> None:
        """Asynchronous evaluation on raw device pointers (e.g. ``tensor.data_ptr()``)."""
        if d_grad is None:
            self._check(self._lib.b2p_loglik(self._h, d_bl, d_logl, stream), "b2p_loglik")
        else:
            self._check(self._lib.b2p_loglik_grad(self._h, d_bl, d_logl, d_grad, stream),
                        "b2p_loglik_grad")

    def loglik_grad_host(self, branch_lengths, want_grad: bool = True) -> Tuple[float, Optional[np.ndarray]]:
        """Host buffers in, host results out (H2D + kernels + D2H + sync inside)."""
        bl = np.ascontiguousarray(branch_lengths, dtype=np.float64)
        if bl.shape != (self.n_branches,):
            raise ValueError(f"branch_lengths must have shape ({self.n_branches},), got {bl.shape}")
        logl = ctypes.c_double()
        grad = np.empty(self.n_branches) if want_grad else None
        self._check(self._lib.b2p_loglik_grad_host(
            self._h, _dptr(bl), ctypes.byref(logl), _dptr(grad) if want_grad else None),
            "b2p_loglik_grad_host")
        return float(logl.value), grad

    def _remember(self, what: str, value) -> None:
        if self.supports_outer_products():
            self._setup[what] = value
            self._version += 1

    def supports_outer_products(self) -> bool:
        """Whether ``branch_outer_products`` runs in one pass: 4 states and 1 or 4 rate categories
        (the outer-products instance itself always works in fp64, also next to an fp32 engine)."""
        return self.n_states == 4 and self.n_categories in (1, 4)

    def branch_outer_products(self, d_bl: int, d_logl: int, d_grad: Optional[int], d_outer: int, stream: int = 0) -> None:
        """G [K][2N-2][4][4] (``b2p_branch_outer_products``) on raw device pointers, in ONE
        evaluation.  An engine created without ``outer_products=True`` keeps a twin instance in
        that mode (same topology, tips, weights and model; set up on first use and whenever one of
        them changes) — the default instance keeps its faster two-columns-per-thread geometry."""
        eng = self if self.outer_products else self._outer_twin()
        eng._check(eng._lib.b2p_branch_outer_products(eng._h, d_bl, d_logl, d_grad, d_outer, stream),
                   "b2p_branch_outer_products")

    def model_gradients(self, d_bl: int, d_outer: int, root_child: int, d_out: int, stream: int = 0) -> None:
        """``b2p_model_gradients``: [16] dlogL/dQ | [4] dlogL/dpi (root prior) | [K] dlogL/dr_k from the
        outer products a ``branch_outer_products`` call with the same branch lengths left in ``d_outer``."""
        eng = self if self.outer_products else self._outer_twin()
        eng._check(eng._lib.b2p_model_gradients(eng._h, d_bl, d_outer, int(root_child), d_out, stream), "b2p_model_gradients")

    def _outer_twin(self) -> "Engine":
        if not self.supports_outer_products():
            raise EngineError("branch_outer_products needs 4 states and 1 or 4 rate categories")
        missing = [k for k in ("topology", "tips", "weights", "eigen", "categories") if k not in self._setup]
        if missing:
            raise EngineError(f"branch_outer_products before {', '.join(missing)} were set")
        if self._twin is None:
            self._twin = Engine(self.n_tips, self.n_patterns, self.n_categories, precision="float64", device=self.device,
                                outer_products=True)
            self._twin_version = -1
        if self._twin_version != self._version:
            t, st = self._twin, self._setup
            if t._setup.get("topology") is not st["topology"]:
                t.set_topology(st["topology"])
            if t._setup.get("tips") is not st["tips"]:
                t.set_tip_states(st["tips"])
            if t._setup.get("weights") is not st["weights"]:
                t.set_pattern_weights(st["weights"])
            if t._setup.get("eigen") is not st["eigen"]:
                t.set_eigen(*st["eigen"])
                t._setup["eigen"] = st["eigen"]
            if t._setup.get("categories") is not st["categories"]:
                t.set_categories(*st["categories"])
                t._setup["categories"] = st["categories"]
            self._twin_version = self._version
        return self._twin

    def set_gradient_operator(self, op=None, scale=None) -> None:
        """Replace (Q, category rates) in the pre-order contraction by a 4x4 operator and a
        per-category scale (``None`` restores the branch-length gradient)."""
        if op is None:
            self._check(self._lib.b2p_set_gradient_operator(self._h, None, None), "b2p_set_gradient_operator")
            return
        a = np.ascontiguousarray(op, dtype=np.float64).reshape(16)
        sc = None if scale is None else np.ascontiguousarray(scale, dtype=np.float64)
        if sc is not None and sc.shape != (self.n_categories,):
            raise ValueError("scale must have one entry per rate category")
        self._check(self._lib.b2p_set_gradient_operator(self._h, _dptr(a), _dptr(sc) if sc is not None else None),
                    "b2p_set_gradient_operator")

    # ------------------------------------------------------------- multi-GPU peers
    def peer_export(self, world: int) -> bytes:
        """Allocate this rank's receive buffer; returns its 64-byte CUDA IPC handle."""
        buf = ctypes.create_string_buffer(64)
        self._check(self._lib.b2p_peer_export(self._h, int(world), buf), "b2p_peer_export")
        return buf.raw

    def peer_attach(self, rank: int, world: int, handles) -> None:
        """``handles``: every rank's export, in rank order.  From here on evaluations return
        results already summed over ranks (the reduction kernel is the all-reduce)."""
        blob = b"".join(handles)
        if len(blob) != 64 * world:
            raise ValueError("need one 64-byte handle per rank")
        self._check(self._lib.b2p_peer_attach(self._h, int(rank), int(world), blob), "b2p_peer_attach")

    def peer_status(self) -> Tuple[bool, int]:
        a, e = ctypes.c_int32(), ctypes.c_int32()
        self._check(self._lib.b2p_peer_status(self._h, ctypes.byref(a), ctypes.byref(e)), "b2p_peer_status")
        return bool(a.value), int(e.value)

    def loglik_grad_host_ptr(self, h_bl: int, h_logl: int, h_grad: Optional[int]) -> None:
        """``b2p_loglik_grad_host`` on raw host addresses (e.g. CPU ``tensor.data_ptr()``):
        [2N-2] doubles in, one double and optionally [2N-2] doubles out; synchronous."""
        rc = _host_raw(self._lib)(self._h, h_bl, h_logl, h_grad)
        if rc:
            self._check(rc, "b2p_loglik_grad_host")

    def set_profiling(self, enabled: bool) -> None:
        self._check(self._lib.b2p_set_profiling(self._h, int(bool(enabled))), "b2p_set_profiling")

    def last_kernel_ms(self) -> float:
        ms = ctypes.c_float()
        self._check(self._lib.b2p_last_kernel_ms(self._h, ctypes.byref(ms)), "b2p_last_kernel_ms")
        return float(ms.value)

    def kernel_ms_stats(self) -> Tuple[float, int]:
        """(total ms, launches) of the traversal kernel since ``set_profiling(True)``."""
        tot, n = ctypes.c_double(), ctypes.c_int32()
        self._check(self._lib.b2p_kernel_ms_stats(self._h, ctypes.byref(tot), ctypes.byref(n)),
                    "b2p_kernel_ms_stats")
        return float(tot.value), int(n.value)

    def branch_partials(self, node: int) -> np.ndarray:
        """P_node x_node of an internal non-root node as [P, K, 4] (debug / tests)."""
        out = np.empty((self.n_patterns, self.n_categories, 4))
        self._check(self._lib.b2p_get_branch_partials(self._h, int(node), _dptr(out)),
                    "b2p_get_branch_partials")
        return out

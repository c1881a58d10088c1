"""Time trees on the GPU: node-height reparameterisation, its log-Jacobian, branch lengths
under a strict clock and the constant-size coalescent.

The classes mirror the objects of the reference's macro model
(/root/reference/flu_H3N2/phylotorch-JC69.template:14-58,75-100 — torchtree's
``ReparameterizedTimeTreeModel``, ``StrictClockModel``, ``ConstantCoalescentModel``) and the
functions its micro benchmark times (/root/reference/plot.Rmd:49-60: ``ratio_transform``,
``ratio_transform_jacobian``, ``coalescent``).  Each function is one kernel launch through
the C ABI (``b2p_timetree_*`` in ``include/b200phylo.h``) wrapped in a
``torch.autograd.Function`` whose backward is another single launch, so a whole
ratios → heights → branch lengths → likelihood → gradient chain stays on the device.
There is no CPU fallback.
"""
from __future__ import annotations

import ctypes
import re
from typing import Optional, Sequence

import numpy as np
import torch

from .engine import EngineError, load_library
from .tree import Tree

__all__ = ["TimeTreeEngine", "ReparameterizedTimeTreeModel", "ConstantCoalescentModel", "JointTimeTreeDensity",
           "tip_heights_from_names", "TIMETREE_SYMBOLS"]

TIMETREE_SYMBOLS = (
    "b2p_timetree_create", "b2p_timetree_destroy", "b2p_timetree_last_error", "b2p_timetree_get_info",
    "b2p_timetree_heights", "b2p_timetree_heights_vjp", "b2p_timetree_log_jacobian",
    "b2p_timetree_branch_lengths", "b2p_timetree_branch_lengths_vjp", "b2p_coalescent_constant",
    "b2p_joint_density_grad",
)
# the no-allocation entry point (csrc/b2p_noalloc.cu; no counterpart in the reference)
NOALLOC_SYMBOLS = ("b2p_coalescent_constant_noalloc", "b2p_noalloc_max_tips", "b2p_noalloc_last_error")

_declared = False


def _lib():
    global _declared
    lib = load_library()
    if not _declared:
        vp = ctypes.c_void_p
        i32p = ctypes.POINTER(ctypes.c_int32)
        lib.b2p_timetree_create.argtypes = [ctypes.c_int32, i32p, ctypes.POINTER(ctypes.c_double), ctypes.c_int32,
                                            ctypes.POINTER(vp)]
        lib.b2p_timetree_destroy.argtypes = [vp]
        lib.b2p_timetree_last_error.argtypes = [vp]
        lib.b2p_timetree_last_error.restype = ctypes.c_char_p
        lib.b2p_timetree_get_info.argtypes = [vp, i32p, i32p, i32p]
        lib.b2p_timetree_heights.argtypes = [vp] * 4
        lib.b2p_timetree_heights_vjp.argtypes = [vp] * 6
        lib.b2p_timetree_log_jacobian.argtypes = [vp] * 5
        lib.b2p_timetree_branch_lengths.argtypes = [vp] * 5
        lib.b2p_timetree_branch_lengths_vjp.argtypes = [vp] * 7
        lib.b2p_coalescent_constant.argtypes = [vp] * 7
        lib.b2p_joint_density_grad.argtypes = [vp, vp, vp, vp, vp, ctypes.c_int32, vp, vp, vp, vp, vp]
        for name in TIMETREE_SYMBOLS:
            if name != "b2p_timetree_last_error":
                getattr(lib, name).restype = ctypes.c_int
        _declared = True
    return lib


_DATE = re.compile(r"_(\d+(?:\.\d+)?)$")


def tip_heights_from_names(names: Sequence[str]) -> np.ndarray:
    """Sampling dates are the suffix after the last underscore of the taxon names
    (/root/reference/bin/helper.py:28-47 appends ``_<date>``); height = latest date - date."""
    dates = []
    for n in names:
        m = _DATE.search(n)
        if not m:
            raise ValueError(f"taxon name {n!r} does not end in _<date>")
        dates.append(float(m.group(1)))
    dates = np.asarray(dates, dtype=np.float64)
    return dates.max() - dates


class TimeTreeEngine:
    """One dated topology on one B200 (``b2p_timetree``)."""

    def __init__(self, triples: np.ndarray, tip_heights: np.ndarray, device: int = 0):
        self._lib = _lib()
        self._h = ctypes.c_void_p()
        tr = np.ascontiguousarray(triples, dtype=np.int32)
        th = np.ascontiguousarray(tip_heights, dtype=np.float64)
        if tr.ndim != 2 or tr.shape[1] != 3 or th.shape != (tr.shape[0] + 1,):
            raise ValueError("triples must be [N-1, 3] and tip_heights [N]")
        self.n_tips = int(th.shape[0])
        self.device = int(device)
        rc = self._lib.b2p_timetree_create(self.n_tips, tr.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)),
                                           th.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), self.device,
                                           ctypes.byref(self._h))
        if rc != 0:
            raise EngineError(f"b2p_timetree_create failed ({rc}): "
                              f"{self._lib.b2p_timetree_last_error(None).decode()}")

    # ------------------------------------------------------------------ plumbing
    def _check(self, rc: int, what: str) -> None:
        if rc != 0:
            raise EngineError(f"{what} failed ({rc}): {self._lib.b2p_timetree_last_error(self._h).decode()}")

    def _vec(self, t: torch.Tensor, n: int, what: str) -> torch.Tensor:
        if not t.is_cuda or t.device.index != self.device:
            raise EngineError(f"{what}: expected a tensor on cuda:{self.device}, got {t.device} (no CPU fallback)")
        if t.numel() != n:
            raise ValueError(f"{what}: expected {n} values, got {tuple(t.shape)}")
        return t.detach().to(torch.float64).contiguous().reshape(-1)

    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def _new(self, n: int) -> torch.Tensor:
        return torch.empty(n, dtype=torch.float64, device=f"cuda:{self.device}")

    def info(self) -> dict:
        a, b, c = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32()
        self._check(self._lib.b2p_timetree_get_info(self._h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c)), "get_info")
        return {"n_levels": a.value, "n_segments": b.value, "in_smem": c.value}

    # ----------------------------------------------------------------- functions
    def heights(self, params: torch.Tensor) -> torch.Tensor:
        p = self._vec(params, self.n_tips - 1, "params")
        out = self._new(self.n_tips - 1)
        self._check(self._lib.b2p_timetree_heights(self._h, p.data_ptr(), out.data_ptr(), self._stream()), "heights")
        return out

    def heights_vjp(self, params, heights, grad_heights) -> torch.Tensor:
        m = self.n_tips - 1
        p, h, g = self._vec(params, m, "params"), self._vec(heights, m, "heights"), self._vec(grad_heights, m, "grad")
        out = self._new(m)
        self._check(self._lib.b2p_timetree_heights_vjp(self._h, p.data_ptr(), h.data_ptr(), g.data_ptr(), out.data_ptr(),
                                                       self._stream()), "heights_vjp")
        return out

    def log_jacobian(self, heights, want_grad: bool = False):
        m = self.n_tips - 1
        h = self._vec(heights, m, "heights")
        out = self._new(1)
        grad = self._new(m) if want_grad else None
        self._check(self._lib.b2p_timetree_log_jacobian(self._h, h.data_ptr(), out.data_ptr(),
                                                        grad.data_ptr() if want_grad else None, self._stream()),
                    "log_jacobian")
        return out, grad

    def branch_lengths(self, heights, rate: Optional[torch.Tensor] = None) -> torch.Tensor:
        h = self._vec(heights, self.n_tips - 1, "heights")
        r = self._vec(rate, 1, "rate") if rate is not None else None
        out = self._new(2 * self.n_tips - 2)
        self._check(self._lib.b2p_timetree_branch_lengths(self._h, h.data_ptr(), r.data_ptr() if r is not None else None,
                                                          out.data_ptr(), self._stream()), "branch_lengths")
        return out

    def branch_lengths_vjp(self, heights, rate, grad_bl):
        h = self._vec(heights, self.n_tips - 1, "heights")
        r = self._vec(rate, 1, "rate") if rate is not None else None
        g = self._vec(grad_bl, 2 * self.n_tips - 2, "grad_bl")
        gh, gr = self._new(self.n_tips - 1), self._new(1)
        self._check(self._lib.b2p_timetree_branch_lengths_vjp(self._h, h.data_ptr(), r.data_ptr() if r is not None else None,
                                                              g.data_ptr(), gh.data_ptr(), gr.data_ptr(), self._stream()),
                    "branch_lengths_vjp")
        return gh, gr

    def constant_coalescent(self, heights, theta, want_grad: bool = True):
        h = self._vec(heights, self.n_tips - 1, "heights")
        th = self._vec(theta, 1, "theta")
        out = self._new(1)
        gh = self._new(self.n_tips - 1) if want_grad else None
        gt = self._new(1) if want_grad else None
        self._check(self._lib.b2p_coalescent_constant(self._h, h.data_ptr(), th.data_ptr(), out.data_ptr(),
                                                      gh.data_ptr() if want_grad else None,
                                                      gt.data_ptr() if want_grad else None, self._stream()), "coalescent")
        return out, gh, gt

    def joint_density_grad(self, like_engine, params, rate=None, theta=None, include_jacobian: bool = True):
        """``b2p_joint_density_grad``: returns ``(out[4], grad_params, grad_rate, grad_theta)``."""
        m = self.n_tips - 1
        p = self._vec(params, m, "params")
        r = self._vec(rate, 1, "rate") if rate is not None else None
        th = self._vec(theta, 1, "theta") if theta is not None else None
        out = torch.zeros(4, dtype=torch.float64, device=p.device)
        gp, gr = self._new(m), self._new(1)
        gt = self._new(1) if th is not None else None
        self._check(self._lib.b2p_joint_density_grad(
            like_engine._h, self._h, p.data_ptr(), r.data_ptr() if r is not None else None,
            th.data_ptr() if th is not None else None, 1 if include_jacobian else 0, out.data_ptr(), gp.data_ptr(),
            gr.data_ptr(), gt.data_ptr() if gt is not None else None, self._stream()), "joint_density_grad")
        return out, gp, gr, gt

    def close(self) -> None:
        if self._h:
            self._lib.b2p_timetree_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ----------------------------------------------------------------------- autograd
class _Heights(torch.autograd.Function):
    @staticmethod
    def forward(ctx, params, eng: TimeTreeEngine):
        h = eng.heights(params)
        ctx.eng = eng
        ctx.save_for_backward(params.detach(), h)
        return h

    @staticmethod
    def backward(ctx, g):
        params, h = ctx.saved_tensors
        return ctx.eng.heights_vjp(params, h, g), None


class _LogJacobian(torch.autograd.Function):
    @staticmethod
    def forward(ctx, heights, eng: TimeTreeEngine):
        out, grad = eng.log_jacobian(heights, want_grad=ctx.needs_input_grad[0])
        if grad is not None:
            ctx.save_for_backward(grad)
        return out[0]

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return g * grad, None


class _BranchLengths(torch.autograd.Function):
    @staticmethod
    def forward(ctx, heights, rate, eng: TimeTreeEngine):
        ctx.eng = eng
        ctx.has_rate = rate is not None
        ctx.save_for_backward(heights.detach(), rate.detach() if rate is not None else heights.new_empty(0))
        return eng.branch_lengths(heights, rate)

    @staticmethod
    def backward(ctx, g):
        heights, rate = ctx.saved_tensors
        gh, gr = ctx.eng.branch_lengths_vjp(heights, rate if ctx.has_rate else None, g)
        return gh, (gr.reshape(rate.shape) if ctx.has_rate else None), None


class _ConstantCoalescent(torch.autograd.Function):
    @staticmethod
    def forward(ctx, heights, theta, eng: TimeTreeEngine):
        need = ctx.needs_input_grad[0] or ctx.needs_input_grad[1]
        out, gh, gt = eng.constant_coalescent(heights, theta, want_grad=need)
        if need:
            ctx.save_for_backward(gh, gt.reshape(theta.shape))
        return out[0]

    @staticmethod
    def backward(ctx, g):
        gh, gt = ctx.saved_tensors
        return g * gh, g * gt, None


class _JointDensity(torch.autograd.Function):
    """value and gradient in the forward call (one C call, nine launches); backward scales."""

    @staticmethod
    def forward(ctx, params, rate, theta, tt: TimeTreeEngine, like_engine, include_jacobian: bool):
        out, gp, gr, gt = tt.joint_density_grad(like_engine, params, rate, theta, include_jacobian)
        ctx.has = (rate is not None, theta is not None)
        ctx.save_for_backward(gp, gr.reshape(rate.shape) if rate is not None else gr,
                              gt.reshape(theta.shape) if theta is not None else gr)
        ctx.mark_non_differentiable(out)
        return out[0].clone(), out

    @staticmethod
    def backward(ctx, g, _unused):
        gp, gr, gt = ctx.saved_tensors
        return g * gp, (g * gr if ctx.has[0] else None), (g * gt if ctx.has[1] else None), None, None, None


# ------------------------------------------------------------------------- models
class ReparameterizedTimeTreeModel:
    """Dated tree parameterised by ratios and the root height
    (/root/reference/flu_H3N2/phylotorch-JC69.template:14-44).

    ``ratios`` [N-2] (internal nodes N..2N-3) and ``root_height`` [1] are torch tensors on the
    GPU — leaves, or the outputs of whatever transform the caller puts in front (the template
    uses a sigmoid and an affine shift).  ``branch_lengths()`` are in time units, indexed by
    child node id; the clock model scales them.
    """

    def __init__(self, id_: Optional[str], tree: Tree, tip_heights: Optional[np.ndarray] = None,
                 ratios: Optional[torch.Tensor] = None, root_height: Optional[torch.Tensor] = None,
                 device: int = 0):
        self.id = id_
        self.tree = tree
        n = tree.n_tips
        if tip_heights is None:
            tip_heights = tip_heights_from_names(tree.names)
        self.tip_heights = np.asarray(tip_heights, dtype=np.float64)
        self.engine = TimeTreeEngine(tree.post_triples, self.tip_heights, device=device)
        dev = torch.device(f"cuda:{device}")
        if ratios is None or root_height is None:
            x = initial_params_from_tree(tree, self.tip_heights)
            if ratios is None:
                ratios = torch.tensor(x[: n - 2], dtype=torch.float64, device=dev)
            if root_height is None:
                root_height = torch.tensor(x[n - 2:], dtype=torch.float64, device=dev)
        self.ratios = ratios            # tensor, or an object with .tensor (Parameter / TransformedParameter)
        self.root_height = root_height

    fuses_clock_rate = True  # branch_lengths(rate=...) applies the strict-clock rate in the same launch

    @property
    def n_tips(self) -> int:
        return self.tree.n_tips

    @property
    def taxa(self):
        return self.tree.names

    @property
    def post_triples(self) -> np.ndarray:
        return self.tree.post_triples

    def params(self) -> torch.Tensor:
        r = getattr(self.ratios, "tensor", self.ratios)
        h = getattr(self.root_height, "tensor", self.root_height)
        return torch.cat([r.reshape(-1), h.reshape(-1)])

    def heights(self) -> torch.Tensor:
        """Internal node heights [N-1] (``ratio_transform``)."""
        return _Heights.apply(self.params(), self.engine)

    def log_jacobian(self, heights: Optional[torch.Tensor] = None) -> torch.Tensor:
        """log|det d heights / d (ratios, root height)| (``ratio_transform_jacobian``)."""
        return _LogJacobian.apply(self.heights() if heights is None else heights, self.engine)

    def branch_lengths(self, rate: Optional[torch.Tensor] = None, heights: Optional[torch.Tensor] = None) -> torch.Tensor:
        if rate is not None and not rate.is_cuda:  # a clock rate read from JSON lives on the host
            rate = rate.to(f"cuda:{self.engine.device}")
        return _BranchLengths.apply(self.heights() if heights is None else heights, rate, self.engine)

    def parameters(self):
        out = []
        for p in (self.ratios, self.root_height):
            out += p.parameters() if hasattr(p, "parameters") else [p]
        return out

    def close(self) -> None:
        self.engine.close()


class ConstantCoalescentModel:
    """Kingman coalescent with one population size
    (/root/reference/flu_H3N2/phylotorch-JC69.template:75-100; micro function ``coalescent``)."""

    def __init__(self, id_: Optional[str], tree_model: ReparameterizedTimeTreeModel, theta: torch.Tensor):
        self.id = id_
        self.tree_model = tree_model
        self.theta = theta

    def __call__(self, heights: Optional[torch.Tensor] = None) -> torch.Tensor:
        h = self.tree_model.heights() if heights is None else heights
        return _ConstantCoalescent.apply(h, getattr(self.theta, "tensor", self.theta), self.tree_model.engine)

    def parameters(self):
        return self.theta.parameters() if hasattr(self.theta, "parameters") else [self.theta]


class JointTimeTreeDensity:
    """log p(alignment | tree, rate) + log p(heights | theta) + log|det J| as ONE engine call —
    the density the reference's macro benchmark differentiates 5000 times per run
    (/root/reference/flu_H3N2/phylotorch-JC69.template:11-100, /root/reference/modules/macro_flu.nf:6,133).
    ``like`` is a ``TreeLikelihoodModel`` built on ``tree_model``; priors on rate / theta are
    plain torch terms the caller adds."""

    def __init__(self, like, tree_model: ReparameterizedTimeTreeModel, coalescent: Optional[ConstantCoalescentModel] = None,
                 include_jacobian: bool = True):
        if like.world_size > 1:
            raise NotImplementedError("sharded likelihoods go through TreeLikelihoodModel (all-reduce per evaluation)")
        self.like, self.tree_model, self.coalescent = like, tree_model, coalescent
        self.include_jacobian = include_jacobian
        self.terms: Optional[torch.Tensor] = None  # [total, logL, coalescent, log-Jacobian] of the last call

    def __call__(self) -> torch.Tensor:
        clock = self.like.clock_model
        rate = clock.rates if clock is not None else None
        dev = f"cuda:{self.tree_model.engine.device}"
        if rate is not None and not rate.is_cuda:
            rate = rate.to(dev)
        theta = None
        if self.coalescent is not None:
            theta = getattr(self.coalescent.theta, "tensor", self.coalescent.theta)
        total, self.terms = _JointDensity.apply(self.tree_model.params(), rate, theta, self.tree_model.engine,
                                                self.like.engine, self.include_jacobian)
        return total


def initial_params_from_tree(tree: Tree, tip_heights: np.ndarray, min_branch: float = 1e-6) -> np.ndarray:
    """(ratios, root height) reproducing the node heights implied by the tree's branch lengths
    (`--heights_init tree`, /root/reference/modules/macro_flu.nf:124-131).  Set-up only (host,
    once per tree): heights are pushed above their bounds where the input tree violates them."""
    n = tree.n_tips
    depth = np.zeros(tree.n_nodes)
    for v in range(tree.n_nodes - 1, n - 1, -1):
        for c in tree.children[v]:
            depth[c] = depth[v] + tree.branch_lengths[c]
    # anchor: the tip whose date fixes the time axis best is the one the root height must clear
    root_h = float(np.max(depth[:n] + tip_heights))
    h = root_h - depth
    bound = np.zeros(tree.n_nodes)
    bound[:n] = tip_heights
    for v in range(n, tree.n_nodes):
        l, r = tree.children[v]
        bound[v] = max(bound[l], bound[r])
        h[v] = max(h[v], bound[v] + min_branch, h[l] + min_branch if l >= n else 0.0, h[r] + min_branch if r >= n else 0.0)
    par = tree.parent
    x = np.empty(n - 1)
    x[n - 2] = h[tree.root]
    for v in range(n, tree.n_nodes - 1):
        x[v - n] = (h[v] - bound[v]) / (h[par[v]] - bound[v])
    return x

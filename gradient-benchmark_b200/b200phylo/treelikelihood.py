"""``TreeLikelihoodModel`` — the plugin surface of the hot path.

In the reference a likelihood engine is swapped in by naming its class in the JSON
``"type"`` of the ``"like"`` object (/root/reference/flu_H3N2/phylotorch-JC69.template:11-74;
``--engine bitorch`` does exactly that for torchtree-bito,
/root/reference/modules/macro_flu.nf:119-122).  This class keeps that surface:

* same five named sub-objects (``tree_model``, ``site_model``, ``branch_model``,
  ``site_pattern``, ``substitution_model``) and a ``from_json(data, dic)`` classmethod;
* ``model()`` returns the log-likelihood as a ``torch.Tensor`` that is differentiable
  w.r.t. ``tree_model.branch_lengths()`` (and, through torch's chain rule, the clock
  rate and anything upstream of the branch lengths);
* like torchtree-bito, ``forward`` asks the native engine for the log-likelihood AND the
  analytic branch gradient in one pass and ``backward`` only scales the cached gradient.

torchtree itself is not installable here, so the sub-objects are
duck-typed: anything with ``branch_lengths()`` / ``rates`` / ``eigen()`` works, and the
small classes in this file stand in for torchtree's when it is absent.

There is no CPU fallback: evaluation goes through ``libb200phylo.so`` on a B200 or raises.
"""
from __future__ import annotations

from typing import Any, Dict, Optional

import numpy as np
import torch

from .alignment import Alignment, SitePattern, compress_alignment
from .diffmodels import (DifferentiableGammaSiteModel, DifferentiableGTR, DifferentiableHKY,
                         DifferentiableWeibullSiteModel, is_differentiable)
from .engine import Engine, EngineError
from .parallel import allreduce_sum_, attach_peers, rank_world, shard_bounds
from .sitemodel import ConstantSiteModel, GammaSiteModel, WeibullSiteModel
from .substitution import GTR, HKY, JC69, SubstitutionModel
from .tree import Tree

__all__ = ["TreeLikelihoodModel", "TreeLikelihoodFunction", "TreeLikelihoodParameterFunction", "UnrootedTreeModel",
           "StrictClockModel", "Parameter", "TransformedParameter"]


class Parameter:
    """Minimal stand-in for torchtree's ``Parameter``: an id + a tensor."""

    def __init__(self, id_: Optional[str], tensor: torch.Tensor):
        self.id = id_
        self.tensor = tensor

    @property
    def requires_grad(self) -> bool:
        return self.tensor.requires_grad

    @requires_grad.setter
    def requires_grad(self, flag: bool) -> None:
        self.tensor.requires_grad_(flag)


class TransformedParameter:
    """``transform(x)`` of an unconstrained ``Parameter`` — the reference's JSON wraps ratios
    in a sigmoid and the root height in an affine shift
    (/root/reference/flu_H3N2/phylotorch-JC69.template:17-42)."""

    _TRANSFORMS = {
        "SigmoidTransform": lambda x, kw: torch.sigmoid(x),
        "ExpTransform": lambda x, kw: torch.exp(x),
        "AffineTransform": lambda x, kw: kw.get("loc", 0.0) + kw.get("scale", 1.0) * x,
    }

    def __init__(self, id_: Optional[str], x: Parameter, transform: str, parameters: Optional[dict] = None):
        name = transform.split(".")[-1]
        if name not in self._TRANSFORMS:
            raise ValueError(f"unsupported transform {transform!r}")
        self.id = id_
        self.x = x
        self._fn = self._TRANSFORMS[name]
        self._kw = dict(parameters or {})

    @property
    def tensor(self) -> torch.Tensor:
        return self._fn(self.x.tensor, self._kw)

    def parameters(self):
        return [self.x]


class UnrootedTreeModel:
    """Tree with free branch lengths — the micro benchmark's parameterisation
    ("input = branch lengths", /root/reference/plot.Rmd:41-48).  ``branch_lengths()``
    has length ``2N-2`` indexed by child node id, the same order as the gradient."""

    def __init__(self, id_: Optional[str], tree: Tree, branch_lengths: Optional[torch.Tensor] = None,
                 device: Optional[torch.device] = None):
        self.id = id_
        self.tree = tree
        if branch_lengths is None:
            branch_lengths = torch.tensor(tree.branch_lengths, dtype=torch.float64, device=device)
        self._bl = Parameter(f"{id_}.blens" if id_ else None, branch_lengths)

    @property
    def n_tips(self) -> int:
        return self.tree.n_tips

    @property
    def taxa(self):
        return self.tree.names

    @property
    def post_triples(self) -> np.ndarray:
        return self.tree.post_triples

    def branch_lengths(self) -> torch.Tensor:
        return self._bl.tensor

    def parameters(self):
        return [self._bl]


class StrictClockModel:
    """One rate for every branch (/root/reference/flu_H3N2/phylotorch-JC69.template:49-58)."""

    def __init__(self, id_: Optional[str], rate):
        self.id = id_
        # a tensor, or an already-built (Transformed)Parameter — the template's "rate" object, which
        # its priors and variational blocks look up by id
        self._rate = rate if hasattr(rate, "tensor") else Parameter(f"{id_}.rate" if id_ else None, rate)

    @property
    def rates(self) -> torch.Tensor:
        return self._rate.tensor

    def parameters(self):
        return [self._rate]


class TreeLikelihoodFunction(torch.autograd.Function):
    """logL(branch lengths) with the analytic gradient computed in ``forward``.

    ``branch_lengths`` may live on the engine's CUDA device (zero-copy, asynchronous on
    the current stream) or on the host (pinned staging inside ``b2p_loglik_grad_host``).
    With an initialised process group the per-rank results are summed over ranks: inside
    the engine's reduction kernel over NVLink peer memory (``model.fused_allreduce``), else
    with one NCCL all-reduce of ``2N-1`` doubles.
    """

    @staticmethod
    def forward(ctx, branch_lengths: torch.Tensor, model: "TreeLikelihoodModel") -> torch.Tensor:
        eng = model.engine
        need_grad = branch_lengths.requires_grad
        bl = branch_lengths.detach()
        if bl.dim() != 1 or bl.shape[0] != eng.n_branches:
            raise ValueError(f"branch_lengths must have shape ({eng.n_branches},), got {tuple(bl.shape)}")
        if bl.is_cuda:
            if bl.device.index != eng.device:
                raise EngineError(f"branch lengths on cuda:{bl.device.index}, engine on cuda:{eng.device}")
            bl64 = bl.to(torch.float64).contiguous()
            out = torch.empty(1 + eng.n_branches, dtype=torch.float64, device=bl.device)
            stream = torch.cuda.current_stream(bl.device).cuda_stream
            eng.loglik_grad_device(bl64.data_ptr(), out.data_ptr(),
                                   out.data_ptr() + 8 if need_grad else None, stream)
            if not need_grad:
                out = out[:1]
        else:
            # host tensors: the C entry point reads the branch lengths and writes [logL, gradient]
            # straight from / into torch's own buffers (no numpy detour, one allocation)
            bl64 = bl if (bl.dtype == torch.float64 and bl.is_contiguous()) else bl.to(torch.float64).contiguous()
            out = torch.empty(1 + (eng.n_branches if need_grad else 0), dtype=torch.float64)
            eng.loglik_grad_host_ptr(bl64.data_ptr(), out.data_ptr(), out.data_ptr() + 8 if need_grad else None)
        if model.world_size > 1 and not model.fused_allreduce:  # fused: the engine already summed over ranks
            if not out.is_cuda:
                out = out.to(f"cuda:{eng.device}")
            allreduce_sum_(out, model.group)
            out = out.to(bl.device)
        if need_grad:
            ctx.save_for_backward(out[1:])
        ctx.in_dtype = branch_lengths.dtype
        logl = out[0]
        return logl if logl.dtype == branch_lengths.dtype else logl.to(branch_lengths.dtype)

    @staticmethod
    def backward(ctx, grad_output):
        (g,) = ctx.saved_tensors
        # d logL / d logL = 1 is the common upstream gradient (``logl.backward()``): hand the
        # gradient buffer over as it is instead of a multiplication by one (host tensors: a
        # 2N-2-element multiply costs as much as a tenth of an evaluation at 1000 taxa)
        if not grad_output.is_cuda and grad_output.dim() == 0 and grad_output.item() == 1.0:
            out = g
        else:
            out = grad_output * g
        return (out if out.dtype == ctx.in_dtype else out.to(ctx.in_dtype)), None


class TreeLikelihoodParameterFunction(torch.autograd.Function):
    """logL(branch lengths, Q, frequencies, category rates): gradients w.r.t. all four from ONE
    engine evaluation.

    ``forward`` calls ``b2p_branch_outer_products`` — log-likelihood, branch gradient and the
    per-branch outer products G_{b,k} = sum_p (w_p / L_p) c_k q m^T in one pass — and
    ``b2p_model_gradients``, which contracts G with dP/d(Q, rates) on the device (divided
    differences of the eigenvalues) and adds the root-prior term for the frequencies;
    ``backward`` only scales.  torchtree gets the same numbers by autograd of the whole pruning
    recursion (/root/reference/modules/micro.nf:52 ``--gtr``)."""

    @staticmethod
    def forward(ctx, branch_lengths, q, freqs, rates, model: "TreeLikelihoodModel"):
        eng = model.engine
        dev = torch.device(f"cuda:{eng.device}")
        bl = branch_lengths.detach().to(dev, torch.float64).contiguous()
        if bl.dim() != 1 or bl.shape[0] != eng.n_branches:
            raise ValueError(f"branch_lengths must have shape ({eng.n_branches},), got {tuple(bl.shape)}")
        out = torch.empty(1 + eng.n_branches, dtype=torch.float64, device=dev)
        g = torch.empty(eng.n_categories, eng.n_branches, 4, 4, dtype=torch.float64, device=dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        one_pass = eng.supports_outer_products()
        if one_pass:
            eng.branch_outer_products(bl.data_ptr(), out.data_ptr(), out.data_ptr() + 8, g.data_ptr(), stream)
        else:
            # other category counts: the outer products entry by entry (16 K evaluations with
            # b2p_set_gradient_operator), then logL and the branch gradient
            from .paramgrad import branch_outer_products
            g = branch_outer_products(eng, bl, one_pass=False)
            eng.loglik_grad_device(bl.data_ptr(), out.data_ptr(), out.data_ptr() + 8, stream)
        if model.world_size > 1:
            # pattern shards: logL, the branch gradient and the outer products are sums over patterns
            # (the outer-products instance is not peer-attached: one NCCL all-reduce each)
            allreduce_sum_(out, model.group)
            allreduce_sum_(g, model.group)
        root_child = int(model.tree_model.post_triples[-1][1])
        if one_pass:
            mg = torch.empty(20 + eng.n_categories, dtype=torch.float64, device=dev)
            eng.model_gradients(bl.data_ptr(), g.data_ptr(), root_child, mg.data_ptr(), stream)
        else:  # ... and the contraction with dP/d(Q, rates) by torch (matrix exponentials of 4x4 matrices)
            from .paramgrad import surrogate
            with torch.enable_grad():
                leaves = [x.detach().to(dev).requires_grad_(True) for x in (q, freqs, rates)]
                grads = torch.autograd.grad(surrogate(g, bl, *leaves, root_child), leaves, allow_unused=True)
            mg = torch.cat([(gx if gx is not None else torch.zeros_like(x)).reshape(-1) for gx, x in zip(grads, leaves)])
        ctx.save_for_backward(out[1:], mg)
        ctx.in_devices = (branch_lengths.device, q.device, freqs.device, rates.device)
        ctx.in_dtype = branch_lengths.dtype
        return out[0].to(branch_lengths.device, branch_lengths.dtype)

    @staticmethod
    def backward(ctx, grad_output):
        g_bl, mg = ctx.saved_tensors
        need = ctx.needs_input_grad
        go = grad_output.to(g_bl.device, torch.float64)
        out = [None, None, None, None, None]
        if need[0]:
            out[0] = (go * g_bl).to(ctx.in_devices[0], ctx.in_dtype)
        if need[1]:
            out[1] = (go * mg[:16]).reshape(4, 4).to(ctx.in_devices[1])
        if need[2]:
            out[2] = (go * mg[16:20]).to(ctx.in_devices[2])
        if need[3]:
            out[3] = (go * mg[20:]).to(ctx.in_devices[3])
        return tuple(out)


class TreeLikelihoodModel:
    """Tree + site patterns + substitution model + site model (+ clock) → logL tensor.

    Constructor argument order follows torchtree's / torchtree-bito's
    ``TreeLikelihoodModel(id_, site_pattern, tree_model, subst_model, site_model,
    clock_model)`` (upstream recall).  ``precision`` is the engine's
    working type for partials ("float64" default, "float32" for ``-d float32``,
    /root/reference/modules/micro.nf:42-45).  When a ``torch.distributed`` group is
    initialised (or ``rank``/``world_size`` are given) the instance keeps only its
    contiguous pattern slice and all-reduces results.
    """

    def __init__(self, id_: Optional[str], site_pattern: SitePattern, tree_model, subst_model,
                 site_model=None, clock_model=None, precision: str = "float64",
                 device: Optional[int] = None, rank: Optional[int] = None,
                 world_size: Optional[int] = None, group=None, warps_per_block: int = 0):
        self.id = id_
        self.tree_model = tree_model
        self.subst_model = subst_model
        self.site_model = site_model if site_model is not None else ConstantSiteModel()
        self.clock_model = clock_model
        self.group = group
        r, w = rank_world(group)
        self.rank = r if rank is None else int(rank)
        self.world_size = w if world_size is None else int(world_size)
        taxa = list(tree_model.taxa) if getattr(tree_model, "taxa", None) else None
        if taxa and list(site_pattern.names) != taxa:
            site_pattern = site_pattern.reorder(taxa)
        self.site_pattern_full = site_pattern
        if self.world_size > 1:
            if site_pattern.n_patterns < self.world_size:
                raise ValueError("fewer site patterns than ranks")
            lo, hi = shard_bounds(site_pattern.n_patterns, self.world_size, self.rank)
            site_pattern = site_pattern.slice(lo, hi)
        self.site_pattern = site_pattern
        if device is None:
            device = torch.cuda.current_device() if torch.cuda.is_available() else 0
        n_cat = len(np.asarray(self.site_model.rates))
        self.engine = Engine(tree_model.n_tips, site_pattern.n_patterns, n_cat,
                             precision=precision, device=device, warps_per_block=warps_per_block,
                             n_states=getattr(site_pattern, "n_states", 4))
        self.engine.set_topology(tree_model.post_triples)
        self.engine.set_tip_states(site_pattern.masks)
        self.engine.set_pattern_weights(site_pattern.weights)
        self.update_model()
        # multi-GPU: the engine's own reduction kernel doubles as the all-reduce over NVLink
        # peer memory; NCCL stays the path only where that cannot be set up
        self.fused_allreduce = False
        if self.world_size > 1 and rank is None:
            self.fused_allreduce = attach_peers(self.engine, group)

    def update_model(self) -> None:
        """Push the substitution / site model to the engine (call after changing them; models whose
        parameters are tensors are followed automatically, see ``_refresh_model``)."""
        self.engine.set_model(self.subst_model, self.site_model)
        self._model_key = self._current_model_key()

    def _current_model_key(self):
        """Identity and in-place version of every tensor a differentiable model is made of."""
        key = []
        for m in (self.subst_model, self.site_model):
            ps = getattr(m, "differentiable_parameters", None)
            if ps is not None:
                for p in ps():
                    t = p.tensor if hasattr(p, "tensor") else p
                    key.append((id(t), t._version))
        return tuple(key)

    def _refresh_model(self) -> None:
        """Tensor-valued model parameters may have changed since the last call (an optimiser step):
        redo the eigen system and push it — only then (a 4x4 eigendecomposition on the host and a
        dropped CUDA graph cost about as much as a small evaluation)."""
        if self._current_model_key() != self._model_key:
            self.update_model()

    # ---------------------------------------------------------------- evaluation
    def _scaled_branch_lengths(self) -> torch.Tensor:
        if self.clock_model is not None and getattr(self.tree_model, "fuses_clock_rate", False):
            return self.tree_model.branch_lengths(rate=self.clock_model.rates)
        bl = self.tree_model.branch_lengths()
        if self.clock_model is not None:
            rates = self.clock_model.rates
            # the rate lives on the engine's device; host-side branch lengths move there with it
            if rates.device != bl.device:
                bl = bl.to(rates.device)
            bl = rates * bl
        return bl

    def __call__(self, *args, **kwargs) -> torch.Tensor:
        if torch.is_grad_enabled() and (is_differentiable(self.subst_model) or is_differentiable(self.site_model)):
            return self._call_with_model_parameters()
        self._refresh_model()
        return TreeLikelihoodFunction.apply(self._scaled_branch_lengths(), self)

    def _call_with_model_parameters(self) -> torch.Tensor:
        """A substitution- or site-model parameter asks for a gradient: one pass that also returns
        the per-branch outer products (``TreeLikelihoodParameterFunction``)."""
        if self.engine.n_states != 4:
            raise EngineError("gradients w.r.t. model parameters are built for 4-state models")
        if self.world_size > 1 and self.fused_allreduce and not self.engine.supports_outer_products():
            raise EngineError("gradients w.r.t. model parameters on pattern shards need 1 or 4 rate categories")
        self._refresh_model()
        dev = torch.device(f"cuda:{self.engine.device}")
        sm, st = self.subst_model, self.site_model
        if hasattr(sm, "q_tensor"):
            q, freqs = sm.q_tensor(), sm.frequencies_tensor()
        else:
            e = sm.eigen()
            q, freqs = torch.tensor(e.q, dtype=torch.float64, device=dev), torch.tensor(e.freqs, dtype=torch.float64, device=dev)
        rates = st.rates_tensor() if hasattr(st, "rates_tensor") else torch.tensor(np.asarray(st.rates), dtype=torch.float64, device=dev)
        return TreeLikelihoodParameterFunction.apply(self._scaled_branch_lengths(), q, freqs, rates, self)

    def log_likelihood(self) -> float:
        with torch.no_grad():
            return float(self())

    def parameters(self):
        out = list(self.tree_model.parameters())
        if self.clock_model is not None:
            out += list(self.clock_model.parameters())
        for m in (self.subst_model, self.site_model):
            if hasattr(m, "parameters"):
                out += list(m.parameters())
        return out

    def close(self) -> None:
        if getattr(self, "fused_allreduce", False):
            # a peer's reduction kernel may still be storing into this rank's receive buffer over
            # NVLink: nobody unmaps or frees before every rank has drained its stream
            import torch.distributed as dist
            torch.cuda.synchronize()
            if dist.is_available() and dist.is_initialized():
                dist.barrier(self.group)
        self.engine.close()

    # ------------------------------------------------------------------- JSON
    @classmethod
    def from_json(cls, data: Dict[str, Any], dic: Dict[str, Any]) -> "TreeLikelihoodModel":
        """Build from the ``"like"`` object of the reference's model JSON
        (/root/reference/flu_H3N2/phylotorch-JC69.template:11-74).

        ``dic`` maps ids to already-built objects (torchtree's registry); sub-objects given
        as strings are looked up there, dict sub-objects are built by the small parsers
        below (newick / Alignment / JC69-HKY-GTR / Constant-Gamma / StrictClock).  Extra
        keys understood: ``"precision"`` ("float64" | "float32"), ``"device"``.
        """
        def resolve(key, builder, required=True):
            spec = data.get(key)
            if spec is None:
                if required:
                    raise KeyError(f"TreeLikelihoodModel JSON needs {key!r}")
                return None
            if isinstance(spec, str):
                if spec not in dic:
                    raise KeyError(f"{key}: id {spec!r} not found")
                return dic[spec]
            obj = builder(spec)
            if isinstance(spec, dict) and "id" in spec:
                dic[spec["id"]] = obj
            return obj

        dev = data.get("device")
        dev = dev if dev is not None else (torch.cuda.current_device() if torch.cuda.is_available() else 0)
        cuda = torch.device(f"cuda:{dev}") if torch.cuda.is_available() else None
        site_pattern = resolve("site_pattern", lambda s: _site_pattern_from_json(s, dic))
        tree_model = resolve("tree_model", lambda s: _tree_model_from_json(s, dic, site_pattern, data.get("device")))
        subst_model = resolve("substitution_model", lambda s: _subst_model_from_json(s, dic))
        site_model = resolve("site_model", lambda s: _site_model_from_json(s, dic), required=False)
        clock_model = resolve("branch_model", lambda s: _clock_from_json(s, dic, cuda), required=False)
        obj = cls(data.get("id"), site_pattern, tree_model, subst_model, site_model, clock_model,
                  precision=data.get("precision", "float64"), device=data.get("device"))
        if "id" in data:
            dic[data["id"]] = obj
        return obj


def _tensor_from_json(spec) -> torch.Tensor:
    if isinstance(spec, dict):  # {"type": "Parameter", "tensor": [...], "full": [n]}
        t = spec.get("tensor")
        if "full" in spec:
            return torch.full(tuple(spec["full"]), float(t), dtype=torch.float64)
        return torch.as_tensor(t, dtype=torch.float64).reshape(-1)
    return torch.as_tensor(spec, dtype=torch.float64).reshape(-1)


def _site_pattern_from_json(spec, dic) -> SitePattern:
    aln = spec["alignment"]
    if isinstance(aln, str):
        aln = dic[aln]
    if isinstance(aln, Alignment):
        return compress_alignment(aln)
    if "file" in aln:
        from .alignment import read_fasta
        return compress_alignment(read_fasta(aln["file"]))
    seqs = aln["sequences"]  # [{"taxon": name, "sequence": "ACGT..."}]
    return compress_alignment(Alignment([s["taxon"] for s in seqs], [s["sequence"] for s in seqs]))


def _parameter_from_json(spec, dic, device=None):
    """``Parameter`` / ``TransformedParameter`` objects of the reference JSON (or an id)."""
    if isinstance(spec, str):
        return dic[spec]
    kind = (spec.get("type") or "Parameter").split(".")[-1] if isinstance(spec, dict) else "Parameter"
    if kind == "TransformedParameter":
        obj = TransformedParameter(spec.get("id"), _parameter_from_json(spec["x"], dic, device), spec["transform"],
                                   spec.get("parameters"))
    else:
        t = _tensor_from_json(spec)
        if device is not None:
            t = t.to(device)
        obj = Parameter(spec.get("id") if isinstance(spec, dict) else None, t)
    if isinstance(spec, dict) and "id" in spec:
        dic[spec["id"]] = obj
    return obj


def _tree_model_from_json(spec, dic, site_pattern: SitePattern, device=None):
    newick = spec.get("newick")
    if newick is None and "file" in spec:
        with open(spec["file"]) as fh:
            newick = fh.read()
    if newick is None:
        raise KeyError("tree_model JSON needs 'newick' or 'file'")
    tree = Tree.from_newick(newick, taxa=list(site_pattern.names))
    if (spec.get("type") or "").split(".")[-1] == "ReparameterizedTimeTreeModel":
        from .timetree import ReparameterizedTimeTreeModel
        dev = device if device is not None else (torch.cuda.current_device() if torch.cuda.is_available() else 0)
        cuda = torch.device(f"cuda:{dev}")
        ratios = _parameter_from_json(spec["ratios"], dic, cuda) if "ratios" in spec else None
        root_height = _parameter_from_json(spec["root_height"], dic, cuda) if "root_height" in spec else None
        return ReparameterizedTimeTreeModel(spec.get("id"), tree, None, ratios, root_height, device=dev)
    bl = None
    if "branch_lengths" in spec:
        bl = _tensor_from_json(spec["branch_lengths"])
    return UnrootedTreeModel(spec.get("id"), tree, bl)


def _values_from_json(spec, dic) -> np.ndarray:
    """Numbers of a model parameter given as a list, a ``Parameter`` object (registered under its
    id, like every nested object of the reference JSON) or the id of one built earlier."""
    if isinstance(spec, (str, dict)):
        if isinstance(spec, str) and spec not in dic:
            raise KeyError(f"parameter id {spec!r} not found")
        return _parameter_from_json(spec, dic).tensor.detach().cpu().numpy().reshape(-1)
    return np.asarray(spec, dtype=np.float64).reshape(-1)


def _object_or_tensor(spec, dic):
    if isinstance(spec, (str, dict)):
        return _parameter_from_json(spec, dic)
    return torch.as_tensor(spec, dtype=torch.float64).reshape(-1)


def _subst_model_from_json(spec, dic=None):
    dic = {} if dic is None else dic
    kind = (spec.get("type") or "").split(".")[-1].upper()
    freqs = spec.get("frequencies")
    if kind in ("GTR", "HKY") and any(isinstance(spec.get(k), (str, dict)) for k in ("rates", "kappa", "frequencies")):
        freqs = None  # handled below: the model keeps the Parameter objects
    freqs = tuple(_values_from_json(freqs, dic)) if freqs is not None else (0.25,) * 4
    if kind == "JC69":
        return JC69()
    # parameters given as Parameter objects (or ids of them): the model keeps the objects, so that
    # whoever holds them (an optimiser, a variational block) changes the model and gets gradients
    is_obj = lambda x: isinstance(x, (str, dict))  # noqa: E731
    if kind == "GTR" and (is_obj(spec.get("rates")) or is_obj(spec.get("frequencies"))):
        return DifferentiableGTR(spec.get("id"), _object_or_tensor(spec.get("rates", [1.0] * 6), dic),
                                 _object_or_tensor(spec.get("frequencies", [0.25] * 4), dic))
    if kind == "HKY" and (is_obj(spec.get("kappa")) or is_obj(spec.get("frequencies"))):
        return DifferentiableHKY(spec.get("id"), _object_or_tensor(spec.get("kappa", [1.0]), dic),
                                 _object_or_tensor(spec.get("frequencies", [0.25] * 4), dic))
    if kind == "HKY":
        return HKY(float(_values_from_json(spec.get("kappa", 1.0), dic)[0]), freqs)
    if kind == "GTR":
        return GTR(tuple(_values_from_json(spec.get("rates", (1.0,) * 6), dic)), freqs)
    raise ValueError(f"unsupported substitution model type {spec.get('type')!r}")


def _site_model_from_json(spec, dic=None):
    dic = {} if dic is None else dic
    kind = (spec.get("type") or "ConstantSiteModel").split(".")[-1]
    if kind == "ConstantSiteModel":
        return ConstantSiteModel()
    if kind == "WeibullSiteModel" and isinstance(spec.get("shape"), (str, dict)):
        return DifferentiableWeibullSiteModel(spec.get("id"), _parameter_from_json(spec["shape"], dic), int(spec.get("categories", 4)))
    if kind == "GammaSiteModel" and isinstance(spec.get("shape"), (str, dict)):
        return DifferentiableGammaSiteModel(spec.get("id"), _parameter_from_json(spec["shape"], dic), int(spec.get("categories", 4)))
    if kind in ("GammaSiteModel", "WeibullSiteModel"):
        shape = float(_values_from_json(spec.get("shape", 0.5), dic)[0])
        if kind == "WeibullSiteModel":
            return WeibullSiteModel(shape, int(spec.get("categories", 4)))
        return GammaSiteModel(shape, int(spec.get("categories", 4)))
    raise ValueError(f"unsupported site model type {spec.get('type')!r}")


def _clock_from_json(spec, dic=None, device=None) -> StrictClockModel:
    dic = {} if dic is None else dic
    kind = (spec.get("type") or "").split(".")[-1]
    if kind != "StrictClockModel":
        raise ValueError(f"unsupported branch model type {spec.get('type')!r}")
    # the rate lives on the engine's device (no per-evaluation host copy) and is registered under
    # its own id: the template's clock.prior and variational blocks refer to "rate"
    return StrictClockModel(spec.get("id"), _parameter_from_json(spec["rate"], dic, device))

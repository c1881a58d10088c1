"""Substitution / site models whose parameters are torch tensors — SURVEY.md §8(f) row 3.

In torchtree the GTR exchangeabilities, the frequencies, HKY's kappa and the site model's
shape are ``Parameter`` objects; the micro benchmark's ``--gtr`` run
(/root/reference/modules/micro.nf:52, ``treelikelihoodGTR`` rows of /root/reference/plot.Rmd:305)
and every variational run that estimates them differentiate the tree likelihood w.r.t. them
by autograd of the whole pruning recursion.  The classes here keep that surface — the model
owns ``Parameter`` objects (``treelikelihood.Parameter``: an id and a tensor), ``parameters()``
lists them — and ``TreeLikelihoodModel`` routes a call that needs their gradient through
``b2p_branch_outer_products`` + ``b2p_model_gradients``: ONE evaluation returns logL, the branch
gradient and the per-branch outer products G, two small kernels contract G with dP/d(Q, rates)
and the root prior, and torch chains the parameterisation (``q_tensor()``, ``rates_tensor()``).

The engine itself consumes an eigen system (``eigen()``, host side, from the current values);
the differentiable rate matrix is only the link between dlogL/dQ and the parameters.
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np
import torch

from .substitution import GTR, EigenSystem

__all__ = ["DifferentiableGTR", "DifferentiableHKY", "DifferentiableWeibullSiteModel", "DifferentiableGammaSiteModel",
           "is_differentiable"]


def _tensor_of(x) -> torch.Tensor:
    return x.tensor if hasattr(x, "tensor") else x


def is_differentiable(model) -> bool:
    """Whether ``model`` carries torch parameters and one of them asks for a gradient."""
    ps = getattr(model, "differentiable_parameters", None)
    return ps is not None and any(_tensor_of(p).requires_grad for p in ps())


class _DifferentiableSubstitutionModel:
    n_states = 4

    def __init__(self, id_: Optional[str], frequencies):
        self.id = id_
        self._frequencies = frequencies

    # torch side -------------------------------------------------------------
    def frequencies_tensor(self) -> torch.Tensor:
        f = _tensor_of(self._frequencies).to(torch.float64)
        return f / f.sum()

    def exchangeabilities_tensor(self) -> torch.Tensor:
        raise NotImplementedError

    def q_tensor(self) -> torch.Tensor:
        """Normalised rate matrix, differentiable w.r.t. the parameters."""
        from .paramgrad import q_matrix
        return q_matrix(self.exchangeabilities_tensor(), self.frequencies_tensor())

    # host side: what Engine.set_model consumes ------------------------------------
    @property
    def freqs(self) -> np.ndarray:
        return self.frequencies_tensor().detach().cpu().numpy()

    @property
    def rates(self) -> np.ndarray:
        return self.exchangeabilities_tensor().detach().cpu().numpy()

    def eigen(self) -> EigenSystem:
        return GTR(tuple(self.rates), tuple(self.freqs)).eigen()

    def q_matrix(self) -> np.ndarray:
        return GTR(tuple(self.rates), tuple(self.freqs)).q_matrix()


class DifferentiableGTR(_DifferentiableSubstitutionModel):
    """GTR with the 6 exchangeabilities (AC, AG, AT, CG, CT, GT) and the 4 frequencies as
    tensors / ``Parameter`` objects (frequencies are renormalised to sum to one)."""

    def __init__(self, id_: Optional[str], rates, frequencies):
        super().__init__(id_, frequencies)
        self._rates = rates
        if _tensor_of(rates).numel() != 6 or _tensor_of(frequencies).numel() != 4:
            raise ValueError("GTR needs 6 exchangeabilities and 4 frequencies")

    def exchangeabilities_tensor(self) -> torch.Tensor:
        return _tensor_of(self._rates).to(torch.float64).reshape(6)

    def differentiable_parameters(self) -> List:
        return [self._rates, self._frequencies]

    def parameters(self) -> List:
        return [p for p in (self._rates, self._frequencies) if hasattr(p, "tensor")]


class DifferentiableHKY(_DifferentiableSubstitutionModel):
    """HKY: transition / transversion ratio kappa and the 4 frequencies."""

    def __init__(self, id_: Optional[str], kappa, frequencies):
        super().__init__(id_, frequencies)
        self._kappa = kappa
        if _tensor_of(kappa).numel() != 1 or _tensor_of(frequencies).numel() != 4:
            raise ValueError("HKY needs one kappa and 4 frequencies")

    @property
    def kappa(self) -> float:
        return float(_tensor_of(self._kappa).detach().reshape(-1)[0])

    def exchangeabilities_tensor(self) -> torch.Tensor:
        from .paramgrad import hky_exchangeabilities
        return hky_exchangeabilities(_tensor_of(self._kappa).to(torch.float64))

    def differentiable_parameters(self) -> List:
        return [self._kappa, self._frequencies]

    def parameters(self) -> List:
        return [p for p in (self._kappa, self._frequencies) if hasattr(p, "tensor")]


class DifferentiableWeibullSiteModel:
    """torchtree's discretised Weibull rate heterogeneity with the shape as a tensor: K
    equal-weight categories at the bin-midpoint quantiles, ``r_k = (-log(1 - (2k+1)/(2K)))^(1/shape)``
    normalised to mean 1 (the same numbers as ``sitemodel.WeibullSiteModel``)."""

    def __init__(self, id_: Optional[str], shape, categories: int = 4):
        self.id = id_
        self._shape = shape
        self.categories = int(categories)
        if _tensor_of(shape).numel() != 1:
            raise ValueError("the Weibull site model needs one shape")
        self.weights = np.full(self.categories, 1.0 / self.categories)

    def rates_tensor(self) -> torch.Tensor:
        a = _tensor_of(self._shape).to(torch.float64).reshape(())
        k = self.categories
        q = (2.0 * torch.arange(k, dtype=torch.float64, device=a.device) + 1.0) / (2.0 * k)
        r = torch.pow(-torch.log1p(-q), 1.0 / a)
        return r / r.mean()

    @property
    def rates(self) -> np.ndarray:
        return self.rates_tensor().detach().cpu().numpy()

    def differentiable_parameters(self) -> List:
        return [self._shape]

    def parameters(self) -> List:
        return [self._shape] if hasattr(self._shape, "tensor") else []


class _GammaRates(torch.autograd.Function):
    """Category rates of Yang's discretised Gamma(alpha, alpha) as a function of the shape.  The
    quantiles have no closed form (that is why autodiff frameworks prefer the Weibull): forward is
    ``sitemodel.GammaSiteModel`` on the host, backward a central difference of it in alpha
    (relative step 1e-6: the rates are smooth in alpha, error ~1e-10 relative)."""

    @staticmethod
    def forward(ctx, alpha, categories):
        from .sitemodel import GammaSiteModel
        a = float(alpha.detach().reshape(-1)[0])
        h = 1e-6 * max(a, 1e-3)
        r = GammaSiteModel(a, categories).rates
        dr = (GammaSiteModel(a + h, categories).rates - GammaSiteModel(a - h, categories).rates) / (2.0 * h)
        ctx.save_for_backward(torch.as_tensor(dr, dtype=torch.float64, device=alpha.device))
        ctx.shape = alpha.shape
        return torch.as_tensor(r, dtype=torch.float64, device=alpha.device)

    @staticmethod
    def backward(ctx, grad_output):
        (dr,) = ctx.saved_tensors
        return (grad_output * dr).sum().reshape(ctx.shape), None


class DifferentiableGammaSiteModel:
    """Discretised Gamma rate heterogeneity (mean-of-bin rates, the numbers of
    ``sitemodel.GammaSiteModel``) with the shape alpha as a tensor / ``Parameter``:
    dlogL/dalpha = sum_k (dlogL/dr_k)(dr_k/dalpha), the first factor from the engine's one pass."""

    def __init__(self, id_: Optional[str], shape, categories: int = 4):
        self.id = id_
        self._shape = shape
        self.categories = int(categories)
        if _tensor_of(shape).numel() != 1:
            raise ValueError("the Gamma site model needs one shape")
        self.weights = np.full(self.categories, 1.0 / self.categories)

    def rates_tensor(self) -> torch.Tensor:
        return _GammaRates.apply(_tensor_of(self._shape).to(torch.float64), self.categories)

    @property
    def rates(self) -> np.ndarray:
        from .sitemodel import GammaSiteModel
        return GammaSiteModel(float(_tensor_of(self._shape).detach().reshape(-1)[0]), self.categories).rates

    def differentiable_parameters(self) -> List:
        return [self._shape]

    def parameters(self) -> List:
        return [self._shape] if hasattr(self._shape, "tensor") else []

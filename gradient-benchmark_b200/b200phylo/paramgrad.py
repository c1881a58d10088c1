"""Gradients of the tree log-likelihood w.r.t. substitution-model and site-model parameters
(SURVEY.md §8(f) row 3): GTR exchangeabilities and frequencies, HKY's kappa, category rates
(Weibull shape directly, Gamma shape through d rates / d alpha).

The micro benchmark's ``--gtr`` run differentiates through these in torchtree by autograd
of the whole pruning recursion (/root/reference/modules/micro.nf:52,
/root/reference/plot.Rmd:305); physher does one analytic pass per parameter.  Here they come
out of the SAME post-/pre-order kernel that produces the branch gradient.  For every branch
b and category k the pre-order pass holds q (everything outside the branch) and m = P x
(the branch's message); the per-pattern likelihood is bilinear in them, so

    dlogL/dtheta = sum_{b,k} < G_{b,k} , dP_{b,k}/dtheta P_{b,k}^{-1} >  +  root-prior term,
    G_{b,k} = sum_p (w_p / L_p) c_k q m^T                                  (4x4 per branch).

``b2p_branch_outer_products`` returns all of G from ONE evaluation (the pre-order pass sums the
outer products next to the branch gradient: 4 states, fp64, 1 or 4 categories).  Elsewhere
``b2p_set_gradient_operator`` makes the kernel's "gradient" output one entry of G (operator
E_ij, one-hot category scale): 16 K evaluations fill G for all branches, whatever the number
of parameters.  The rest is 4x4 algebra on a few thousand matrices, done here with torch so
that ANY differentiable parameterisation of Q, pi and the rates gets its gradient from one
``backward()`` of a surrogate  S(theta) = sum <G P^{-T}, P(theta)> + sum_i c_i pi_i(theta)
whose gradient equals dlogL/dtheta at the current parameters.  The inverse transition
matrices limit this to r t |lambda|_max of a few tens (any realistic branch).  No CPU
fallback: G comes from the engine.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional, Sequence

import numpy as np
import torch

from .engine import Engine

__all__ = ["branch_outer_products", "q_matrix", "gtr_exchangeabilities", "hky_exchangeabilities",
           "parameter_gradients", "surrogate"]


def branch_outer_products(engine: Engine, bl: torch.Tensor, one_pass: Optional[bool] = None) -> torch.Tensor:
    """G [K, 2N-2, 4, 4] on the device.

    One evaluation (``b2p_branch_outer_products``: the pre-order pass sums the outer products
    itself) where the engine supports it — 4 states, fp64, 1 or 4 rate categories — else 16 K
    evaluations with operators E_ij / one-hot scales (``one_pass=False`` forces that route)."""
    if not bl.is_cuda:
        raise ValueError("branch lengths must live on the engine's device")
    k_cat, n_br = engine.n_categories, engine.n_branches
    bl = bl.detach().to(torch.float64).contiguous()
    stream = torch.cuda.current_stream(bl.device).cuda_stream
    if one_pass is None:
        one_pass = engine.supports_outer_products()
    if one_pass:
        g = torch.empty(k_cat, n_br, 4, 4, dtype=torch.float64, device=bl.device)
        logl = torch.empty(1, dtype=torch.float64, device=bl.device)
        engine.branch_outer_products(bl.data_ptr(), logl.data_ptr(), None, g.data_ptr(), stream)
        return g
    out = torch.empty(k_cat, 4, 4, 1 + n_br, dtype=torch.float64, device=bl.device)
    op = np.zeros((4, 4))
    try:
        for k in range(k_cat):
            scale = np.zeros(k_cat)
            scale[k] = 1.0
            for i in range(4):
                for j in range(4):
                    op[:] = 0.0
                    op[i, j] = 1.0
                    engine.set_gradient_operator(op, scale)
                    o = out[k, i, j]
                    engine.loglik_grad_device(bl.data_ptr(), o.data_ptr(), o.data_ptr() + 8, stream)
    finally:
        engine.set_gradient_operator(None)
    return out[..., 1:].permute(0, 3, 1, 2).contiguous()


def q_matrix(exch: torch.Tensor, freqs: torch.Tensor) -> torch.Tensor:
    """Normalised reversible rate matrix from the 6 exchangeabilities (AC, AG, AT, CG, CT, GT)
    and frequencies — the torch twin of ``substitution.SubstitutionModel.q_matrix``."""
    iu = torch.triu_indices(4, 4, 1, device=exch.device)
    r = torch.zeros(4, 4, dtype=exch.dtype, device=exch.device)
    r = r.index_put((iu[0], iu[1]), exch)
    r = r + r.T
    q = r * freqs[None, :]
    q = q - torch.diag(torch.diag(q))
    q = q - torch.diag(q.sum(dim=1))
    norm = -(freqs * torch.diag(q)).sum()
    return q / norm


def gtr_exchangeabilities(rates: torch.Tensor) -> torch.Tensor:
    return rates


def hky_exchangeabilities(kappa: torch.Tensor) -> torch.Tensor:
    one = torch.ones((), dtype=kappa.dtype, device=kappa.device)
    k = kappa.reshape(())
    return torch.stack([one, k, one, one, k, one])


def surrogate(g: torch.Tensor, t: torch.Tensor, q: torch.Tensor, freqs: torch.Tensor, rates: torch.Tensor,
              root_child: int) -> torch.Tensor:
    """S(q, freqs, rates, t) = sum <G P^{-T}, P(q, rates, t)> + sum_i c_i pi_i with G held constant:
    its gradient at the parameters G was computed at is dlogL/d(q, freqs, rates) (and, through P,
    the branch gradient).  ``g`` [K, 2N-2, 4, 4], ``t`` [2N-2], ``q`` [4, 4], ``rates`` [K]."""
    # P_{k,b} = exp(Q r_k t_b), differentiable; its inverse only as a constant
    qt = q[None, None] * (rates[:, None] * t[None, :])[..., None, None]
    p = torch.linalg.matrix_exp(qt)
    with torch.no_grad():
        p_inv_t = torch.linalg.matrix_exp(-qt).transpose(-1, -2)
        w = g @ p_inv_t                       # G P^{-T}
        c = torch.diagonal(g[:, int(root_child)], dim1=-2, dim2=-1).sum(0) / freqs   # root prior: dlogL/dpi_i
    return (w * p).sum() + (c * freqs).sum()


def parameter_gradients(engine: Engine, bl: torch.Tensor, root_children: Sequence[int],
                        model: Callable[[], Dict[str, torch.Tensor]], g: Optional[torch.Tensor] = None,
                        algebra: Optional[str] = None) -> torch.Tensor:
    """Surrogate whose ``backward()`` leaves dlogL/dtheta in the ``.grad`` of every leaf tensor
    ``model()`` depends on.

    ``model()`` returns the CURRENT model as differentiable torch tensors on the engine's
    device: ``{"q": [4,4] rate matrix, "freqs": [4], "rates": [K] category rates}`` (and may
    also depend on ``bl``, which then receives the ordinary branch gradient).  The engine must
    already hold the same model (``Engine.set_model``).  Returns S (call ``.backward()``).

    ``algebra``: "device" — ``b2p_model_gradients`` contracts G with dP/d(Q, rates) in two small
    launches and S is linear in (q, freqs, rates, bl); "torch" — S = sum <G P^{-T}, P(theta)> with
    torch's matrix exponentials (the round-1 route; the only one when the engine cannot return
    G in one pass).  Default: "device" where available and ``g`` is not given.
    """
    if algebra is None:
        algebra = "device" if (g is None and engine.supports_outer_products()) else "torch"
    m = model()
    t = bl if bl.requires_grad else bl.detach()
    if algebra == "device":
        if g is not None:
            raise ValueError("algebra='device' computes the outer products itself")
        k_cat, n_br = engine.n_categories, engine.n_branches
        b = bl.detach().to(torch.float64).contiguous()
        stream = torch.cuda.current_stream(b.device).cuda_stream
        g = torch.empty(k_cat, n_br, 4, 4, dtype=torch.float64, device=b.device)
        out = torch.empty(1 + n_br, dtype=torch.float64, device=b.device)
        mg = torch.empty(20 + k_cat, dtype=torch.float64, device=b.device)
        engine.branch_outer_products(b.data_ptr(), out.data_ptr(), out.data_ptr() + 8, g.data_ptr(), stream)
        engine.model_gradients(b.data_ptr(), g.data_ptr(), int(root_children[0]), mg.data_ptr(), stream)
        s = (mg[:16].reshape(4, 4) * m["q"]).sum() + (mg[16:20] * m["freqs"]).sum() + (mg[20:] * m["rates"]).sum()
        return s + (out[1:] * t).sum() if t.requires_grad else s
    if g is None:
        g = branch_outer_products(engine, bl)
    return surrogate(g, t, m["q"], m["freqs"], m["rates"], root_children[0])

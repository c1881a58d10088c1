"""CPU ORACLE — test infrastructure only; never on the product path.

PARITY UNPINNED: the reference (/root/reference) contains no likelihood code, no
golden vectors and no expected values for this path.  The
arithmetic it times lives in third-party programs that are not vendored and not
installable here: physher@b19ff2f, torchtree@f383165, torchtree-bito@e2a95ce,
bito@cc0806a + BEAGLE (/root/reference/README.md:65-73, /root/reference/Dockerfile:25-68).
This file restates their published algorithm (Felsenstein 1981 pruning; Yang 1994
discrete Gamma; Ji et al. 2020 linear-time pre-order branch gradient — the algorithm
behind ``beagleUpdatePartials`` / ``beagleUpdatePrePartials`` /
``beagleCalculateEdgeDerivatives`` and physher's lower/upper partials) in three
independent formulations that the tests require to agree:

  (A) ``loglik_autograd``     torchtree-style pruning as batched mat-muls + autograd
                              (call site: /root/reference/modules/micro.nf:47-53)
  (B) ``loglik_and_gradient`` analytic post-order + pre-order, physher/BEAGLE-style
                              (call sites: /root/reference/modules/micro.nf:20-24,
                              /root/reference/modules/macro_flu.nf:104,119-122)
  (C) ``finite_difference``   central differences (small trees only)

Model semantics follow /root/reference/flu_H3N2/phylotorch-JC69.template:11-74 and
/root/reference/flu_H3N2/physher-JC69.template:6-57; inputs/outputs follow the
micro-benchmark description in /root/reference/plot.Rmd:41-60 (branch lengths in,
log-likelihood and d/d branch lengths out).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.
"""
from __future__ import annotations

import numpy as np

__all__ = [
    "transition_matrices", "q_from_eigen", "post_order", "log_likelihood",
    "loglik_and_gradient", "loglik_autograd", "finite_difference", "tip_partials_from_masks",
]


def tip_partials_from_masks(masks: np.ndarray, n_states: int = 4) -> np.ndarray:
    """uint8 [N, P] IUPAC bit masks → float64 [N, P, S] 0/1 tip partials."""
    bits = (masks[..., None] >> np.arange(n_states, dtype=np.uint8)) & 1
    return bits.astype(np.float64)


def q_from_eigen(evec, ivec, evals) -> np.ndarray:
    return (evec * evals[None, :]) @ ivec


def transition_matrices(bl, evec, ivec, evals, rates) -> np.ndarray:
    """P[b, k] = V diag(exp(lambda * r_k * t_b)) V^-1  → [B, K, S, S]  (row a2)."""
    bl = np.asarray(bl, dtype=np.float64)
    rates = np.asarray(rates, dtype=np.float64)
    ex = np.exp(bl[:, None, None] * rates[None, :, None] * evals[None, None, :])
    return np.einsum("ij,bkj,jl->bkil", evec, ex, ivec)


def post_order(triples, tip_partials, pmats):
    """Felsenstein pruning with per-pattern rescaling at every node (row a3).

    Returns (post [2N-1, K, P, S] scaled, logscale [2N-1, P] cumulative per subtree).
    The scale factor is shared across states *and* categories (SURVEY hard part 2).
    """
    n, p, s = tip_partials.shape
    k = pmats.shape[1]
    post = np.empty((2 * n - 1, k, p, s))
    post[:n] = tip_partials[:, None, :, :]
    logscale = np.zeros((2 * n - 1, p))
    for v, l, r in np.asarray(triples):
        ml = np.einsum("kij,kpj->kpi", pmats[l], post[l])
        mr = np.einsum("kij,kpj->kpi", pmats[r], post[r])
        x = ml * mr
        c = x.max(axis=(0, 2))
        c = np.where(c > 0, c, 1.0)
        post[v] = x / c[None, :, None]
        logscale[v] = logscale[l] + logscale[r] + np.log(c)
    return post, logscale


def _site_likelihoods(post_root, freqs, cat_weights):
    return np.einsum("k,kpi,i->p", cat_weights, post_root, freqs)


def log_likelihood(triples, tip_partials, weights, bl, evec, ivec, evals, freqs,
                   rates, cat_weights) -> float:
    """Rows a3 + a4: logL = sum_p w_p (log L_p + logscale_p)."""
    pm = transition_matrices(bl, evec, ivec, evals, rates)
    post, logscale = post_order(triples, tip_partials, pm)
    root = int(np.asarray(triples)[-1, 0])
    lp = _site_likelihoods(post[root], freqs, np.asarray(cat_weights))
    with np.errstate(divide="ignore"):
        return float(np.sum(weights * (np.log(lp) + logscale[root])))


def loglik_and_gradient(triples, tip_partials, weights, bl, evec, ivec, evals, freqs,
                        rates, cat_weights):
    """Rows a3-a6: analytic two-pass gradient (formulation B).

    pre_root = pi;  for parent v, child c, sibling s:
        pre_c = P_c^T (pre_v ∘ (P_s post_s))
        g_c   = sum_p w_p [sum_k w_k r_k pre_c·(Q post_c)] / [sum_k w_k pre_c·post_c]
    Pre partials are renormalised per pattern at every node; the ratio is scale-free
    because numerator and denominator use the same pre_c and post_c.
    """
    triples = np.asarray(triples)
    rates = np.asarray(rates, dtype=np.float64)
    cw = np.asarray(cat_weights, dtype=np.float64)
    pm = transition_matrices(bl, evec, ivec, evals, rates)
    post, logscale = post_order(triples, tip_partials, pm)
    n, p, s = tip_partials.shape
    k = len(rates)
    root = int(triples[-1, 0])
    lp = _site_likelihoods(post[root], freqs, cw)
    with np.errstate(divide="ignore"):
        logl = float(np.sum(weights * (np.log(lp) + logscale[root])))
    q = q_from_eigen(evec, ivec, evals)
    pre = np.empty((2 * n - 1, k, p, s))
    pre[root] = np.broadcast_to(freqs, (k, p, s))
    grad = np.zeros(2 * n - 2)
    for v, l, r in triples[::-1]:
        for c, sib in ((l, r), (r, l)):
            msib = np.einsum("kij,kpj->kpi", pm[sib], post[sib])
            qc = pre[v] * msib
            pc = np.einsum("kij,kpi->kpj", pm[c], qc)
            mx = pc.max(axis=(0, 2))
            mx = np.where(mx > 0, mx, 1.0)
            pc = pc / mx[None, :, None]
            pre[c] = pc
            qpost = np.einsum("ij,kpj->kpi", q, post[c])
            num = np.einsum("k,kpi,kpi->p", cw * rates, pc, qpost)
            den = np.einsum("k,kpi,kpi->p", cw, pc, post[c])
            grad[c] = np.sum(weights * num / den)
    return logl, grad


def branch_outer_products(triples, tip_partials, weights, bl, evec, ivec, evals, freqs,
                          rates, cat_weights) -> np.ndarray:
    """G [K, 2N-2, S, S]:  G_{c,k} = sum_p (w_p / L_p) c_k q_c m_c^T  with  q_c = pre_v o (P_s post_s)
    (everything outside branch c) and m_c = P_c post_c (the branch's message) — the per-branch sums
    every gradient w.r.t. the substitution / site model is a contraction of (SURVEY.md §8(f) row 3;
    b2p_branch_outer_products computes them on the GPU).  Same two passes as loglik_and_gradient;
    the ratio is scale-free for the same reason.  Identities the tests use:
    sum_k r_k <Q, G_{c,k}> = dlogL/dt_c  and  sum_{k,i} G_{c,k}[i][i] = sum_p w_p for every c."""
    triples = np.asarray(triples)
    rates = np.asarray(rates, dtype=np.float64)
    cw = np.asarray(cat_weights, dtype=np.float64)
    pm = transition_matrices(bl, evec, ivec, evals, rates)
    post, _ = post_order(triples, tip_partials, pm)
    n, p, s = tip_partials.shape
    k = len(rates)
    root = int(triples[-1, 0])
    pre = np.empty((2 * n - 1, k, p, s))
    pre[root] = np.broadcast_to(freqs, (k, p, s))
    g = np.zeros((k, 2 * n - 2, s, s))
    for v, l, r in triples[::-1]:
        for c, sib in ((l, r), (r, l)):
            msib = np.einsum("kij,kpj->kpi", pm[sib], post[sib])
            qc = pre[v] * msib
            mc = np.einsum("kij,kpj->kpi", pm[c], post[c])
            den = np.einsum("k,kpi,kpi->p", cw, qc, mc)
            with np.errstate(divide="ignore", invalid="ignore"):
                a = np.where(den > 0, weights / den, 0.0)
            g[:, c] = np.einsum("k,p,kpi,kpj->kij", cw, a, qc, mc)
            pc = np.einsum("kij,kpi->kpj", pm[c], qc)
            mx = pc.max(axis=(0, 2))
            mx = np.where(mx > 0, mx, 1.0)
            pre[c] = pc / mx[None, :, None]
    return g


def finite_difference(fn, bl, eps: float = 1e-6) -> np.ndarray:
    """Central differences of ``fn(bl) -> float`` (formulation C)."""
    bl = np.asarray(bl, dtype=np.float64)
    g = np.zeros_like(bl)
    for i in range(bl.size):
        h = eps * max(1.0, abs(bl[i]))
        up = bl.copy(); up[i] += h
        dn = bl.copy(); dn[i] = max(dn[i] - h, 0.0)
        g[i] = (fn(up) - fn(dn)) / (up[i] - dn[i])
    return g


def loglik_autograd(triples, tip_partials, weights, bl, evec, ivec, evals, freqs,
                    rates, cat_weights):
    """Formulation A: pruning as batched mat-muls, gradient by reverse-mode AD."""
    import torch

    t = lambda a: torch.as_tensor(np.asarray(a), dtype=torch.float64)
    blt = t(bl).clone().requires_grad_(True)
    ev, iv, lam, fr, rt, cw, w = map(t, (evec, ivec, evals, freqs, rates, cat_weights, weights))
    ex = torch.exp(blt[:, None, None] * rt[None, :, None] * lam[None, None, :])
    pm = torch.einsum("ij,bkj,jl->bkil", ev, ex, iv)
    tips = t(tip_partials)
    n, p, s = tip_partials.shape
    k = rt.shape[0]
    part = [tips[i].T.unsqueeze(0).expand(k, s, p) for i in range(n)] + [None] * (n - 1)
    logscale = [torch.zeros(p, dtype=torch.float64) for _ in range(2 * n - 1)]
    for v, l, r in np.asarray(triples):
        x = (pm[l] @ part[l]) * (pm[r] @ part[r])
        c = x.detach().amax(dim=(0, 1)).clamp_min(1e-300)
        part[v] = x / c
        logscale[v] = logscale[l] + logscale[r] + torch.log(c)
    root = int(np.asarray(triples)[-1, 0])
    lp = torch.einsum("k,kip,i->p", cw, part[root], fr)
    logl = torch.sum(w * (torch.log(lp) + logscale[root]))
    (g,) = torch.autograd.grad(logl, blt)
    return float(logl.detach()), g.numpy()

"""CPU restatement (torch, float64) of the tree log-likelihood as a differentiable function of
the substitution-model parameters, branch lengths and category rates — TEST INFRASTRUCTURE
ONLY.  This is the formulation torchtree itself uses for these gradients: Felsenstein pruning
written with tensor ops, ``logL.backward()`` (torchtree
``evolution/tree_likelihood.py``, pinned f3831650a807e74cc2e9478009e57a41f47bed8d,
/root/reference/README.md:71; upstream recall — PARITY UNPINNED, the reference pins no values).
Transition matrices come from ``torch.linalg.matrix_exp`` (no eigendecomposition), so the
gradients are independent of the engine's eigen route."""
import numpy as np
import torch


def log_likelihood(triples, tip_partials, weights, bl, q, freqs, rates, cat_w):
    """tip_partials [N, P, 4] (0/1), bl [2N-2], q [4,4], freqs [4], rates [K], cat_w [K] -> scalar."""
    n = tip_partials.shape[0]
    p_mats = torch.linalg.matrix_exp(q[None, None] * (rates[:, None] * bl[None, :])[..., None, None])  # [K, B, 4, 4]
    part = [None] * (2 * n - 1)
    for i in range(n):
        part[i] = torch.as_tensor(tip_partials[i], dtype=torch.float64)[None].expand(rates.shape[0], -1, -1)  # [K, P, 4]
    logscale = torch.zeros(tip_partials.shape[1], dtype=torch.float64)
    root = int(triples[-1][0])
    for v, l, r in triples:
        a = torch.einsum("kij,kpj->kpi", p_mats[:, int(l)], part[int(l)])
        b = torch.einsum("kij,kpj->kpi", p_mats[:, int(r)], part[int(r)])
        x = a * b
        s = x.detach().amax(dim=(0, 2)).clamp_min(1e-300)
        part[int(v)] = x / s[None, :, None]
        logscale = logscale + torch.log(s)
    site = torch.einsum("k,kpi,i->p", torch.as_tensor(cat_w, dtype=torch.float64), part[root], freqs)
    return (torch.as_tensor(weights, dtype=torch.float64) * (torch.log(site) + logscale)).sum()

"""The time-tree oracle (oracle/timetree_oracle.py) against independent restatements:
torch autograd of a literal forward pass, finite differences, and hand-worked cases."""
import numpy as np
import pytest
import torch

from b200phylo.tree import simulate_tree
from oracle import timetree_oracle as tto


def dated_case(n, shape, seed, contemporaneous=False):
    tree = simulate_tree(n, shape=shape, seed=seed)
    rng = np.random.default_rng(seed + 1)
    tips = np.zeros(n) if contemporaneous else rng.uniform(0.0, 12.0, n)
    if not contemporaneous:
        tips[rng.integers(n)] = 0.0
    params = rng.uniform(0.05, 0.95, n - 1)
    params[n - 2] = tips.max() + rng.uniform(0.5, 5.0)
    return tree, tips, params


def torch_heights(n, triples, tips, params):
    """Literal forward restatement with torch scalars (autograd gives the reference gradients)."""
    par = {}
    for v, l, r in triples:
        par[int(l)] = int(v)
        par[int(r)] = int(v)
    b = tto.bounds(n, triples, tips)
    h = [None] * (n - 1)
    h[n - 2] = params[n - 2]
    for v in range(2 * n - 3, n - 1, -1):
        h[v - n] = b[v] + params[v - n] * (h[par[v] - n] - b[v])
    return torch.stack(h), par, b


@pytest.mark.parametrize("n,shape,seed", [(2, "balanced", 1), (3, "caterpillar", 2), (17, "coalescent", 3),
                                         (64, "balanced", 4), (50, "caterpillar", 5), (200, "coalescent", 6)])
def test_heights_and_vjp_against_autograd(n, shape, seed):
    tree, tips, params = dated_case(n, shape, seed)
    tr = tree.post_triples
    x = torch.tensor(params, dtype=torch.float64, requires_grad=True)
    h, par, b = torch_heights(n, tr, tips, x)
    got = tto.heights_from_ratios(n, tr, tips, params)
    np.testing.assert_allclose(got, h.detach().numpy(), rtol=0, atol=0)
    # heights respect the bounds and the parent
    for v in range(n, 2 * n - 2):
        assert b[v] <= got[v - n] <= got[par[v] - n]
    g = np.random.default_rng(seed).normal(size=n - 1)
    (h * torch.tensor(g)).sum().backward()
    np.testing.assert_allclose(tto.heights_vjp(n, tr, tips, params, g), x.grad.numpy(), rtol=1e-12, atol=1e-13)
    # inverse map (deep chains lose the ratios to rounding, the heights they give back do not)
    back = tto.ratios_from_heights(n, tr, tips, got)
    np.testing.assert_allclose(tto.heights_from_ratios(n, tr, tips, back), got, rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("n,shape,seed", [(3, "caterpillar", 2), (33, "coalescent", 7), (30, "caterpillar", 8), (120, "balanced", 14)])
def test_log_jacobian_is_the_log_determinant(n, shape, seed):
    tree, tips, params = dated_case(n, shape, seed)
    tr = tree.post_triples
    x = torch.tensor(params, dtype=torch.float64)
    jac = torch.autograd.functional.jacobian(lambda p: torch_heights(n, tr, tips, p)[0], x)
    sign, logdet = np.linalg.slogdet(jac.numpy())
    h = tto.heights_from_ratios(n, tr, tips, params)
    assert sign > 0
    assert tto.log_jacobian(n, tr, tips, h) == pytest.approx(logdet, rel=1e-11, abs=1e-11)
    # gradient w.r.t. heights: autograd over the same sum
    ht = torch.tensor(h, dtype=torch.float64, requires_grad=True)
    _, par, b = torch_heights(n, tr, tips, x)
    sum(torch.log(ht[par[v] - n] - b[v]) for v in range(n, 2 * n - 2)).backward()
    np.testing.assert_allclose(tto.log_jacobian_grad(n, tr, tips, h), ht.grad.numpy(), rtol=1e-13, atol=0)


@pytest.mark.parametrize("n,shape,seed", [(2, "balanced", 1), (40, "coalescent", 9), (90, "caterpillar", 10)])
def test_branch_lengths_and_vjp(n, shape, seed):
    tree, tips, params = dated_case(n, shape, seed)
    tr = tree.post_triples
    h = tto.heights_from_ratios(n, tr, tips, params)
    rate = 0.0031
    bl = tto.branch_lengths(n, tr, tips, h, rate)
    assert bl.shape == (2 * n - 2,) and (bl >= 0).all()
    # every root-to-tip path spans root height - tip height
    par = tto._parent(n, tr)
    for tip in range(0, n, max(1, n // 7)):
        s, v = 0.0, tip
        while par[v] >= 0:
            s += bl[v]
            v = par[v]
        assert s == pytest.approx(rate * (h[n - 2] - tips[tip]), rel=1e-12)
    g = np.random.default_rng(seed).normal(size=2 * n - 2)
    ht = torch.tensor(h, requires_grad=True)
    rt = torch.tensor(rate, dtype=torch.float64, requires_grad=True)
    allh = torch.cat([torch.tensor(tips), ht])
    blt = rt * (allh[torch.tensor(par[: 2 * n - 2])] - allh[: 2 * n - 2])
    np.testing.assert_allclose(bl, blt.detach().numpy(), rtol=1e-15)
    (blt * torch.tensor(g)).sum().backward()
    gh, gr = tto.branch_lengths_vjp(n, tr, tips, h, rate, g)
    np.testing.assert_allclose(gh, ht.grad.numpy(), rtol=1e-12, atol=1e-15)
    assert gr == pytest.approx(float(rt.grad), rel=1e-12)


def test_coalescent_hand_case():
    # three contemporaneous tips, coalescences at 1 and 3, theta = 2:
    # 3 lineages for 1 unit, 2 lineages for 2 units: -(3*1 + 1*2)/2 - 2 log 2
    logp, gh, gt = tto.constant_coalescent(3, np.zeros(3), np.array([1.0, 3.0]), 2.0)
    assert logp == pytest.approx(-2.5 - 2 * np.log(2.0), rel=1e-15)
    np.testing.assert_allclose(gh, [-(3 - 1) / 2.0, -(2 - 1) / 2.0])
    assert gt == pytest.approx(5.0 / 4.0 - 2 / 2.0)


@pytest.mark.parametrize("n,shape,seed,contemp", [(12, "coalescent", 11, True), (60, "coalescent", 12, False),
                                                  (150, "caterpillar", 13, False)])
def test_coalescent_gradient_finite_differences(n, shape, seed, contemp):
    tree, tips, params = dated_case(n, shape, seed, contemp)
    tr = tree.post_triples
    h = tto.heights_from_ratios(n, tr, tips, params)
    theta = 3.7
    logp, gh, gt = tto.constant_coalescent(n, tips, h, theta)
    # torch restatement with the event order frozen
    allh = np.concatenate([tips, h])
    order = np.argsort(allh, kind="stable")
    ht = torch.tensor(h, requires_grad=True)
    th = torch.tensor(theta, dtype=torch.float64, requires_grad=True)
    ah = torch.cat([torch.tensor(tips), ht])[torch.tensor(order)]
    k = np.cumsum(np.where(order >= n, -1, 1))[:-1]
    lp = -(torch.tensor(0.5 * k * (k - 1)) * (ah[1:] - ah[:-1])).sum() / th - (n - 1) * torch.log(th)
    assert logp == pytest.approx(float(lp.detach()), rel=1e-13)
    lp.backward()
    np.testing.assert_allclose(gh, ht.grad.numpy(), rtol=1e-12, atol=1e-14)
    assert gt == pytest.approx(float(th.grad), rel=1e-12)
    eps = 1e-6
    fd = (tto.constant_coalescent(n, tips, h, theta + eps)[0] - tto.constant_coalescent(n, tips, h, theta - eps)[0]) / (2 * eps)
    assert gt == pytest.approx(fd, rel=1e-6)

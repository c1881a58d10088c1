"""CPU restatement of the time-tree functions the micro benchmark times next to the tree
likelihood — TEST INFRASTRUCTURE ONLY (tests/, __graft_entry__.smoke(), bench.py's
cpu_baseline leg).  Nothing in the product path imports this file.

What it restates.  /root/reference holds no implementation of these functions either; it
times them in external programs and defines them in prose:

  * ratio_transform            "input: ratios/root height, output: internal node heights,
                                derivatives wrt ratios"          /root/reference/plot.Rmd:53-56
  * ratio_transform_jacobian   "input: internal node heights, output: log determinant of
                                the Jacobian, derivatives wrt internal node heights"
                                                                 /root/reference/plot.Rmd:57-60
  * coalescent                 "input: node heights and pop size, output: log likelihood,
                                derivatives wrt constant pop size and node heights"
                                                                 /root/reference/plot.Rmd:49-52
  * the model they feed: ReparameterizedTimeTreeModel (ratios through a sigmoid, root height
    through an affine shift) + StrictClockModel + ConstantCoalescentModel
                                 /root/reference/flu_H3N2/phylotorch-JC69.template:14-58,75-100

The algorithms live in torchtree (pinned f3831650a807e74cc2e9478009e57a41f47bed8d,
/root/reference/README.md:71, not vendored, not installable here): its
GeneralNodeHeightTransform and ConstantCoalescent.  Their published definitions, restated:

  node heights   h_root = root height;  for every other internal node i with parent p:
                 h_i = b_i + r_i (h_p - b_i),  b_i = the largest sampling height among the
                 tips below i (a node cannot be younger than its descendants).
  log|det J|     the map (ratios, root height) -> heights is triangular in pre-order with
                 diagonal dh_i/dr_i = h_p - b_i, so log|det J| = sum_i log(h_p(i) - b_i).
  branch lengths t_c = h_parent(c) - h_c in time; substitutions = rate * t_c (strict clock).
  coalescent     Kingman's constant-size density of the sorted node heights: with k lineages
                 over an interval of length d, -C(k,2) d / theta; every coalescent event adds
                 -log(theta).

PARITY UNPINNED: the reference carries no golden vectors for these functions and torchtree
cannot be imported on this box.  The restatement is pinned only to itself: analytic
gradients here are checked against torch autograd of a literal forward restatement and
against finite differences (tests/test_timetree_oracle.py).

Conventions (same as include/b200phylo.h): tips 0..N-1, internal nodes N..2N-2 with
children before parents, root = 2N-2.  ``params`` has length N-1: ratios of nodes N..2N-3,
then the root height.  ``heights`` are the N-1 internal heights in node order.
"""
from __future__ import annotations

import numpy as np

__all__ = ["bounds", "heights_from_ratios", "heights_vjp", "ratios_from_heights", "log_jacobian",
           "log_jacobian_grad", "branch_lengths", "branch_lengths_vjp", "constant_coalescent"]


def _parent(n_tips: int, triples: np.ndarray) -> np.ndarray:
    par = np.full(2 * n_tips - 1, -1, dtype=np.int64)
    for v, l, r in triples:
        par[l] = v
        par[r] = v
    return par


def bounds(n_tips: int, triples: np.ndarray, tip_heights: np.ndarray) -> np.ndarray:
    """b_v for every node (tips: their own sampling height)."""
    b = np.zeros(2 * n_tips - 1)
    b[:n_tips] = tip_heights
    for v, l, r in triples:  # children before parents
        b[v] = max(b[l], b[r])
    return b


def heights_from_ratios(n_tips, triples, tip_heights, params) -> np.ndarray:
    n = n_tips
    par = _parent(n, triples)
    b = bounds(n, triples, tip_heights)
    h = np.empty(n - 1)
    h[n - 2] = params[n - 2]
    for v in range(2 * n - 3, n - 1, -1):  # parents before children
        i = v - n
        h[i] = b[v] + params[i] * (h[par[v] - n] - b[v])
    return h


def ratios_from_heights(n_tips, triples, tip_heights, heights) -> np.ndarray:
    """Inverse map (used to initialise ratios from a dated tree, `--heights_init tree`,
    /root/reference/modules/macro_flu.nf:124-131)."""
    n = n_tips
    par = _parent(n, triples)
    b = bounds(n, triples, tip_heights)
    x = np.empty(n - 1)
    x[n - 2] = heights[n - 2]
    for v in range(n, 2 * n - 2):
        x[v - n] = (heights[v - n] - b[v]) / (heights[par[v] - n] - b[v])
    return x


def heights_vjp(n_tips, triples, tip_heights, params, grad_heights) -> np.ndarray:
    """J^T g: gradient w.r.t. (ratios, root height) of a function with dF/dh = g."""
    n = n_tips
    par = _parent(n, triples)
    b = bounds(n, triples, tip_heights)
    h = heights_from_ratios(n, triples, tip_heights, params)
    a = np.array(grad_heights, dtype=np.float64)  # total derivative w.r.t. each height
    out = np.empty(n - 1)
    for v in range(n, 2 * n - 2):  # children before parents
        i, p = v - n, par[v] - n
        out[i] = a[i] * (h[p] - b[v])
        a[p] += params[i] * a[i]
    out[n - 2] = a[n - 2]
    return out


def log_jacobian(n_tips, triples, tip_heights, heights) -> float:
    n = n_tips
    par = _parent(n, triples)
    b = bounds(n, triples, tip_heights)
    return float(sum(np.log(heights[par[v] - n] - b[v]) for v in range(n, 2 * n - 2)))


def log_jacobian_grad(n_tips, triples, tip_heights, heights) -> np.ndarray:
    n = n_tips
    par = _parent(n, triples)
    b = bounds(n, triples, tip_heights)
    g = np.zeros(n - 1)
    for v in range(n, 2 * n - 2):
        g[par[v] - n] += 1.0 / (heights[par[v] - n] - b[v])
    return g


def branch_lengths(n_tips, triples, tip_heights, heights, rate: float = 1.0) -> np.ndarray:
    n = n_tips
    par = _parent(n, triples)
    allh = np.concatenate([np.asarray(tip_heights, dtype=np.float64), heights])
    return np.array([rate * (allh[par[c]] - allh[c]) for c in range(2 * n - 2)])


def branch_lengths_vjp(n_tips, triples, tip_heights, heights, rate, grad_bl):
    """(gradient w.r.t. internal heights, gradient w.r.t. the rate)."""
    n = n_tips
    par = _parent(n, triples)
    allh = np.concatenate([np.asarray(tip_heights, dtype=np.float64), heights])
    gh = np.zeros(n - 1)
    grate = 0.0
    for c in range(2 * n - 2):
        gh[par[c] - n] += rate * grad_bl[c]
        if c >= n:
            gh[c - n] -= rate * grad_bl[c]
        grate += grad_bl[c] * (allh[par[c]] - allh[c])
    return gh, grate


def constant_coalescent(n_tips, tip_heights, heights, theta: float):
    """(log density, gradient w.r.t. internal heights, gradient w.r.t. theta).

    Events are sorted by height (stable: at equal heights tips come before internal nodes,
    lower node id first).  Going back in time a tip adds a lineage, an internal node merges
    two."""
    n = n_tips
    allh = np.concatenate([np.asarray(tip_heights, dtype=np.float64), heights])
    order = np.argsort(allh, kind="stable")
    k = 0
    s = 0.0
    gh = np.zeros(n - 1)
    prev = allh[order[0]]
    for e in order:
        s += 0.5 * k * (k - 1) * (allh[e] - prev)
        prev = allh[e]
        if e >= n:
            gh[e - n] = -(k - 1) / theta  # C(k,2) - C(k-1,2) = k-1
            k -= 1
        else:
            k += 1
    logp = -s / theta - (n - 1) * np.log(theta)
    gtheta = s / theta ** 2 - (n - 1) / theta
    return float(logp), gh, float(gtheta)

"""CPU tests of the oracle itself (no GPU).

The reference pins no values for this path ("parity unpinned"), so
the oracle is pinned by (i) closed-form known answers, (ii) agreement of three
independent formulations (autograd-of-pruning, analytic two-pass, finite differences),
(iii) model invariances, and (iv) committed golden vectors (tests/golden/).
"""
import json
import math
import os

import numpy as np
import pytest

from b200phylo.alignment import SitePattern
from b200phylo.sitemodel import ConstantSiteModel, GammaSiteModel
from b200phylo.substitution import GTR, HKY, JC69
from b200phylo.tree import Tree, simulate_tree
from b200phylo.workloads import Workload, make_workload, random_patterns
from helpers import grad_err, oracle_ad, oracle_c, oracle_np, rel_err
from oracle import oracle


def _two_taxon(t1, t2, s1, s2):
    tree = Tree(2, np.array([[-1, -1], [-1, -1], [0, 1]], dtype=np.int32), np.array([t1, t2]), ["a", "b"])
    sp = SitePattern(["a", "b"], np.array([[1 << s1], [1 << s2]], dtype=np.uint8), np.ones(1))
    return Workload("two", tree, sp, JC69(), ConstantSiteModel())


def test_known_answer_two_taxa_jc69():
    # L = 1/4 * P_{xy}(t1+t2);  JC69: P_same = 1/4 + 3/4 e^{-4t/3}, P_diff = 1/4 - 1/4 e^{-4t/3}
    t1, t2 = 0.1, 0.25
    e = math.exp(-4.0 * (t1 + t2) / 3.0)
    for same in (True, False):
        w = _two_taxon(t1, t2, 0, 0 if same else 2)
        p = 0.25 + 0.75 * e if same else 0.25 - 0.25 * e
        dp = -e if same else e / 3.0
        for fn in (oracle_np, oracle_c, oracle_ad):
            logl, g = fn(w)
            assert logl == pytest.approx(math.log(0.25 * p), rel=1e-14)
            assert g == pytest.approx([dp / p, dp / p], rel=1e-12)


def test_known_answer_three_taxa_star_jc69():
    # ((a:t,b:t):0,c:t): with a zero internal branch this is a star; sum over the root state
    t = 0.3
    tree = Tree.from_newick("((a:0.3,b:0.3):0.0,c:0.3);")
    e = math.exp(-4.0 * t / 3.0)
    ps, pd = 0.25 + 0.75 * e, 0.25 - 0.25 * e
    masks = np.array([[1, 1, 1], [1, 1, 2], [1, 4, 8]], dtype=np.uint8)  # AAA, AAG, ACT columns
    sp = SitePattern(["a", "b", "c"], masks, np.array([2.0, 1.0, 3.0]))
    w = Workload("star", tree, sp, JC69(), ConstantSiteModel())
    l_aaa = 0.25 * (ps ** 3 + 3 * pd ** 3)
    l_aag = 0.25 * (ps * ps * pd + pd * pd * ps + 2 * pd ** 3)
    l_act = 0.25 * (3 * ps * pd * pd + pd ** 3)
    expect = 2 * math.log(l_aaa) + math.log(l_aag) + 3 * math.log(l_act)
    for fn in (oracle_np, oracle_c, oracle_ad):
        assert fn(w)[0] == pytest.approx(expect, rel=1e-13)


@pytest.mark.parametrize("name,n,p", [("jc69_20x1000", 20, 300), ("hky_g4_flu", 17, 120),
                                      ("gtr_g4_5000x10k", 33, 64)])
def test_three_formulations_agree(name, n, p):
    w = make_workload(name, n_tips=n, n_patterns=p)
    la, ga = oracle_np(w)
    lb, gb = oracle_ad(w)
    lc, gc = oracle_c(w)
    assert rel_err(lb, la) < 1e-13 and rel_err(lc, la) < 1e-13
    assert grad_err(gb, ga) < 1e-10 and grad_err(gc, ga) < 1e-10


def test_finite_differences_small_tree():
    w = make_workload("gtr_g4_5000x10k", n_tips=7, n_patterns=40)
    tp = w.patterns.partials()
    f = lambda bl: oracle.log_likelihood(w.tree.post_triples, tp, w.patterns.weights, bl, *w.eigen_args())
    g_fd = oracle.finite_difference(f, w.tree.branch_lengths, eps=1e-6)
    _, g = oracle_c(w)
    assert grad_err(g, g_fd) < 1e-6


def test_gamma_rates_yang1994():
    sm = GammaSiteModel(0.5, 4)
    assert sm.rates == pytest.approx([0.0334, 0.2519, 0.8203, 2.8944], abs=5e-5)
    assert sm.rates.mean() == pytest.approx(1.0, rel=1e-12)


def test_q_normalisation_and_reversibility():
    for m in (JC69(), HKY(3.0, (0.3, 0.2, 0.2, 0.3)), GTR((1, 2, 1, 1, 2, 1), (0.1, 0.2, 0.3, 0.4))):
        e = m.eigen()
        assert -(e.freqs * np.diag(e.q)).sum() == pytest.approx(1.0, rel=1e-13)
        flux = e.freqs[:, None] * e.q
        assert np.allclose(flux, flux.T, atol=1e-14)
        assert np.allclose(e.evec @ np.diag(e.evals) @ e.ivec, e.q, atol=1e-13)
        assert np.allclose(m.p_t(np.array(0.0)), np.eye(4), atol=1e-13)


def test_pulley_principle_root_branches():
    """Reversible model: logL depends on the two root branches only through their sum,
    and both root-adjacent gradients are equal (SURVEY.md §8(c) hazard iii)."""
    w = make_workload("hky_g4_flu", n_tips=15, n_patterns=80)
    l0, g0 = oracle_c(w)
    a, b = w.tree.children[w.tree.root]
    assert g0[a] == pytest.approx(g0[b], rel=1e-9)
    bl = w.tree.branch_lengths.copy()
    tot = bl[a] + bl[b]
    bl[a], bl[b] = 0.3 * tot, 0.7 * tot
    assert oracle_c(w, bl)[0] == pytest.approx(l0, rel=1e-13)


def test_weights_equal_duplicated_patterns():
    w = make_workload("jc69_20x1000", n_tips=10, n_patterns=200)
    sp = w.patterns
    reps = sp.weights.astype(int)
    dup = SitePattern(sp.names, np.repeat(sp.masks, reps, axis=1), np.ones(int(reps.sum())))
    w2 = Workload("dup", w.tree, dup, w.subst_model, w.site_model)
    l1, g1 = oracle_c(w)
    l2, g2 = oracle_c(w2)
    assert rel_err(l2, l1) < 1e-12 and grad_err(g2, g1) < 1e-10
    perm = np.random.default_rng(0).permutation(sp.n_patterns)
    w3 = Workload("perm", w.tree, SitePattern(sp.names, sp.masks[:, perm], sp.weights[perm]),
                  w.subst_model, w.site_model)
    assert rel_err(oracle_c(w3)[0], l1) < 1e-12


def test_rescaling_deep_tree_no_underflow():
    """1500 taxa underflow fp64 without rescaling (0.25^1500); numpy rescales at every
    node, the C oracle lazily by 2^200 — they must agree (hazard v)."""
    tree = simulate_tree(1500, shape="caterpillar", seed=5, mean_branch=0.05)
    sp = random_patterns(1500, 24, seed=3, p_change=0.3)
    w = Workload("deep", tree, sp, JC69(), ConstantSiteModel())
    la, ga = oracle_np(w)
    lc, gc = oracle_c(w)
    assert np.isfinite(la) and la < -1e4
    assert rel_err(lc, la) < 1e-12 and grad_err(gc, ga) < 1e-9


def test_ambiguity_and_gaps():
    w = make_workload("hky_g4_flu", n_tips=9, n_patterns=60)
    m = w.patterns.masks.copy()
    rng = np.random.default_rng(1)
    idx = rng.random(m.shape) < 0.2
    m[idx] = rng.choice([15, 5, 10, 3, 12, 7], size=int(idx.sum()))
    m[:, 0] = 15  # an all-gap column has likelihood 1
    w2 = Workload("amb", w.tree, SitePattern(w.patterns.names, m, w.patterns.weights), w.subst_model, w.site_model)
    la, ga = oracle_np(w2)
    lc, gc = oracle_c(w2)
    assert rel_err(lc, la) < 1e-13 and grad_err(gc, ga) < 1e-10
    only = Workload("gap", w.tree, SitePattern(w.patterns.names, m[:, :1], np.ones(1)), w.subst_model, w.site_model)
    assert abs(oracle_c(only)[0]) < 1e-12


def test_zero_length_branches():
    w = make_workload("jc69_20x1000", n_tips=12, n_patterns=100)
    bl = w.tree.branch_lengths.copy()
    bl[::3] = 0.0
    la, ga = oracle_np(w, bl)
    lc, gc = oracle_c(w, bl)
    assert np.isfinite(la) and np.all(np.isfinite(ga))
    assert rel_err(lc, la) < 1e-13 and grad_err(gc, ga) < 1e-10


def test_oracle_threads_agree():
    w = make_workload("jc69_1000x10k", n_tips=60, n_patterns=700)
    l1, g1 = oracle_c(w, n_threads=1)
    l4, g4 = oracle_c(w, n_threads=4)
    assert rel_err(l4, l1) < 1e-13 and grad_err(g4, g1) < 1e-12


def test_golden_vectors(golden_dir):
    """Committed golden vectors (generated by tests/golden/make_golden.py from the numpy
    oracle) — guard against silent drift of either oracle implementation."""
    files = sorted(f for f in os.listdir(golden_dir) if f.endswith(".npz") and not f.startswith("ref_"))
    assert files, "no golden fixtures"
    for f in files:
        z = np.load(os.path.join(golden_dir, f))
        meta = json.loads(str(z["meta"]))
        logl, g = oracle.loglik_and_gradient(
            z["triples"], oracle.tip_partials_from_masks(z["masks"]), z["weights"], z["bl"],
            z["evec"], z["ivec"], z["evals"], z["freqs"], z["rates"], z["cat_w"])
        assert rel_err(logl, z["logl"]) < 1e-12, meta
        assert grad_err(g, z["grad"]) < 1e-10, meta
        from oracle import c_oracle
        lc, gc = c_oracle.loglik_and_gradient(z["triples"], z["masks"], z["weights"], z["bl"], z["evec"],
                                              z["ivec"], z["evals"], z["freqs"], z["rates"], z["cat_w"])
        assert rel_err(lc, z["logl"]) < 1e-12 and grad_err(gc, z["grad"]) < 1e-10, meta


def test_param_oracle_value_and_branch_gradient_agree_with_the_analytic_oracle():
    """oracle/param_oracle.py (torch pruning, matrix_exp) against the C two-pass oracle: same
    value, and its autograd branch gradient equals the analytic one."""
    import torch
    from oracle import param_oracle
    w = make_workload("hky_g4_flu", n_tips=25, n_patterns=120)
    e = w.subst_model.eigen()
    bl = torch.tensor(w.tree.branch_lengths, requires_grad=True)
    ll = param_oracle.log_likelihood(w.tree.post_triples, w.patterns.partials(), w.patterns.weights, bl,
                                     torch.tensor(e.q), torch.tensor(e.freqs),
                                     torch.tensor(np.asarray(w.site_model.rates)), np.asarray(w.site_model.weights))
    ll.backward()
    ol, og = oracle_c(w)
    assert rel_err(float(ll.detach()), ol) < 1e-12
    assert grad_err(bl.grad.numpy(), og) < 1e-9


def test_dense_c_oracle_matches_numpy_oracle():
    """oracle/phylo_oracle_dense.c (run-time state count) against the numpy restatement: the 61-state
    codon model, and a 4-state model with rate categories against the 4-state C oracle."""
    from oracle import c_oracle
    w = make_workload("codon_500x2000", n_tips=24, n_patterns=50)
    codes = w.patterns.masks.copy()
    codes[np.random.default_rng(1).random(codes.shape) < 0.05] = 255  # missing codons
    from b200phylo.alignment import SitePattern
    sp = SitePattern(w.patterns.names, codes, w.patterns.weights, 61)
    ln, gn = oracle.loglik_and_gradient(w.tree.post_triples, sp.partials(), sp.weights, w.tree.branch_lengths, *w.eigen_args())
    lc, gc = c_oracle.dense_loglik_and_gradient(w.tree.post_triples, codes, sp.weights, w.tree.branch_lengths, *w.eigen_args())
    assert abs(ln - lc) <= 1e-12 * abs(ln)
    assert np.max(np.abs(gn - gc)) <= 1e-10 * np.max(np.abs(gn))
    w = make_workload("hky_g4_flu", n_tips=40, n_patterns=150)
    m = w.patterns.masks
    codes = np.select([m == 1, m == 2, m == 4, m == 8], [0, 1, 2, 3], 9).astype(np.uint8)
    l4, g4 = c_oracle.loglik_and_gradient(w.tree.post_triples, m, w.patterns.weights, w.tree.branch_lengths, *w.eigen_args())
    ld, gd = c_oracle.dense_loglik_and_gradient(w.tree.post_triples, codes, w.patterns.weights, w.tree.branch_lengths, *w.eigen_args())
    assert abs(l4 - ld) <= 1e-13 * abs(l4) and np.max(np.abs(g4 - gd)) <= 1e-11 * np.max(np.abs(g4))
    lo, _ = c_oracle.dense_loglik_and_gradient(w.tree.post_triples, codes, w.patterns.weights, w.tree.branch_lengths, *w.eigen_args(),
                                               want_grad=False, n_threads=1)
    assert abs(lo - ld) <= 1e-13 * abs(ld)


def test_branch_outer_products_identities():
    """oracle.branch_outer_products: contracting G with r_k Q gives the branch gradient, its traces
    sum to the number of sites (sum_k c_k q.m = L), and it follows the analytic two-pass gradient's
    partials — for one and four categories, with ambiguity codes in the tips."""
    from b200phylo.sitemodel import ConstantSiteModel, GammaSiteModel
    from b200phylo.substitution import GTR
    from b200phylo.workloads import Workload, make_workload
    from oracle import oracle
    for site in (ConstantSiteModel(), GammaSiteModel(0.6, 4)):
        base = make_workload("hky_g4_flu", n_tips=14, n_patterns=60)
        w = Workload("g", base.tree, base.patterns, GTR((1.3, 2.9, 0.7, 1.1, 3.4, 1.0), (0.31, 0.19, 0.22, 0.28)), site)
        args = (w.tree.post_triples, w.patterns.partials(), w.patterns.weights, w.tree.branch_lengths, *w.eigen_args())
        g = oracle.branch_outer_products(*args)
        _, grad = oracle.loglik_and_gradient(*args)
        e = w.eigen_args()
        q = oracle.q_from_eigen(e[0], e[1], e[2])
        rates = np.asarray(w.site_model.rates)
        assert g.shape == (len(rates), 2 * w.n_tips - 2, 4, 4)
        np.testing.assert_allclose(np.einsum("k,ij,kbij->b", rates, q, g), grad, rtol=1e-11, atol=1e-9)
        np.testing.assert_allclose(np.einsum("kbii->b", g), np.full(2 * w.n_tips - 2, w.patterns.weights.sum()), rtol=1e-12)

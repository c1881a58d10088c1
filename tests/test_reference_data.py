"""Parity hardening with the reference's own data files.

The reference pins no VALUES for this path, but it holds two real inputs and the model template:
  * /root/reference/treetime_validation/resources/beast/bedford_et_al_2015__h3_small_tree.xml
    (1391 taxa x 1698 sites -> 1113 patterns, IUPAC R/Y/W/K/M/S/D/?/-),
  * /root/reference/treetime_validation/resources/flu_H3N2/H3N2_HA_2011_2013.nwk
    (3561 tips, polytomies, ~1811 zero-length branches),
  * /root/reference/flu_H3N2/phylotorch-JC69.template (substituted as /root/reference/SConstruct:75-122 does).
`tests/golden/make_reference_fixtures.py` derives small fixtures from them; the tests that need the
reference tree itself (absent on the GPU box) say so in their skip reason, the others run from
the fixtures everywhere.
"""
import json
import os

import numpy as np
import pytest

from b200phylo.alignment import IUPAC_MASK, SitePattern, compress_alignment, read_beast_xml, simulate_alignment
from b200phylo.sitemodel import ConstantSiteModel, GammaSiteModel
from b200phylo.substitution import HKY, JC69
from b200phylo.tree import Tree, simulate_tree
from b200phylo.workloads import Workload
from helpers import grad_err, grad_report, make_engine, oracle_c, oracle_np, rel_err

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REF = "/root/reference"
H3_XML = f"{REF}/treetime_validation/resources/beast/bedford_et_al_2015__h3_small_tree.xml"
H3N2_NWK = f"{REF}/treetime_validation/resources/flu_H3N2/H3N2_HA_2011_2013.nwk"
TEMPLATE = f"{REF}/flu_H3N2/phylotorch-JC69.template"
needs_reference = pytest.mark.skipif(not os.path.exists(REF), reason="/root/reference is not on this box (GPU box): "
                                     "the committed fixtures under tests/golden/ref_* stand in")


def h3_patterns() -> SitePattern:
    z = np.load(os.path.join(GOLDEN, "ref_h3_patterns.npz"))
    return SitePattern([f"t{i}" for i in range(z["masks"].shape[0])], z["masks"], z["weights"])


def h3n2_tree() -> Tree:
    z = np.load(os.path.join(GOLDEN, "ref_h3n2_tree.npz"))
    tr, bl = z["post_triples"], z["branch_lengths"]
    n = tr.shape[0] + 1
    children = np.full((2 * n - 1, 2), -1, dtype=np.int32)
    children[tr[:, 0]] = tr[:, 1:]
    t = Tree(n, children, bl.copy(), [f"t{i}" for i in range(n)])
    t.validate()
    return t


# ------------------------------------------------------------------ CPU: fixtures vs the reference files
@needs_reference
def test_h3_alignment_of_the_reference():
    aln = read_beast_xml(H3_XML)
    assert len(aln.names) == 1391 and {len(s) for s in aln.sequences} == {1698}
    assert set("".join(aln.sequences)) == set("ACGT") | set("RYWKMSD?-")
    sp = compress_alignment(aln)
    assert sp.n_patterns == 1113 and sp.weights.sum() == 1698
    fx = h3_patterns()
    assert np.array_equal(sp.masks, fx.masks) and np.array_equal(sp.weights, fx.weights)


@needs_reference
def test_h3n2_tree_of_the_reference():
    with open(H3N2_NWK) as fh:
        t = Tree.from_newick(fh.read())
    t.validate()
    bl = t.branch_lengths
    assert t.n_tips == 3561
    assert int((bl == 0).sum()) == 1811 and int(np.isclose(bl, 1e-5).sum()) == 2297
    fx = h3n2_tree()
    assert np.array_equal(t.post_triples, fx.post_triples) and np.array_equal(bl, fx.branch_lengths)


@needs_reference
def test_template_fixture_is_the_reference_template_substituted():
    import sys
    sys.path.insert(0, GOLDEN)
    import make_reference_fixtures as mk
    doc = json.load(open(os.path.join(GOLDEN, "ref_phylotorch_jc69.json")))
    taxa = doc[0]["taxa"]
    like = doc[1]["distributions"][0]
    names = [t["id"] for t in taxa]
    text = mk.substituted_template(open(TEMPLATE).read(), names, [t["attributes"]["date"] for t in taxa],
                                   [s["sequence"] for s in like["site_pattern"]["alignment"]["sequences"]],
                                   like["tree_model"]["newick"])
    assert json.loads(text) == doc


def test_reference_fixtures_shape():
    sp = h3_patterns()
    assert sp.masks.shape == (1391, 1113) and sp.weights.sum() == 1698
    assert set(np.unique(sp.masks)) == {IUPAC_MASK[c] for c in "ACGTRYWKMSD-"}
    t = h3n2_tree()
    assert t.n_tips == 3561 and t.levels().max() > 50


# ------------------------------------------------------------------ workloads built on the reference data
def h3_workload(model="hky_g4") -> Workload:
    """The reference's real H3 patterns (ambiguity-heavy) on a simulated 1391-tip tree."""
    sp = h3_patterns()
    tree = simulate_tree(1391, seed=11, mean_branch=0.004)
    tree.names = sp.names
    if model == "jc69":
        return Workload("ref_h3_jc69", tree, sp, JC69(), ConstantSiteModel())
    return Workload("ref_h3_hky_g4", tree, sp, HKY(3.0, (0.33, 0.19, 0.23, 0.25)), GammaSiteModel(0.5, 4))


def h3n2_workload(n_sites=600) -> Workload:
    """Sequences simulated on the reference's real 3561-tip tree (1811 zero-length branches)."""
    tree = h3n2_tree()
    model, sm = JC69(), ConstantSiteModel()
    pm = model.p_t(tree.branch_lengths[:, None] * np.asarray(sm.rates)[None, :])
    sp = compress_alignment(simulate_alignment(tree, pm, model.freqs, n_sites, seed=5))
    return Workload("ref_h3n2_tree_jc69", tree, sp, model, sm)


def test_oracles_agree_on_the_reference_patterns():
    w = h3_workload("jc69")
    lc, gc = oracle_c(w)
    ln, gn = oracle_np(w)
    assert np.isfinite(lc) and rel_err(lc, ln) < 1e-12
    assert grad_err(gc, gn) < 1e-9


def test_oracle_on_the_reference_tree_zero_length_branches():
    """Hazard (iv) of SURVEY.md §8(c): half the branches of the real tree have length 0 (P = I)."""
    w = h3n2_workload(200)
    bl = w.tree.branch_lengths
    l0, g = oracle_c(w)
    assert np.isfinite(l0) and np.all(np.isfinite(g))
    # one-sided differences on zero-length branches, central ones on polytomy-resolution branches
    rng = np.random.default_rng(0)
    for b in rng.choice(np.flatnonzero(bl == 0), 3, replace=False):
        h = 1e-7
        up = bl.copy(); up[b] += h
        fd = (oracle_c(w, up, want_grad=False)[0] - l0) / h
        assert abs(fd - g[b]) <= 1e-3 * max(1.0, abs(g[b])), (b, fd, g[b])
    for b in rng.choice(np.flatnonzero(np.isclose(bl, 1e-5)), 3, replace=False):
        h = 2e-6
        up, dn = bl.copy(), bl.copy(); up[b] += h; dn[b] -= h
        fd = (oracle_c(w, up, want_grad=False)[0] - oracle_c(w, dn, want_grad=False)[0]) / (2 * h)
        assert abs(fd - g[b]) <= 1e-4 * max(1.0, abs(g[b])), (b, fd, g[b])


# ------------------------------------------------------------------ GPU parity on the reference data
@pytest.mark.gpu
@pytest.mark.parametrize("model", ["jc69", "hky_g4"])
def test_gpu_reference_h3_patterns(model):
    w = h3_workload(model)
    eng = make_engine(w)
    logl, grad = eng.loglik_grad_host(w.tree.branch_lengths)
    ol, og = oracle_c(w)
    rep = grad_report(grad, og)
    print(f"{w.name}: logL {logl:.6f} rel {rel_err(logl, ol):.1e}  gradient {rep}")
    assert rel_err(logl, ol) < 1e-10
    assert rep["floored"] < 1e-8
    eng.close()


@pytest.mark.gpu
def test_gpu_reference_h3n2_tree():
    w = h3n2_workload()
    eng = make_engine(w)
    logl, grad = eng.loglik_grad_host(w.tree.branch_lengths)
    ol, og = oracle_c(w)
    rep = grad_report(grad, og)
    print(f"{w.name}: logL {logl:.6f} rel {rel_err(logl, ol):.1e}  gradient {rep}")
    assert np.isfinite(logl) and np.all(np.isfinite(grad))
    assert rel_err(logl, ol) < 1e-10
    assert rep["floored"] < 1e-8
    eng.close()


@pytest.mark.gpu
def test_gpu_from_json_reference_template():
    """The "like" object of the reference's own template, after SConstruct's substitution."""
    import torch

    from b200phylo.treelikelihood import TreeLikelihoodModel
    from oracle import c_oracle
    from oracle import timetree_oracle as tto
    doc = json.load(open(os.path.join(GOLDEN, "ref_phylotorch_jc69.json")))
    assert doc[0]["id"] == "taxa" and doc[1]["id"] == "joint"
    like = doc[1]["distributions"][0]
    assert like["id"] == "like" and like["type"] == "TreeLikelihoodModel"
    dic = {"taxa": doc[0]}
    m = TreeLikelihoodModel.from_json(like, dic)
    n = len(doc[0]["taxa"])
    for key in ("tree", "sp", "clock", "rate", "ratios", "ratios.unres", "root_height", "root_height.unres", "like"):
        assert key in dic, key
    x = dic["ratios.unres"].tensor
    assert x.is_cuda and x.shape == (n - 2,)
    x.requires_grad_(True)
    logl = m()
    logl.backward()
    assert torch.isfinite(logl) and torch.isfinite(x.grad).all() and float(x.grad.abs().sum()) > 0
    tm = m.tree_model
    tips = tm.tip_heights
    loc = like["tree_model"]["root_height"]["parameters"]["loc"]
    params = np.concatenate([np.full(n - 2, 1 / (1 + np.exp(-0.5))), [loc + 0.5]])
    tr = tm.post_triples
    oh = tto.heights_from_ratios(n, tr, tips, params)
    sp = m.site_pattern
    e = JC69().eigen()
    ol, _ = c_oracle.loglik_and_gradient(tr, sp.masks, sp.weights, tto.branch_lengths(n, tr, tips, oh, 0.001),
                                         e.evec, e.ivec, e.evals, e.freqs, [1.0], [1.0])
    assert rel_err(float(logl), ol) < 1e-10
    m.close()

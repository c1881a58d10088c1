"""CPU tests of the host-side layer: Newick/FASTA I/O, pattern compression, sharding,
JSON plugin parsing, CLI argument contract.  No compute calls (no GPU here)."""
import os
import re

import numpy as np
import pytest

from b200phylo.alignment import (IUPAC_MASK, Alignment, SitePattern, compress_alignment, read_fasta,
                                 write_fasta)
from b200phylo.parallel import shard_bounds
from b200phylo.tree import Tree, simulate_tree
from b200phylo.workloads import make_workload


def test_newick_roundtrip_and_numbering():
    t = simulate_tree(37, seed=3)
    t2 = Tree.from_newick(t.to_newick(precision=17), taxa=t.names)
    assert t2.n_tips == 37
    assert np.array_equal(t2.children, t.children)
    assert np.allclose(t2.branch_lengths, t.branch_lengths, rtol=1e-15)
    tr = t.post_triples
    assert tr.shape == (36, 3) and tr[-1, 0] == t.root == 72
    assert np.all(tr[:, 1] < tr[:, 0]) and np.all(tr[:, 2] < tr[:, 0])


def test_newick_polytomy_unary_comments_and_dates():
    # polytomies become binary splits joined by 1e-5 branches, as the reference's data
    # preparation does (/root/reference/treetime_validation/utility_functions_general.py:18-46)
    txt = "((A_2001.5:0.1,B_2002:0.2,C_2003:0.3)[&date=1999]:0.05,((D_2000:0.4):0.1,'E_2004':0.2):0.3);"
    t = Tree.from_newick(txt)
    assert t.n_tips == 5 and sorted(t.names) == ["A_2001.5", "B_2002", "C_2003", "D_2000", "E_2004"]
    t.validate()
    assert np.isclose(t.branch_lengths, 1e-5).sum() == 1
    d = t.names.index("D_2000")
    assert t.branch_lengths[d] == pytest.approx(0.5)  # unary node merged


def test_tree_shapes_and_levels():
    for shape in ("coalescent", "yule", "balanced", "caterpillar"):
        t = simulate_tree(33, shape=shape, seed=1)
        t.validate()
        assert t.branch_lengths.shape == (64,) and np.all(t.branch_lengths >= 0)
    assert simulate_tree(33, shape="caterpillar").levels().max() == 32
    assert simulate_tree(32, shape="balanced").levels().max() == 5


def test_fasta_and_compression(tmp_path):
    aln = Alignment(["a_2000", "b_2001", "c_2002"], ["ACGTACGTNN-R", "ACGTACGAnn-R", "ACCTACCT??-Y"])
    p = tmp_path / "x.fasta"
    write_fasta(str(p), aln)
    back = read_fasta(str(p))
    assert back.names == aln.names and back.sequences == aln.sequences
    sp = compress_alignment(back)
    assert sp.weights.sum() == 12
    assert sp.masks.shape[0] == 3
    cols = {tuple(sp.masks[:, j]) for j in range(sp.n_patterns)}
    assert len(cols) == sp.n_patterns  # unique
    # gaps / N / ? are all-ones, IUPAC R = A|G, Y = C|T
    assert IUPAC_MASK["-"] == IUPAC_MASK["N"] == IUPAC_MASK["?"] == 15
    j = [k for k in range(sp.n_patterns) if sp.masks[0, k] == 5][0]
    assert tuple(sp.masks[:, j]) == (5, 5, 10)
    # first-occurrence order, deterministic
    assert tuple(sp.masks[:, 0]) == (1, 1, 1) and sp.weights[0] == 2


def test_partials_and_reorder():
    sp = SitePattern(["x", "y"], np.array([[1, 15], [10, 4]], dtype=np.uint8), np.array([1.0, 2.0]))
    tp = sp.partials()
    assert tp.shape == (2, 2, 4)
    assert tp[0, 0].tolist() == [1, 0, 0, 0] and tp[0, 1].tolist() == [1, 1, 1, 1]
    assert tp[1, 0].tolist() == [0, 1, 0, 1]
    ro = sp.reorder(["y", "x"])
    assert ro.masks[0].tolist() == [10, 4]


@pytest.mark.parametrize("p,w", [(10, 1), (10, 3), (100000, 8), (7, 7), (1701, 4)])
def test_shard_bounds_cover_exactly(p, w):
    spans = [shard_bounds(p, w, r) for r in range(w)]
    assert spans[0][0] == 0 and spans[-1][1] == p
    for a, b in zip(spans, spans[1:]):
        assert a[1] == b[0]
    sizes = [b - a for a, b in spans]
    assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_bounds(p, w, w)


def test_workload_seeded_and_sized():
    a = make_workload("jc69_20x1000")
    b = make_workload("jc69_20x1000")
    assert a.n_tips == 20 and a.patterns.weights.sum() == 1000
    assert np.array_equal(a.patterns.masks, b.patterns.masks)
    assert np.array_equal(a.tree.branch_lengths, b.tree.branch_lengths)
    c = make_workload("jc69_1000x10k", n_tips=50, n_patterns=128)
    assert c.n_patterns == 128 and set(np.unique(c.patterns.masks)) <= {1, 2, 4, 8}


def test_cli_contract_flags():
    """Flags the pipeline passes (/root/reference/modules/micro.nf:47-53) must parse."""
    from b200phylo import benchmark
    src = open(benchmark.__file__).read()
    for flag in ('"-i"', '"-t"', '"-r"', '"-s"', '"-o"', '"-d"', '"--gtr"'):
        assert flag in src
    assert "function,mode,JIT,time" in src
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    assert os.access(os.path.join(root, "bin", "b200phylo-benchmark"), os.X_OK)


def test_json_subobject_parsers():
    """The sub-objects of the reference's "like" JSON
    (/root/reference/flu_H3N2/phylotorch-JC69.template:11-74) parse without a GPU."""
    from b200phylo import treelikelihood as tl
    sp = tl._site_pattern_from_json({"id": "sp", "type": "SitePattern", "alignment": {
        "id": "alignment", "type": "Alignment", "datatype": "nucleotide", "taxa": "taxa",
        "sequences": [{"taxon": "A_2000", "sequence": "ACGTAC"}, {"taxon": "B_2001", "sequence": "ACGTAA"},
                      {"taxon": "C_2002", "sequence": "ACCTAC"}]}}, {})
    assert sp.names == ["A_2000", "B_2001", "C_2002"] and sp.weights.sum() == 6
    tm = tl._tree_model_from_json({"id": "tree", "newick": "((B_2001:1,C_2002:2):0.5,A_2000:3);"}, {}, sp)
    assert tm.taxa == sp.names and tm.branch_lengths().shape == (4,)
    assert float(tm.branch_lengths()[0]) == 3.0  # indexed by taxon order of the alignment
    assert isinstance(tl._subst_model_from_json({"id": "m", "type": "JC69"}), tl.JC69)
    hky = tl._subst_model_from_json({"type": "HKY", "kappa": {"tensor": [2.5]},
                                    "frequencies": {"tensor": [0.1, 0.2, 0.3, 0.4]}})
    assert hky.kappa == 2.5 and hky.freqs[3] == pytest.approx(0.4)
    assert len(tl._site_model_from_json({"type": "GammaSiteModel", "shape": 0.5, "categories": 4}).rates) == 4
    assert len(tl._site_model_from_json({"id": "sm", "type": "ConstantSiteModel"}).rates) == 1
    clock = tl._clock_from_json({"id": "clock", "type": "StrictClockModel", "tree_model": "tree",
                                 "rate": {"id": "rate", "type": "Parameter", "tensor": [0.001]}})
    assert float(clock.rates[0]) == 0.001
    with pytest.raises(ValueError):
        tl._subst_model_from_json({"type": "WAG"})


def test_read_lsd_nexus(tmp_path):
    """lsd2's dated Nexus (/root/reference/bin/helper.py:169-176): annotations are dropped,
    the tree and its branch lengths survive."""
    from b200phylo.tree import read_lsd_nexus
    p = tmp_path / "lsd.out.date.nexus"
    p.write_text('#NEXUS\nBegin trees;\ntree 1 = ((a_2001[&date="2001"]:1.5,b_2003.5[&date="2003.5"]:4)[&date="1999.5"]:2,'
                 'c_2000[&date="2000"]:2.5)[&date="1997.5"];\nEnd;\n')
    t = read_lsd_nexus(str(p))
    assert sorted(t.names) == ["a_2001", "b_2003.5", "c_2000"]
    bl = {t.names[i]: t.branch_lengths[i] for i in range(3)}
    assert bl == {"a_2001": 1.5, "b_2003.5": 4.0, "c_2000": 2.5}
    from b200phylo.timetree import tip_heights_from_names
    np.testing.assert_allclose(tip_heights_from_names(["a_2001", "b_2003.5", "c_2000"]), [2.5, 0.0, 3.5])


def test_weibull_site_model():
    from b200phylo.sitemodel import WeibullSiteModel
    from b200phylo import treelikelihood as tl
    m = WeibullSiteModel(1.0, 4)  # shape 1 = exponential: quantiles -log(1-q)
    q = -np.log1p(-np.array([0.125, 0.375, 0.625, 0.875]))
    np.testing.assert_allclose(m.rates, q / q.mean(), rtol=1e-15)
    assert m.rates.mean() == pytest.approx(1.0) and np.all(np.diff(m.rates) > 0) and m.weights.sum() == pytest.approx(1.0)
    j = tl._site_model_from_json({"type": "WeibullSiteModel", "shape": {"tensor": [0.7]}, "categories": 6})
    assert len(j.rates) == 6 and j.rates.mean() == pytest.approx(1.0)


def test_differentiable_models_match_their_host_twins():
    """b200phylo.diffmodels: the torch-parameter models give the numbers of the numpy ones (the
    engine is fed from ``eigen()``, ``backward`` from ``q_tensor()``)."""
    import torch
    from b200phylo import treelikelihood as tl
    from b200phylo.diffmodels import DifferentiableGTR, DifferentiableHKY, DifferentiableWeibullSiteModel, is_differentiable
    from b200phylo.sitemodel import WeibullSiteModel
    from b200phylo.substitution import GTR, HKY
    ex = tl.Parameter("gtr.rates", torch.tensor([1.3, 2.9, 0.7, 1.1, 3.4, 1.0], dtype=torch.float64))
    fr = tl.Parameter("gtr.freqs", torch.tensor([0.31, 0.19, 0.22, 0.28], dtype=torch.float64))
    m = DifferentiableGTR("gtr", ex, fr)
    ref = GTR((1.3, 2.9, 0.7, 1.1, 3.4, 1.0), (0.31, 0.19, 0.22, 0.28))
    np.testing.assert_allclose(m.q_tensor().numpy(), ref.q_matrix(), rtol=1e-14, atol=1e-16)
    np.testing.assert_allclose(m.eigen().evals, ref.eigen().evals, rtol=1e-13, atol=1e-15)
    assert m.parameters() == [ex, fr] and not is_differentiable(m)
    ex.requires_grad = True
    assert is_differentiable(m) and m.q_tensor().requires_grad
    assert not is_differentiable(ref)
    h = DifferentiableHKY("hky", torch.tensor([2.7], dtype=torch.float64), torch.tensor([0.1, 0.2, 0.3, 0.4], dtype=torch.float64))
    np.testing.assert_allclose(h.q_tensor().numpy(), HKY(2.7, (0.1, 0.2, 0.3, 0.4)).q_matrix(), rtol=1e-14, atol=1e-16)
    assert h.kappa == 2.7 and h.parameters() == []
    w = DifferentiableWeibullSiteModel("site", tl.Parameter("shape", torch.tensor([0.7], dtype=torch.float64, requires_grad=True)), 6)
    np.testing.assert_allclose(w.rates, WeibullSiteModel(0.7, 6).rates, rtol=1e-14)
    assert is_differentiable(w) and len(w.weights) == 6
    # the JSON parsers keep Parameter objects given by id or inline
    dic = {}
    g = tl._subst_model_from_json({"id": "gtr", "type": "GTR", "rates": {"id": "r", "type": "Parameter", "tensor": [1, 2, 1, 1, 2, 1]},
                                   "frequencies": {"id": "f", "type": "Parameter", "tensor": [0.3, 0.2, 0.2, 0.3]}}, dic)
    assert isinstance(g, DifferentiableGTR) and g.parameters() == [dic["r"], dic["f"]]
    np.testing.assert_allclose(g.q_matrix(), GTR((1, 2, 1, 1, 2, 1), (0.3, 0.2, 0.2, 0.3)).q_matrix(), rtol=1e-14, atol=1e-16)
    assert isinstance(tl._subst_model_from_json({"type": "GTR", "rates": [1, 2, 1, 1, 2, 1], "frequencies": [0.3, 0.2, 0.2, 0.3]}), GTR)


def test_differentiable_gamma_site_model():
    """Gamma shape as a tensor: the rates of sitemodel.GammaSiteModel, d rates / d alpha by central
    difference inside a custom autograd function (checked against a wider difference here)."""
    import torch
    from b200phylo import treelikelihood as tl
    from b200phylo.diffmodels import DifferentiableGammaSiteModel, is_differentiable
    from b200phylo.sitemodel import GammaSiteModel
    alpha = tl.Parameter("alpha", torch.tensor([0.6], dtype=torch.float64, requires_grad=True))
    m = DifferentiableGammaSiteModel("site", alpha, 4)
    np.testing.assert_allclose(m.rates, GammaSiteModel(0.6, 4).rates, rtol=1e-15)
    assert is_differentiable(m) and m.parameters() == [alpha]
    c = torch.tensor([0.3, -1.0, 2.0, 0.5], dtype=torch.float64)
    (m.rates_tensor() * c).sum().backward()
    h = 1e-4
    want = float(((GammaSiteModel(0.6 + h, 4).rates - GammaSiteModel(0.6 - h, 4).rates) / (2 * h) * c.numpy()).sum())
    assert float(alpha.tensor.grad) == pytest.approx(want, rel=1e-6)
    j = tl._site_model_from_json({"type": "GammaSiteModel", "shape": {"id": "a", "type": "Parameter", "tensor": [0.5]}, "categories": 4}, {})
    assert isinstance(j, DifferentiableGammaSiteModel) and len(j.rates) == 4

"""GPU tests of the plugin surface (TreeLikelihoodModel / autograd / CLI) and of
size-independent properties at BASELINE.json's full sizes."""
import csv
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from b200phylo.alignment import Alignment, write_fasta
from b200phylo.sitemodel import ConstantSiteModel, GammaSiteModel
from b200phylo.substitution import HKY, JC69
from b200phylo.workloads import Workload, make_workload
from helpers import grad_err, make_engine, oracle_c, rel_err

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _model(w, bl, clock=None, **kw):
    from b200phylo.treelikelihood import TreeLikelihoodModel, UnrootedTreeModel
    tm = UnrootedTreeModel("tree", w.tree, bl)
    return TreeLikelihoodModel("like", w.patterns, tm, w.subst_model, w.site_model, clock, **kw)


@pytest.mark.parametrize("where", ["cuda", "cpu"])
def test_model_call_and_autograd(where):
    w = make_workload("hky_g4_flu", n_tips=50, n_patterns=400)
    bl = torch.tensor(w.tree.branch_lengths, dtype=torch.float64, device=where, requires_grad=True)
    m = _model(w, bl)
    logl = m()
    assert logl.dim() == 0 and logl.dtype == torch.float64 and logl.device.type == where
    (2.0 * logl).backward()
    ol, og = oracle_c(w)
    assert rel_err(float(logl), ol) < 1e-10
    assert grad_err(bl.grad.cpu().numpy(), 2.0 * og) < 1e-8
    with torch.no_grad():
        assert rel_err(float(m()), ol) < 1e-10
    m.close()


def test_strict_clock_chain_rule():
    """branch_model = StrictClockModel: d logL / d rate = sum_b t_b * dlogL/d(rate*t_b)
    (/root/reference/flu_H3N2/phylotorch-JC69.template:49-58)."""
    from b200phylo.treelikelihood import StrictClockModel
    w = make_workload("jc69_20x1000")
    times = torch.tensor(w.tree.branch_lengths / 0.001, dtype=torch.float64, device="cuda", requires_grad=True)
    rate = torch.tensor([0.001], dtype=torch.float64, device="cuda", requires_grad=True)
    m = _model(w, times, StrictClockModel("clock", rate))
    m().backward()
    ol, og = oracle_c(w)
    assert float(rate.grad) == pytest.approx(float(np.sum(og * w.tree.branch_lengths / 0.001)), rel=1e-8)
    assert grad_err(times.grad.cpu().numpy(), og * 0.001) < 1e-8
    m.close()


def test_from_json_reference_template_shape():
    from b200phylo.treelikelihood import TreeLikelihoodModel
    w = make_workload("jc69_20x1000", n_tips=8, n_patterns=200)
    from b200phylo.alignment import simulate_alignment
    pm = w.subst_model.p_t(w.tree.branch_lengths[:, None] * np.ones(1)[None, :])
    aln = simulate_alignment(w.tree, pm, w.subst_model.freqs, 300, seed=9)
    data = {"id": "like", "type": "b200phylo.TreeLikelihoodModel",
            "tree_model": {"id": "tree", "type": "UnrootedTreeModel", "newick": w.tree.to_newick(17)},
            "site_model": {"id": "sm", "type": "ConstantSiteModel"},
            "branch_model": {"id": "clock", "type": "StrictClockModel", "tree_model": "tree",
                             "rate": {"id": "rate", "type": "Parameter", "tensor": [1.0]}},
            "site_pattern": {"id": "sp", "type": "SitePattern", "alignment": {
                "id": "alignment", "type": "Alignment", "datatype": "nucleotide", "taxa": "taxa",
                "sequences": [{"taxon": n, "sequence": s} for n, s in zip(aln.names, aln.sequences)]}},
            "substitution_model": {"id": "m", "type": "JC69"}}
    dic = {}
    m = TreeLikelihoodModel.from_json(data, dic)
    assert {"like", "tree", "sm", "clock", "sp", "m"} <= set(dic)
    from b200phylo.alignment import compress_alignment
    w2 = Workload("j", w.tree, compress_alignment(aln), JC69(), ConstantSiteModel())
    assert rel_err(float(m()), oracle_c(w2)[0]) < 1e-10
    m.close()


def test_cli_contract_end_to_end(tmp_path):
    """`b200phylo-benchmark -i F -t T -r R -s S -o out.csv --gtr [-d float32]` writes the
    CSV the pipeline's awk + plot.Rmd expect (/root/reference/modules/micro.nf:47-56)."""
    w = make_workload("jc69_20x1000")
    from b200phylo.alignment import simulate_alignment
    pm = w.subst_model.p_t(w.tree.branch_lengths[:, None] * np.ones(1)[None, :])
    aln = simulate_alignment(w.tree, pm, w.subst_model.freqs, 1000, seed=9)
    fa, nw, out = tmp_path / "a.fasta", tmp_path / "t.nwk", tmp_path / "out.csv"
    write_fasta(str(fa), aln)
    nw.write_text(w.tree.scaled(1000.0).to_newick(17))  # a "time tree"; -s 0.001 rescales it
    for extra in ([], ["-d", "float32"]):
        subprocess.check_call([sys.executable, os.path.join(ROOT, "bin", "b200phylo-benchmark"), "-i", str(fa),
                               "-t", str(nw), "-r", "20", "-s", "0.001", "-o", str(out), "--gtr"] + extra)
        rows = list(csv.reader(open(out)))
        assert rows[0] == ["function", "mode", "JIT", "time"]
        assert [r[:3] for r in rows[1:5]] == [
            ["treelikelihoodJC69", "evaluation", "off"], ["treelikelihoodJC69", "gradient", "off"],
            ["treelikelihoodGTR", "evaluation", "off"], ["treelikelihoodGTR", "gradient", "off"]]
        assert all(float(r[3]) > 0 for r in rows[1:])
        # dated tip labels: the time-tree functions follow (/root/reference/plot.Rmd:49-60)
        assert [r[0] for r in rows[5:]] == ["ratio_transform"] * 2 + ["ratio_transform_jacobian"] * 2 + ["coalescent"] * 2
        side = json.load(open(str(out) + ".b200.json"))
        from b200phylo.alignment import compress_alignment
        w2 = Workload("c", w.tree, compress_alignment(aln), JC69(), ConstantSiteModel())
        tol = 1e-4 if extra else 1e-9  # newick text round-trip of the branch lengths
        assert rel_err(side[0]["logl"], oracle_c(w2)[0]) < tol


# ------------------------------------------------------------- full-size properties
def _full(name):
    w = make_workload(name)
    eng = make_engine(w)
    return w, eng


def test_config2_full_size_properties():
    """JC69, 1000 taxa x 10k patterns, fp64 (BASELINE configs[1]): oracle parity at full
    size (the C oracle takes ~0.2 s) plus determinism, pulley principle, shard additivity."""
    w, eng = _full("jc69_1000x10k")
    bl = w.tree.branch_lengths
    l1, g1 = eng.loglik_grad_host(bl)
    l2, g2 = eng.loglik_grad_host(bl)
    assert l1 == l2 and np.array_equal(g1, g2)  # deterministic reductions: bit-identical
    ol, og = oracle_c(w)
    assert rel_err(l1, ol) < 1e-10 and grad_err(g1, og) < 1e-8
    a, b = w.tree.children[w.tree.root]
    assert g1[a] == pytest.approx(g1[b], rel=1e-8)
    bl2 = bl.copy()
    bl2[a], bl2[b] = 0.25 * (bl[a] + bl[b]), 0.75 * (bl[a] + bl[b])
    assert rel_err(eng.loglik_grad_host(bl2, want_grad=False)[0], l1) < 1e-11
    eng.close()
    # pattern shards add up (the multi-GPU decomposition)
    from b200phylo.parallel import shard_bounds
    tot_l, tot_g = 0.0, np.zeros_like(g1)
    for r in range(3):
        lo, hi = shard_bounds(w.n_patterns, 3, r)
        ws = Workload("s", w.tree, w.patterns.slice(lo, hi), w.subst_model, w.site_model)
        e = make_engine(ws)
        l, g = e.loglik_grad_host(bl)
        e.close()
        tot_l += l
        tot_g += g
    assert rel_err(tot_l, l1) < 1e-12 and grad_err(tot_g, g1) < 1e-10


def test_config3_fp32_vs_fp64():
    """HKY+G4 flu-shaped (1704 taxa x 1701 sites): fp32 within 1e-4 of fp64 on logL."""
    w = make_workload("hky_g4_flu")
    e64, e32 = make_engine(w), make_engine(w, precision="float32")
    l64, g64 = e64.loglik_grad_host(w.tree.branch_lengths)
    l32, g32 = e32.loglik_grad_host(w.tree.branch_lengths)
    e64.close(); e32.close()
    ol, og = oracle_c(w)
    assert rel_err(l64, ol) < 1e-10 and grad_err(g64, og) < 1e-8
    assert rel_err(l32, l64) < 1e-4
    assert grad_err(g32, g64) < 5e-3


def test_gate_shape_gtr_g4_5000x10k_linearity():
    """GTR+G4, 5000 taxa x 10k patterns: weights are linear — doubling every weight
    doubles logL and the gradient exactly; one tenth of the patterns is checked against
    the oracle."""
    from b200phylo.workloads import random_patterns
    from b200phylo.tree import simulate_tree
    base = make_workload("gtr_g4_5000x10k", n_tips=16, n_patterns=16)
    tree = simulate_tree(5000, seed=20261104)
    sp = random_patterns(5000, 10000)
    w = Workload("gate", tree, sp, base.subst_model, base.site_model)
    eng = make_engine(w)
    l1, g1 = eng.loglik_grad_host(tree.branch_lengths)
    eng.set_pattern_weights(2.0 * sp.weights)
    l2, g2 = eng.loglik_grad_host(tree.branch_lengths)
    eng.close()
    assert l2 == pytest.approx(2.0 * l1, rel=1e-14) and np.allclose(g2, 2.0 * g1, rtol=1e-13)
    ws = Workload("gate10", tree, sp.slice(0, 1000), base.subst_model, base.site_model)
    e = make_engine(ws)
    l, g = e.loglik_grad_host(tree.branch_lengths)
    e.close()
    ol, og = oracle_c(ws)
    assert rel_err(l, ol) < 1e-10 and grad_err(g, og) < 1e-8


def test_two_live_instances_of_different_size():
    """Kernel attributes (dynamic shared memory) are process-wide: creating a second, smaller
    instance must not break launches of the first."""
    wa = make_workload("jc69_20x1000", n_tips=300, n_patterns=3000)
    wb = make_workload("jc69_20x1000", n_tips=12, n_patterns=40)
    ea = make_engine(wa)
    eb = make_engine(wb)
    la, ga = ea.loglik_grad_host(wa.tree.branch_lengths)
    lb, gb = eb.loglik_grad_host(wb.tree.branch_lengths)
    la2, _ = ea.loglik_grad_host(wa.tree.branch_lengths)
    assert la == la2
    assert rel_err(la, oracle_c(wa)[0]) < 1e-10 and rel_err(lb, oracle_c(wb)[0]) < 1e-10
    ea.close()
    eb.close()

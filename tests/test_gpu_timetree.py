"""GPU parity of the time-tree kernels (b2p_timetree_*, b2p_coalescent_constant) against
oracle/timetree_oracle.py, through the C ABI (ctypes) and through the model classes.

Tolerances: the height recursion, its adjoint, the branch lengths and every gradient that
is a fixed-order expression are compared bit for bit; reductions (log-Jacobian, coalescent
sum, d/d rate) to 1e-12 relative — the GPU sums in a tree, the oracle serially."""
import numpy as np
import pytest
import torch

from b200phylo.tree import simulate_tree
from b200phylo.workloads import make_workload
from helpers import grad_err, oracle_c, rel_err
from oracle import timetree_oracle as tto

pytestmark = pytest.mark.gpu


def dated_case(n, shape, seed, contemporaneous=False, ties=False):
    tree = simulate_tree(n, shape=shape, seed=seed)
    rng = np.random.default_rng(seed + 1)
    tips = np.zeros(n) if contemporaneous else rng.uniform(0.0, 12.0, n)
    if ties:
        tips = np.round(tips)  # many tips share a sampling date
    if not contemporaneous:
        tips[rng.integers(n)] = 0.0
    params = rng.uniform(0.05, 0.95, n - 1)
    if ties:
        params[rng.integers(0, max(1, n - 2), size=max(1, n // 10))] = 0.0  # nodes sitting on their bound
    params[n - 2] = tips.max() + rng.uniform(0.5, 5.0)
    return tree, tips, params


def cu(a):
    return torch.tensor(np.asarray(a, dtype=np.float64), device="cuda")


CASES = [(2, "balanced", 1), (3, "caterpillar", 2), (33, "coalescent", 3), (64, "balanced", 4),
         (1000, "coalescent", 5), (3000, "caterpillar", 6), (4096, "balanced", 7), (5000, "coalescent", 8),
         (9000, "coalescent", 9)]  # 9000: level arrays do not fit shared memory (global fallback)


@pytest.mark.parametrize("n,shape,seed", CASES)
def test_heights_and_adjoint_bit_exact(n, shape, seed):
    from b200phylo.timetree import TimeTreeEngine
    tree, tips, params = dated_case(n, shape, seed)
    tr = tree.post_triples
    eng = TimeTreeEngine(tr, tips)
    info = eng.info()
    assert info["in_smem"] == (1 if n <= 6000 else 0)
    h = eng.heights(cu(params))
    want = tto.heights_from_ratios(n, tr, tips, params)
    np.testing.assert_array_equal(h.cpu().numpy(), want)
    g = np.random.default_rng(seed).normal(size=n - 1)
    got = eng.heights_vjp(cu(params), h, cu(g)).cpu().numpy()
    np.testing.assert_array_equal(got, tto.heights_vjp(n, tr, tips, params, g))
    eng.close()


@pytest.mark.parametrize("n,shape,seed", [(2, "balanced", 1), (3, "caterpillar", 2), (200, "coalescent", 3),
                                         (1000, "coalescent", 5), (5000, "balanced", 8)])
def test_log_jacobian_and_branch_lengths(n, shape, seed):
    from b200phylo.timetree import TimeTreeEngine
    tree, tips, params = dated_case(n, shape, seed)
    tr = tree.post_triples
    eng = TimeTreeEngine(tr, tips)
    h = tto.heights_from_ratios(n, tr, tips, params)
    logj, grad = eng.log_jacobian(cu(h), want_grad=True)
    if n > 2:
        assert rel_err(float(logj), tto.log_jacobian(n, tr, tips, h)) < 1e-12
    else:
        assert float(logj) == 0.0
    np.testing.assert_array_equal(grad.cpu().numpy(), tto.log_jacobian_grad(n, tr, tips, h))
    rate = 0.0031
    bl = eng.branch_lengths(cu(h), cu([rate]))
    np.testing.assert_array_equal(bl.cpu().numpy(), tto.branch_lengths(n, tr, tips, h, rate))
    np.testing.assert_array_equal(eng.branch_lengths(cu(h)).cpu().numpy(), tto.branch_lengths(n, tr, tips, h, 1.0))
    g = np.random.default_rng(seed).normal(size=2 * n - 2)
    gh, gr = eng.branch_lengths_vjp(cu(h), cu([rate]), cu(g))
    wh, wr = tto.branch_lengths_vjp(n, tr, tips, h, rate, g)
    np.testing.assert_array_equal(gh.cpu().numpy(), wh)
    assert float(gr) == pytest.approx(wr, rel=1e-11, abs=1e-11 * np.abs(g).sum() * h.max())
    eng.close()


@pytest.mark.parametrize("n,shape,seed,contemp,ties", [(2, "balanced", 1, True, False), (3, "caterpillar", 2, False, False),
                                                       (40, "coalescent", 3, True, False), (333, "coalescent", 4, False, True),
                                                       (1000, "coalescent", 5, False, False), (2000, "caterpillar", 6, False, True),
                                                       (5000, "coalescent", 8, False, False)])
def test_constant_coalescent(n, shape, seed, contemp, ties):
    from b200phylo.timetree import TimeTreeEngine
    tree, tips, params = dated_case(n, shape, seed, contemp, ties)
    tr = tree.post_triples
    eng = TimeTreeEngine(tr, tips)
    h = tto.heights_from_ratios(n, tr, tips, params)
    theta = 3.7
    for _ in range(2):  # second launch: the completion ticket was reset by the first
        logp, gh, gt = eng.constant_coalescent(cu(h), cu([theta]))
    wl, wh, wt = tto.constant_coalescent(n, tips, h, theta)
    assert rel_err(float(logp), wl) < 1e-12
    np.testing.assert_array_equal(gh.cpu().numpy(), wh)
    assert rel_err(float(gt), wt) < 1e-11
    eng.close()


def test_cpu_tensor_is_an_error_not_a_fallback():
    from b200phylo.engine import EngineError
    from b200phylo.timetree import TimeTreeEngine
    tree, tips, params = dated_case(10, "coalescent", 1)
    eng = TimeTreeEngine(tree.post_triples, tips)
    with pytest.raises(EngineError):
        eng.heights(torch.tensor(params))
    with pytest.raises(ValueError):
        eng.heights(cu(params[:-1]))
    eng.close()
    bad = tree.post_triples.copy()
    bad[0, 1] = bad[0, 0]
    with pytest.raises(EngineError):
        TimeTreeEngine(bad, tips)


def test_joint_density_gradient_chain():
    """The macro model's joint (/root/reference/flu_H3N2/phylotorch-JC69.template:11-100):
    tree likelihood (JC69, strict clock) + constant coalescent + log-Jacobian, differentiated
    w.r.t. ratios, root height, clock rate and population size in one backward pass."""
    from b200phylo.timetree import ConstantCoalescentModel, ReparameterizedTimeTreeModel
    from b200phylo.treelikelihood import StrictClockModel, TreeLikelihoodModel
    w = make_workload("jc69_20x1000", n_tips=120, n_patterns=600)
    n = w.n_tips
    rng = np.random.default_rng(5)
    tips = rng.uniform(0.0, 8.0, n)
    tips[3] = 0.0
    x = rng.uniform(0.1, 0.9, n - 1)
    x[n - 2] = 15.0
    ratios = cu(x[: n - 2]).requires_grad_(True)
    root = cu(x[n - 2:]).requires_grad_(True)
    rate = cu([0.004]).requires_grad_(True)
    theta = cu([6.0]).requires_grad_(True)
    tm = ReparameterizedTimeTreeModel("tree", w.tree, tips, ratios, root)
    like = TreeLikelihoodModel("like", w.patterns, tm, w.subst_model, w.site_model, StrictClockModel("clock", rate))
    coal = ConstantCoalescentModel("coalescent", tm, theta)
    h = tm.heights()
    joint = like() + coal(h) + tm.log_jacobian(h)
    joint.backward()
    # oracle chain
    tr = w.tree.post_triples
    oh = tto.heights_from_ratios(n, tr, tips, x)
    obl = tto.branch_lengths(n, tr, tips, oh, 0.004)
    ol, og = oracle_c(w, bl=obl)
    cl, cgh, cgt = tto.constant_coalescent(n, tips, oh, 6.0)
    gh, grate = tto.branch_lengths_vjp(n, tr, tips, oh, 0.004, og)
    gh = gh + cgh + tto.log_jacobian_grad(n, tr, tips, oh)
    gx = tto.heights_vjp(n, tr, tips, x, gh)
    assert rel_err(float(joint.detach()), ol + cl + tto.log_jacobian(n, tr, tips, oh)) < 1e-10
    assert grad_err(ratios.grad.cpu().numpy(), gx[: n - 2]) < 1e-8
    assert float(root.grad) == pytest.approx(gx[n - 2], rel=1e-8)
    assert float(rate.grad) == pytest.approx(grate, rel=1e-8)
    assert float(theta.grad) == pytest.approx(cgt, rel=1e-10)
    like.close()
    tm.close()


def test_time_tree_from_json_template():
    """tree_model of type ReparameterizedTimeTreeModel with transformed parameters
    (/root/reference/flu_H3N2/phylotorch-JC69.template:14-44)."""
    from b200phylo.treelikelihood import TreeLikelihoodModel
    w = make_workload("jc69_20x1000", n_tips=12, n_patterns=60)
    n = w.n_tips
    dates = np.linspace(1990.0, 2011.0, n)
    names = [f"A/x{i}_{dates[i]:.3f}" for i in range(n)]
    tree = w.tree
    tree.names = names
    seqs = []
    part = w.patterns.masks
    code = {1: "A", 2: "C", 4: "G", 8: "T"}
    for i in range(n):
        seqs.append({"taxon": names[i], "sequence": "".join(code.get(int(m), "N") for m in part[i])})
    spec = {
        "id": "like", "type": "b200phylo.treelikelihood.TreeLikelihoodModel",
        "tree_model": {"id": "tree", "type": "ReparameterizedTimeTreeModel", "newick": tree.to_newick(),
                       "ratios": {"id": "ratios", "type": "TransformedParameter",
                                  "transform": "torch.distributions.SigmoidTransform",
                                  "x": {"id": "ratios.unres", "type": "Parameter", "tensor": 0.5, "full": [n - 2]}},
                       "root_height": {"id": "root_height", "type": "TransformedParameter",
                                       "transform": "torch.distributions.AffineTransform",
                                       "parameters": {"loc": 21.0, "scale": 1.0},
                                       "x": {"id": "root_height.unres", "type": "Parameter", "tensor": [0.5]}},
                       "taxa": "taxa"},
        "site_model": {"id": "sm", "type": "ConstantSiteModel"},
        "branch_model": {"id": "clock", "type": "StrictClockModel", "tree_model": "tree",
                         "rate": {"id": "rate", "type": "Parameter", "tensor": [0.001]}},
        "site_pattern": {"id": "sp", "type": "SitePattern", "alignment": {"id": "aln", "sequences": seqs}},
        "substitution_model": {"id": "m", "type": "JC69"},
    }
    dic = {}
    m = TreeLikelihoodModel.from_json(spec, dic)
    x = dic["ratios.unres"].tensor
    assert x.is_cuda and x.shape == (n - 2,)
    x.requires_grad_(True)
    logl = m()
    logl.backward()
    assert torch.isfinite(logl) and torch.isfinite(x.grad).all() and float(x.grad.abs().sum()) > 0
    # same value from the oracle chain
    tm = m.tree_model
    tips = tm.tip_heights
    params = np.concatenate([np.full(n - 2, 1 / (1 + np.exp(-0.5))), [21.5]])
    tr = tm.post_triples
    oh = tto.heights_from_ratios(n, tr, tips, params)
    from oracle import c_oracle
    sp = m.site_pattern
    ol, _ = c_oracle.loglik_and_gradient(tr, sp.masks, sp.weights, tto.branch_lengths(n, tr, tips, oh, 0.001),
                                         *w.eigen_args())
    assert rel_err(float(logl), ol) < 1e-10
    m.close()


def test_cli_writes_time_function_rows(tmp_path):
    """Dated tip labels -> the CLI also times ratio_transform, ratio_transform_jacobian and
    coalescent (/root/reference/plot.Rmd:49-60,301-303) in the same CSV vocabulary."""
    import csv
    import os
    import subprocess
    import sys
    from b200phylo.alignment import simulate_alignment, write_fasta
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    w = make_workload("jc69_20x1000")
    n = w.n_tips
    tree = w.tree
    # a clock-like dated tree: tip date = latest date - (height implied by the branch lengths)
    time_tree = tree.scaled(1000.0)
    heights = time_tree.node_heights()
    tree.names = [f"t{i}_{2015.0 - heights[i]:.6f}" for i in range(n)]
    time_tree.names = list(tree.names)
    pm = w.subst_model.p_t(tree.branch_lengths[:, None] * np.ones(1)[None, :])
    aln = simulate_alignment(tree, pm, w.subst_model.freqs, 500, seed=9)
    fa, nw, out = tmp_path / "a.fasta", tmp_path / "t.nwk", tmp_path / "out.csv"
    write_fasta(str(fa), aln)
    nw.write_text(time_tree.to_newick(17))
    subprocess.check_call([sys.executable, os.path.join(root, "bin", "b200phylo-benchmark"), "-i", str(fa),
                           "-t", str(nw), "-r", "20", "-s", "0.001", "-o", str(out)])
    rows = list(csv.reader(open(out)))
    assert rows[0] == ["function", "mode", "JIT", "time"]
    assert [r[:2] for r in rows[1:]] == [
        ["treelikelihoodJC69", "evaluation"], ["treelikelihoodJC69", "gradient"],
        ["ratio_transform", "evaluation"], ["ratio_transform", "gradient"],
        ["ratio_transform_jacobian", "evaluation"], ["ratio_transform_jacobian", "gradient"],
        ["coalescent", "evaluation"], ["coalescent", "gradient"]]
    assert all(float(r[3]) > 0 and r[2] == "off" for r in rows[1:])


def test_initial_params_reproduce_the_dated_tree():
    from b200phylo.timetree import ReparameterizedTimeTreeModel
    w = make_workload("jc69_20x1000", n_tips=80, n_patterns=50)
    tree = w.tree.scaled(1000.0)
    hh = tree.node_heights()
    tm = ReparameterizedTimeTreeModel("tree", tree, hh[: tree.n_tips].copy())
    got = tm.heights().cpu().numpy()
    np.testing.assert_allclose(got, hh[tree.n_tips:], rtol=1e-9, atol=1e-9)
    bl = tm.branch_lengths().cpu().numpy()
    np.testing.assert_allclose(bl, tree.branch_lengths, rtol=1e-7, atol=1e-9)
    tm.close()


@pytest.mark.parametrize("with_coalescent,with_jacobian", [(True, True), (False, True), (True, False)])
def test_joint_density_single_call_matches_the_chain(with_coalescent, with_jacobian):
    """b2p_joint_density_grad (one C call) against the oracle chain and against the
    function-by-function autograd path."""
    from b200phylo.timetree import ConstantCoalescentModel, JointTimeTreeDensity, ReparameterizedTimeTreeModel
    from b200phylo.treelikelihood import StrictClockModel, TreeLikelihoodModel
    w = make_workload("hky_g4_flu", n_tips=90, n_patterns=500)
    n = w.n_tips
    rng = np.random.default_rng(15)
    tips = rng.uniform(0.0, 8.0, n)
    tips[7] = 0.0
    x = rng.uniform(0.1, 0.9, n - 1)
    x[n - 2] = 15.0
    ratios = cu(x[: n - 2]).requires_grad_(True)
    root = cu(x[n - 2:]).requires_grad_(True)
    rate = cu([0.004]).requires_grad_(True)
    theta = cu([6.0]).requires_grad_(True)
    tm = ReparameterizedTimeTreeModel("tree", w.tree, tips, ratios, root)
    like = TreeLikelihoodModel("like", w.patterns, tm, w.subst_model, w.site_model, StrictClockModel("clock", rate))
    coal = ConstantCoalescentModel("coalescent", tm, theta) if with_coalescent else None
    joint = JointTimeTreeDensity(like, tm, coal, include_jacobian=with_jacobian)
    total = joint()
    (3.0 * total).backward()
    got = [t.grad.clone() for t in (ratios, root, rate)] + ([theta.grad.clone()] if with_coalescent else [])
    terms = joint.terms.cpu().numpy()
    # oracle chain
    tr = w.tree.post_triples
    oh = tto.heights_from_ratios(n, tr, tips, x)
    ol, og = oracle_c(w, bl=tto.branch_lengths(n, tr, tips, oh, 0.004))
    gh, grate = tto.branch_lengths_vjp(n, tr, tips, oh, 0.004, og)
    want = ol
    assert rel_err(terms[1], ol) < 1e-10
    if with_coalescent:
        cl, cgh, cgt = tto.constant_coalescent(n, tips, oh, 6.0)
        gh, want = gh + cgh, want + cl
        assert rel_err(terms[2], cl) < 1e-12
        assert float(got[3]) == pytest.approx(3.0 * cgt, rel=1e-10)
    if with_jacobian:
        gh, want = gh + tto.log_jacobian_grad(n, tr, tips, oh), want + tto.log_jacobian(n, tr, tips, oh)
    gx = tto.heights_vjp(n, tr, tips, x, gh)
    assert rel_err(float(total.detach()), want) < 1e-10 and rel_err(terms[0], want) < 1e-10
    assert grad_err(got[0].cpu().numpy(), 3.0 * gx[: n - 2]) < 1e-8
    assert float(got[1]) == pytest.approx(3.0 * gx[n - 2], rel=1e-8)
    assert float(got[2]) == pytest.approx(3.0 * grate, rel=1e-8)
    # the function-by-function path gives the same numbers
    for t in (ratios, root, rate, theta):
        t.grad = None
    h = tm.heights()
    sep = like()
    if with_coalescent:
        sep = sep + coal(h)
    if with_jacobian:
        sep = sep + tm.log_jacobian(h)
    (3.0 * sep).backward()
    assert rel_err(float(sep.detach()), float(total.detach())) < 1e-13
    assert grad_err(ratios.grad.cpu().numpy(), got[0].cpu().numpy()) < 1e-12
    like.close()
    tm.close()


def test_advi_cli_runs_and_prints_parsable_progress(tmp_path):
    """b200phylo-advi with the macro pipeline's flags (/root/reference/modules/macro_flu.nf:124-137):
    the ELBO goes up and the progress lines parse with bin/time-parser.py's torchtree pattern."""
    import os
    import re
    import subprocess
    import sys
    from b200phylo.alignment import simulate_alignment, write_fasta
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    w = make_workload("jc69_20x1000")
    n = w.n_tips
    tree = w.tree
    time_tree = tree.scaled(1000.0)
    heights = time_tree.node_heights()
    tree.names = [f"t{i}_{2015.0 - heights[i]:.6f}" for i in range(n)]
    time_tree.names = list(tree.names)
    pm = w.subst_model.p_t(tree.branch_lengths[:, None] * np.ones(1)[None, :])
    aln = simulate_alignment(tree, pm, w.subst_model.freqs, 800, seed=9)
    fa, nw = tmp_path / "a.fasta", tmp_path / "t.nwk"
    write_fasta(str(fa), aln)
    nw.write_text(time_tree.to_newick(17))
    txt = subprocess.check_output([sys.executable, os.path.join(root, "bin", "b200phylo-advi"), "-i", str(fa), "-t", str(nw),
                                   "--clock", "strict", "--coalescent", "constant", "--heights_init", "tree",
                                   "--rate_init", "0.001", "--clockpr", "exponential", "--lr", "0.01", "--elbo_samples", "1",
                                   "--tol_rel_obj", "0", "--iter", "400", "--samples", "0", "--print_every", "50"], text=True)
    pat = re.compile(r"\s+(\d+)\s+(-\d+\.\d+).+")  # /root/reference/bin/time-parser.py:60-61
    rows = [(int(m.group(1)), float(m.group(2))) for m in map(pat.match, txt.splitlines()) if m]
    assert rows[0][0] == 1 and rows[-1][0] == 400 and len(rows) == 9
    assert all(np.isfinite(e) for _, e in rows)
    assert np.mean([e for _, e in rows[-3:]]) > rows[0][1]  # variational fit improves on the start
    assert "Ran inference for 400 iterations" in txt

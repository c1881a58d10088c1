"""GPU parity tests: the CUDA path, called through the C ABI (ctypes), against the CPU
oracle on the same seeded inputs, against the committed golden vectors, and — at the
BASELINE.json sizes — through size-independent properties.

Tolerances are BASELINE.json north_star's: log-likelihood 1e-10 relative in fp64 (1e-4
in fp32), gradients 1e-8 relative in fp64.
"""
import json
import os

import numpy as np
import pytest

from b200phylo.alignment import SitePattern
from b200phylo.sitemodel import ConstantSiteModel, GammaSiteModel
from b200phylo.substitution import GTR, HKY, JC69
from b200phylo.tree import Tree, simulate_tree
from b200phylo.workloads import Workload, make_workload, random_patterns
from helpers import grad_err, make_engine, oracle_c, oracle_np, rel_err

pytestmark = pytest.mark.gpu

TOL_LOGL_F64 = 1e-10
TOL_GRAD_F64 = 1e-8
TOL_LOGL_F32 = 1e-4
TOL_GRAD_F32 = 5e-3  # not a north_star gate; fp32 partials, fp64 reductions


def _check(w, precision="float64", **kw):
    eng = make_engine(w, precision=precision, **kw)
    try:
        logl, grad = eng.loglik_grad_host(w.tree.branch_lengths)
        logl_only, none = eng.loglik_grad_host(w.tree.branch_lengths, want_grad=False)
    finally:
        eng.close()
    assert none is None
    ol, og = oracle_c(w)
    tl, tg = (TOL_LOGL_F64, TOL_GRAD_F64) if precision == "float64" else (TOL_LOGL_F32, TOL_GRAD_F32)
    assert rel_err(logl, ol) < tl, (logl, ol)
    assert rel_err(logl_only, ol) < tl
    assert grad_err(grad, og) < tg
    return logl, grad


@pytest.mark.parametrize("name,n,p", [
    ("jc69_20x1000", None, None),        # BASELINE configs[0], full size
    ("jc69_1000x10k", 200, 3000),
    ("hky_g4_flu", 150, 700),
    ("gtr_g4_5000x10k", 257, 1031),      # ragged: P*K not a multiple of the tile width
])
def test_parity_fp64(name, n, p):
    _check(make_workload(name, n_tips=n, n_patterns=p))


@pytest.mark.parametrize("name,n,p", [("jc69_20x1000", None, None), ("hky_g4_flu", 150, 700)])
def test_parity_fp32(name, n, p):
    _check(make_workload(name, n_tips=n, n_patterns=p), precision="float32")


def test_golden_vectors_gpu(golden_dir):
    from b200phylo.engine import Engine
    for f in sorted(x for x in os.listdir(golden_dir) if x.endswith(".npz") and not x.startswith("ref_")):
        z = np.load(os.path.join(golden_dir, f))
        n, p = z["masks"].shape
        eng = Engine(n, p, len(z["rates"]))
        eng.set_topology(z["triples"])
        eng.set_tip_states(z["masks"])
        eng.set_pattern_weights(z["weights"])
        eng.set_eigen(z["evec"], z["ivec"], z["evals"], z["freqs"])
        eng.set_categories(z["rates"], z["cat_w"])
        logl, grad = eng.loglik_grad_host(z["bl"])
        eng.close()
        assert rel_err(logl, z["logl"]) < TOL_LOGL_F64, f
        assert grad_err(grad, z["grad"]) < TOL_GRAD_F64, f


@pytest.mark.parametrize("n,p", [(2, 1), (2, 33), (3, 1), (5, 31), (9, 64), (9, 65)])
def test_tiny_and_ragged(n, p):
    base = make_workload("gtr_g4_5000x10k", n_tips=4, n_patterns=4)
    tree = simulate_tree(n, seed=n, mean_branch=0.1)
    sp = random_patterns(n, p, seed=p, p_change=0.4)  # columns need not be unique
    _check(Workload("tiny", tree, sp, base.subst_model, base.site_model))


@pytest.mark.parametrize("k", [1, 2, 3, 4, 5, 8])
def test_category_counts(k):
    base = make_workload("hky_g4_flu", n_tips=40, n_patterns=300)
    sm = GammaSiteModel(0.7, k) if k > 1 else ConstantSiteModel()
    _check(Workload("k", base.tree, base.patterns, base.subst_model, sm))


@pytest.mark.parametrize("shape", ["caterpillar", "balanced", "yule"])
def test_tree_shapes(shape):
    tree = simulate_tree(130, shape=shape, seed=11, mean_branch=0.05)
    sp = random_patterns(130, 500, seed=2, p_change=0.2)
    _check(Workload(shape, tree, sp, HKY(2.0, (0.3, 0.2, 0.2, 0.3)), GammaSiteModel(0.5, 4)))


def test_rescaling_deep_tree_fp64_and_fp32():
    """Partials underflow without rescaling: 0.25^2000 (fp64) — hazard (v)."""
    tree = simulate_tree(2000, shape="caterpillar", seed=5, mean_branch=0.05)
    sp = random_patterns(2000, 96, seed=3, p_change=0.3)
    w = Workload("deep", tree, sp, JC69(), ConstantSiteModel())
    logl, _ = _check(w)
    assert np.isfinite(logl) and logl < -1e4
    _check(w, precision="float32")


def test_rescaling_divergent_balanced_tree_fp32():
    """ADVICE.md (round 1): on a balanced tree of very divergent sequences the two children's large
    entries sit on different states, so a node's partial can be far more than ONE rescale factor
    (2^40 in fp32) below the threshold; with a single fixed step per node the deficit doubles per
    level and fp32 underflows to logL = -inf within a few levels.  The rescale takes as many steps
    as the exponent asks for; the C oracle (fp64, one step suffices there) is the reference."""
    tree = simulate_tree(512, shape="balanced", seed=4, mean_branch=0.002)
    sp = random_patterns(512, 200, seed=8, p_change=0.75)  # near-random tips on near-zero branches
    w = Workload("divergent", tree, sp, JC69(), ConstantSiteModel())
    for prec in ("float64", "float32"):
        logl, grad = _check(w, precision=prec)
        assert np.isfinite(logl) and np.all(np.isfinite(grad)) and logl < -1e5


def test_ambiguity_codes_and_gaps():
    w = make_workload("hky_g4_flu", n_tips=60, n_patterns=400)
    m = w.patterns.masks.copy()
    rng = np.random.default_rng(1)
    idx = rng.random(m.shape) < 0.2
    m[idx] = rng.choice([15, 5, 10, 3, 12, 7, 14], size=int(idx.sum()))
    m[:, 0] = 15
    _check(Workload("amb", w.tree, SitePattern(w.patterns.names, m, w.patterns.weights), w.subst_model,
                    w.site_model))


def test_zero_length_branches():
    w = make_workload("jc69_1000x10k", n_tips=100, n_patterns=500)
    w.tree.branch_lengths[::3] = 0.0
    _check(w)


def test_branch_partials_match_oracle():
    """What a node sends to its parent, P_v x_v, against the oracle's post-order partials."""
    from oracle import oracle
    w = make_workload("hky_g4_flu", n_tips=12, n_patterns=90)
    eng = make_engine(w)
    eng.loglik_grad_host(w.tree.branch_lengths)
    e = w.eigen_args()
    pm = oracle.transition_matrices(w.tree.branch_lengths, e[0], e[1], e[2], e[4])
    post, _ = oracle.post_order(w.tree.post_triples, w.patterns.partials(), pm)
    for node in (w.n_tips, w.n_tips + 3, 2 * w.n_tips - 3):
        got = eng.branch_partials(node)  # [P, K, 4]
        ref = np.einsum("kij,kpj->pki", pm[node], post[node])
        ratio = got / ref
        assert np.allclose(ratio, ratio[:, :1, :1], rtol=1e-11)  # equal up to the per-pattern scale
    eng.close()


def test_setup_errors_are_reported():
    from b200phylo.engine import Engine, EngineError
    eng = Engine(4, 8, 1)
    with pytest.raises(EngineError, match="before setup"):
        eng.loglik_grad_host(np.ones(6))
    with pytest.raises(EngineError, match="post-order|range|parents"):
        eng.set_topology(np.array([[4, 0, 1], [5, 4, 4], [6, 5, 3]], dtype=np.int32))
    with pytest.raises(EngineError, match="mask"):
        eng.set_tip_states(np.full((4, 8), 16, dtype=np.uint8))
    eng.close()
    with pytest.raises(EngineError, match="4-state|n_states|unsupported|one rate category"):
        import ctypes
        from b200phylo import engine as E
        h = ctypes.c_void_p()
        cfg = E._Config(4, 8, 1, 200, 64, 0, 0, 0)
        rc = E.load_library().b2p_create(ctypes.byref(cfg), ctypes.byref(h))
        if rc != 0:
            raise EngineError(E.load_library().b2p_last_error(None).decode())


# ---- dense-state path: 61-state codon model (BASELINE.json configs[4]) --------------------
@pytest.mark.parametrize("n,p", [(2, 3), (12, 40), (60, 130), (150, 333)])
def test_codon_parity(n, p):
    """logL and branch gradient of the 61-state path against the numpy oracle (analytic
    two-pass formulation, generic in the state count)."""
    w = make_workload("codon_500x2000", n_tips=n, n_patterns=p)
    assert w.patterns.n_states == 61
    eng = make_engine(w)
    try:
        logl, grad = eng.loglik_grad_host(w.tree.branch_lengths)
        logl_only, _ = eng.loglik_grad_host(w.tree.branch_lengths, want_grad=False)
    finally:
        eng.close()
    ol, og = oracle_np(w)
    assert rel_err(logl, ol) < TOL_LOGL_F64, (logl, ol)
    assert rel_err(logl_only, ol) < TOL_LOGL_F64
    assert grad_err(grad, og) < TOL_GRAD_F64


def test_codon_missing_states_and_rescaling():
    """Missing codons (code 255) are all-ones tips; a deep caterpillar tree underflows
    61-state partials without rescaling (61^-400)."""
    from b200phylo.alignment import simulate_patterns
    base = make_workload("codon_500x2000", n_tips=8, n_patterns=8)
    tree = simulate_tree(400, shape="caterpillar", seed=3, mean_branch=0.05)
    # sequences evolved on the tree: with random codes the likelihood is dominated by the
    # ~1e-9 entries of P (three-nucleotide changes), whose eigen-form rounding error makes
    # any two implementations disagree at 1e-9 — an ill-conditioned input, not a parity case
    pm = base.subst_model.p_t(tree.branch_lengths[:, None] * np.ones(1)[None, :])
    sp = simulate_patterns(tree, pm, base.subst_model.freqs, 70, seed=9)
    rng = np.random.default_rng(5)
    codes = sp.masks.copy()
    codes[rng.random(codes.shape) < 0.1] = 255
    sp = SitePattern(sp.names, codes, sp.weights, 61)
    w = Workload("codon_deep", tree, sp, base.subst_model, base.site_model)
    eng = make_engine(w)
    try:
        logl, grad = eng.loglik_grad_host(w.tree.branch_lengths)
    finally:
        eng.close()
    ol, og = oracle_np(w)
    assert np.isfinite(logl) and logl < -1e4
    assert rel_err(logl, ol) < TOL_LOGL_F64, (logl, ol)
    assert grad_err(grad, og) < TOL_GRAD_F64


def test_codon_full_size_properties():
    """BASELINE.json configs[4] at its full size (500 taxa x 2000 codon patterns), where the numpy
    oracle takes minutes: size-independent properties instead.  Two evaluations are bit-identical
    (every sum has a fixed order), and doubling the pattern weights doubles logL and every
    derivative (a power of two commutes with every rounding)."""
    w = make_workload("codon_500x2000")
    assert w.n_tips == 500 and w.n_patterns == 2000
    eng = make_engine(w)
    try:
        logl, grad = eng.loglik_grad_host(w.tree.branch_lengths)
        logl2, grad2 = eng.loglik_grad_host(w.tree.branch_lengths)
    finally:
        eng.close()
    assert np.isfinite(logl) and np.all(np.isfinite(grad))
    assert logl == logl2 and np.array_equal(grad, grad2)
    sp = w.patterns
    w2 = Workload("codon_w2", w.tree, SitePattern(sp.names, sp.masks, sp.weights * 2.0, 61), w.subst_model, w.site_model)
    eng = make_engine(w2)
    try:
        logl_w, grad_w = eng.loglik_grad_host(w2.tree.branch_lengths)
    finally:
        eng.close()
    # exact in exact-scaling arithmetic; 1e-14 leaves room for a fused multiply-add contraction
    assert rel_err(logl_w, 2.0 * logl) < 1e-14 and grad_err(grad_w, 2.0 * grad) < 1e-14


def test_codon_full_size_against_the_c_oracle():
    """BASELINE.json configs[4] at its full size against oracle/phylo_oracle_dense.c (the two-pass
    algorithm with a run-time state count; ~2 s on eight cores): logL and every branch derivative."""
    from oracle import c_oracle
    w = make_workload("codon_500x2000")
    eng = make_engine(w)
    try:
        logl, grad = eng.loglik_grad_host(w.tree.branch_lengths)
    finally:
        eng.close()
    ol, og = c_oracle.dense_loglik_and_gradient(w.tree.post_triples, w.patterns.masks, w.patterns.weights,
                                                w.tree.branch_lengths, *w.eigen_args())
    assert rel_err(logl, ol) < TOL_LOGL_F64, (logl, ol)
    assert grad_err(grad, og) < TOL_GRAD_F64


@pytest.mark.parametrize("env", [{"B2P_NP": "2"}, {"B2P_NP": "1"}, {"B2P_NP": "4"}, {"B2P_NT": "640"}, {"B2P_NT": "768"},
                                 {"B2P_CPT": "1"}, {"B2P_DEPTH": "1"}, {"B2P_DEPTH": "3"}, {"B2P_CTAS": "2"},
                                 {"B2P_CT_MAX": "64"}, {"B2P_MAT_ROWS": "14"}, {"B2P_PLAN_SORT": "0"}])
def test_kernel_knobs_keep_parity(env, monkeypatch):
    """The tuning knobs (copy warps, who issues the loads, columns per thread, prefetch depth)
    select different schedules of the same arithmetic: results must not move."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    for name, nt, npat in (("jc69_1000x10k", 120, 900), ("hky_g4_flu", 90, 333)):
        w = make_workload(name, n_tips=nt, n_patterns=npat)
        eng = make_engine(w)
        logl, grad = eng.loglik_grad_host(w.tree.branch_lengths)
        ol, og = oracle_c(w)
        assert rel_err(logl, ol) < 1e-10 and grad_err(grad, og) < 1e-8, (env, name)
        eng.close()

"""Gradients w.r.t. substitution-model / site-model parameters (b200phylo.paramgrad over
b2p_set_gradient_operator) against torch autograd of the whole pruning recursion
(oracle/param_oracle.py).  Tolerance 1e-8 relative to the gradient's scale."""
import numpy as np
import pytest
import torch

from b200phylo.sitemodel import ConstantSiteModel, GammaSiteModel
from b200phylo.substitution import GTR, HKY
from b200phylo.workloads import Workload, make_workload
from helpers import grad_err, make_engine, oracle_c, rel_err
from oracle import param_oracle

pytestmark = pytest.mark.gpu


def _case(kind, n_tips, n_patterns, gamma):
    base = make_workload("hky_g4_flu", n_tips=n_tips, n_patterns=n_patterns)
    pi = np.array([0.31, 0.19, 0.22, 0.28])
    if kind == "gtr":
        exch = np.array([1.3, 2.9, 0.7, 1.1, 3.4, 1.0])
        model = GTR(tuple(exch), tuple(pi))
    else:
        exch = np.array([2.7])  # kappa
        model = HKY(2.7, tuple(pi))
    site = GammaSiteModel(0.6, 4) if gamma else ConstantSiteModel()
    return Workload(kind, base.tree, base.patterns, model, site), exch, pi


@pytest.mark.parametrize("kind,gamma", [("gtr", True), ("hky", True), ("gtr", False)])
def test_parameter_gradients_match_autograd_of_pruning(kind, gamma):
    from b200phylo import paramgrad as pg
    w, exch0, pi0 = _case(kind, 36, 240, gamma)
    eng = make_engine(w)
    dev = torch.device("cuda")
    tree = w.tree

    def leaves(device):
        ex = torch.tensor(exch0, dtype=torch.float64, device=device, requires_grad=True)
        fr = torch.tensor(pi0, dtype=torch.float64, device=device, requires_grad=True)   # unnormalised
        rt = torch.tensor(np.asarray(w.site_model.rates), dtype=torch.float64, device=device, requires_grad=True)
        bl = torch.tensor(tree.branch_lengths, dtype=torch.float64, device=device, requires_grad=True)
        return ex, fr, rt, bl

    def build(ex, fr):
        f = fr / fr.sum()
        e = pg.gtr_exchangeabilities(ex) if kind == "gtr" else pg.hky_exchangeabilities(ex)
        return pg.q_matrix(e, f), f

    # engine route
    ex, fr, rt, bl = leaves(dev)

    def model():
        q, f = build(ex, fr)
        return {"q": q, "freqs": f, "rates": rt}

    root_children = tree.children[tree.root]
    s = pg.parameter_gradients(eng, bl, root_children, model)  # G in one pass, b2p_model_gradients
    s.backward()
    # the round-1 route (16 K evaluations, torch's matrix exponentials) gives the same gradients
    dev_grads = [x.grad.clone() for x in (ex, fr, rt, bl)]
    for x in (ex, fr, rt, bl):
        x.grad = None
    pg.parameter_gradients(eng, bl, root_children, model, g=pg.branch_outer_products(eng, bl, one_pass=False)).backward()
    for a, b in zip(dev_grads, (ex, fr, rt, bl)):
        assert grad_err(a.cpu().numpy(), b.grad.cpu().numpy()) < 1e-10
    # oracle route
    oex, ofr, ort, obl = leaves("cpu")
    q, f = build(oex, ofr)
    ll = param_oracle.log_likelihood(tree.post_triples, w.patterns.partials(), w.patterns.weights, obl, q, f, ort,
                                     np.asarray(w.site_model.weights))
    ll.backward()
    assert rel_err(float(ll.detach()), oracle_c(w)[0]) < 1e-10   # the two oracles agree on the value
    for name, got, want in (("exchangeabilities", ex.grad, oex.grad), ("frequencies", fr.grad, ofr.grad),
                            ("category rates", rt.grad, ort.grad), ("branch lengths", bl.grad, obl.grad)):
        assert grad_err(got.cpu().numpy(), want.numpy()) < 1e-8, (kind, gamma, name, got.cpu().numpy(), want.numpy())
    # the engine is back to branch-length gradients afterwards
    logl, g = eng.loglik_grad_host(tree.branch_lengths)
    assert grad_err(g, obl.grad.numpy()) < 1e-8
    eng.close()


@pytest.mark.parametrize("name,n_tips,n_patterns,gamma", [
    ("hky_g4_flu", 36, 240, True), ("hky_g4_flu", 36, 240, False), ("hky_g4_flu", 7, 5, True), ("hky_g4_flu", 2, 33, False),
    ("hky_g4_flu", 120, 3000, True),      # several tiles per CTA: the parts are added across tiles
    ("jc69_1000x10k", 300, 20011, False),  # more tiles than SMs at one category, ragged last tile
    ("jc69_1000x10k", 50, 5476, False),    # 37 columns per tile: slot boundaries inside a group of four lanes
    ("jc69_1000x10k", 90, 9000, True),     # 4 categories, wide tile
    ("hky_g4_flu", 200, 6007, True)])
def test_outer_products_in_one_pass_match_the_operator_route(name, n_tips, n_patterns, gamma):
    """``b2p_branch_outer_products`` (one evaluation) against the 16 K evaluations of
    ``b2p_set_gradient_operator``: the same sums in another order — 1e-11 of the largest entry
    of each branch's matrix — plus logL / the branch gradient of the outer-products instance
    against the default one (other tile widths: 1e-11 / 1e-9)."""
    from b200phylo import paramgrad as pg
    base = make_workload(name, n_tips=n_tips, n_patterns=n_patterns)
    site = GammaSiteModel(0.45, 4) if gamma else ConstantSiteModel()
    w = Workload("outer", base.tree, base.patterns, HKY(3.1, (0.33, 0.17, 0.24, 0.26)), site)
    eng = make_engine(w)
    bl = torch.tensor(w.tree.branch_lengths, dtype=torch.float64, device="cuda")
    g16 = pg.branch_outer_products(eng, bl, one_pass=False)
    g1 = pg.branch_outer_products(eng, bl, one_pass=True)
    assert g1.shape == g16.shape == (w.n_categories, 2 * n_tips - 2, 4, 4)
    scale = g16.abs().amax(dim=(2, 3), keepdim=True).clamp_min(1e-300)
    assert float(((g1 - g16).abs() / scale).max()) < 1e-11
    # twice in a row gives the same bits (fixed summation order), also after a model change
    assert torch.equal(pg.branch_outer_products(eng, bl), g1)
    eng.set_model(HKY(1.9, (0.33, 0.17, 0.24, 0.26)), site)
    g2 = pg.branch_outer_products(eng, bl)
    g2_16 = pg.branch_outer_products(eng, bl, one_pass=False)
    assert float(((g2 - g2_16).abs() / g2_16.abs().amax(dim=(2, 3), keepdim=True).clamp_min(1e-300)).max()) < 1e-11
    assert not torch.equal(g2, g1)
    # the outer-products instance's own logL and branch gradient
    twin = eng._outer_twin()
    logl0, grad0 = eng.loglik_grad_host(w.tree.branch_lengths)
    logl1, grad1 = twin.loglik_grad_host(w.tree.branch_lengths)
    assert rel_err(logl1, logl0) < 1e-11 and grad_err(grad1, grad0) < 1e-9
    out = torch.zeros(2 * n_tips - 1, dtype=torch.float64, device="cuda")
    g3 = torch.empty_like(g2)
    twin.branch_outer_products(bl.data_ptr(), out.data_ptr(), out.data_ptr() + 8, g3.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert torch.equal(g3, g2)
    assert rel_err(float(out[0]), logl0) < 1e-11 and grad_err(out[1:].cpu().numpy(), grad0) < 1e-9
    eng.close()


@pytest.mark.parametrize("k_cat,ct", [(1, 33), (1, 41), (1, 50), (1, 63), (1, 65), (1, 97), (1, 129), (1, 143),
                                      (4, 36), (4, 44), (4, 52), (4, 68), (4, 100), (4, 132), (4, 140)])
def test_outer_products_for_every_kind_of_slot_boundary(k_cat, ct):
    """Tile widths that put the boundary between two slots on every kind of lane (inside a group of
    four columns at one category, inside either 16-lane half at four): a warp that holds columns of
    two tasks keeps two accumulators and the products that straddle the boundary run once per slot."""
    from b200phylo import paramgrad as pg
    n_patterns = 148 * ct // k_cat  # one tile of ct columns per SM
    base = make_workload("jc69_1000x10k", n_tips=23, n_patterns=n_patterns)
    site = GammaSiteModel(0.7, 4) if k_cat == 4 else ConstantSiteModel()
    w = Workload("outer", base.tree, base.patterns, GTR((1.3, 2.9, 0.7, 1.1, 3.4, 1.0), (0.31, 0.19, 0.22, 0.28)), site)
    eng = make_engine(w)
    assert eng._outer_twin().info()["cols_per_tile"] == ct
    bl = torch.tensor(w.tree.branch_lengths, dtype=torch.float64, device="cuda")
    g1 = pg.branch_outer_products(eng, bl, one_pass=True)
    g16 = pg.branch_outer_products(eng, bl, one_pass=False)
    scale = g16.abs().amax(dim=(2, 3), keepdim=True).clamp_min(1e-300)
    assert float(((g1 - g16).abs() / scale).max()) < 1e-11
    eng.close()


@pytest.mark.parametrize("n_tips,n_patterns,gamma", [(14, 60, True), (14, 60, False), (60, 700, True), (33, 5000, False)])
def test_outer_products_match_the_cpu_oracle(n_tips, n_patterns, gamma):
    """b2p_branch_outer_products against oracle.branch_outer_products (numpy, the analytic two passes
    with q and m kept): 1e-10 of the largest entry of each branch's matrix."""
    from b200phylo import paramgrad as pg
    from oracle import oracle
    base = make_workload("hky_g4_flu", n_tips=n_tips, n_patterns=n_patterns)
    site = GammaSiteModel(0.6, 4) if gamma else ConstantSiteModel()
    w = Workload("g", base.tree, base.patterns, GTR((1.3, 2.9, 0.7, 1.1, 3.4, 1.0), (0.31, 0.19, 0.22, 0.28)), site)
    eng = make_engine(w)
    bl = torch.tensor(w.tree.branch_lengths, dtype=torch.float64, device="cuda")
    got = pg.branch_outer_products(eng, bl, one_pass=True).cpu().numpy()
    want = oracle.branch_outer_products(w.tree.post_triples, w.patterns.partials(), w.patterns.weights, w.tree.branch_lengths,
                                        *w.eigen_args())
    scale = np.abs(want).max(axis=(2, 3), keepdims=True)
    assert np.max(np.abs(got - want) / scale) < 1e-10
    eng.close()


def test_outer_products_unsupported_shapes_say_so():
    from b200phylo.engine import Engine, EngineError
    with pytest.raises(EngineError, match="OUTER_PRODUCTS"):
        Engine(10, 100, 3, outer_products=True)
    with pytest.raises(EngineError, match="OUTER_PRODUCTS"):
        Engine(10, 100, 4, precision="float32", outer_products=True)
    assert Engine(10, 100, 4, precision="float32").supports_outer_products()  # through an fp64 twin
    eng = Engine(10, 100, 4)
    with pytest.raises(EngineError, match="before"):
        eng.branch_outer_products(0, 0, None, 0)
    eng.close()


def test_gamma_shape_gradient_by_chain_rule():
    """d logL / d alpha = sum_k (d logL / d r_k) (d r_k / d alpha), the rates' derivative taken
    numerically on the host (the discretised-Gamma quantiles have no closed form)."""
    from b200phylo import paramgrad as pg
    w, exch0, pi0 = _case("hky", 30, 200, True)
    eng = make_engine(w)
    bl = torch.tensor(w.tree.branch_lengths, dtype=torch.float64, device="cuda")
    rt = torch.tensor(np.asarray(w.site_model.rates), dtype=torch.float64, device="cuda", requires_grad=True)
    e = w.subst_model.eigen()
    qd, fd = torch.tensor(e.q, device="cuda"), torch.tensor(e.freqs, device="cuda")
    pg.parameter_gradients(eng, bl, w.tree.children[w.tree.root], lambda: {"q": qd, "freqs": fd, "rates": rt}).backward()
    h = 1e-6
    dr = (GammaSiteModel(0.6 + h, 4).rates - GammaSiteModel(0.6 - h, 4).rates) / (2 * h)
    d_alpha = float((rt.grad.cpu().numpy() * dr).sum())
    # reference: central difference of the engine's own log-likelihood in alpha

    def logl(alpha):
        eng.set_model(w.subst_model, GammaSiteModel(alpha, 4))
        return eng.loglik_grad_host(w.tree.branch_lengths, want_grad=False)[0]

    fd_alpha = (logl(0.6 + 1e-4) - logl(0.6 - 1e-4)) / 2e-4
    assert d_alpha == pytest.approx(fd_alpha, rel=1e-5)
    eng.close()


@pytest.mark.parametrize("where,site", [("cuda", "weibull"), ("cpu", "weibull"), ("cuda", "constant"), ("cuda", "weibull3")])
def test_tree_likelihood_model_differentiates_model_parameters(where, site):
    """The plugin surface: ``TreeLikelihoodModel`` with ``DifferentiableGTR`` / a Weibull site model
    whose parameters are tensors — ``model().backward()`` fills the gradients w.r.t. branch
    lengths, exchangeabilities, (unnormalised) frequencies and the shape from ONE engine
    evaluation; against autograd of the whole pruning recursion (oracle/param_oracle.py), 1e-8."""
    from b200phylo import paramgrad as pg
    from b200phylo.diffmodels import DifferentiableGTR, DifferentiableWeibullSiteModel
    from b200phylo.treelikelihood import Parameter, TreeLikelihoodModel, UnrootedTreeModel
    n_cat = {"weibull": 4, "weibull3": 3, "constant": 1}[site]  # 3 categories: no one-pass kernel, the operator route
    w, exch0, pi0 = _case("gtr", 40, 300, site == "weibull")
    if site == "weibull3":
        from b200phylo.sitemodel import WeibullSiteModel
        w = Workload("gtr", w.tree, w.patterns, w.subst_model, WeibullSiteModel(0.8, 3))
    tree = w.tree

    def leaves(device):
        mk = lambda v: torch.tensor(v, dtype=torch.float64, device=device, requires_grad=True)  # noqa: E731
        return mk(exch0), mk(pi0), mk([0.8]), mk(tree.branch_lengths)

    ex, fr, shape, bl = leaves(where)
    subst = DifferentiableGTR("gtr", Parameter("gtr.rates", ex), Parameter("gtr.freqs", fr))
    site_model = DifferentiableWeibullSiteModel("site", Parameter("shape", shape), n_cat) if n_cat > 1 else ConstantSiteModel()
    m = TreeLikelihoodModel("like", w.patterns, UnrootedTreeModel("tree", tree, bl), subst, site_model)
    assert [p.id for p in m.parameters()][-3 if n_cat > 1 else -2:][:2] == ["gtr.rates", "gtr.freqs"]
    logl = m()
    assert logl.device.type == where
    (3.0 * logl).backward()
    # oracle: the same parameterisation on the CPU, pruning differentiated by torch
    oex, ofr, oshape, obl = leaves("cpu")
    f = ofr / ofr.sum()
    q = pg.q_matrix(oex, f)
    if n_cat > 1:
        orates = DifferentiableWeibullSiteModel(None, oshape, n_cat).rates_tensor()
        weights = np.full(n_cat, 1.0 / n_cat)
    else:
        orates, weights = torch.ones(1, dtype=torch.float64), np.ones(1)
    ll = param_oracle.log_likelihood(tree.post_triples, w.patterns.partials(), w.patterns.weights, obl, q, f, orates, weights)
    (3.0 * ll).backward()
    assert rel_err(float(logl.detach()), float(ll.detach())) < 1e-10
    pairs = [("exchangeabilities", ex.grad, oex.grad), ("frequencies", fr.grad, ofr.grad), ("branch lengths", bl.grad, obl.grad)]
    if n_cat > 1:
        pairs.append(("shape", shape.grad, oshape.grad))
    for name, got, want in pairs:
        assert got is not None and got.device.type == where, name
        assert grad_err(got.cpu().numpy(), want.numpy()) < 1e-8, (name, got.cpu().numpy(), want.numpy())
    # a changed parameter is seen by the next call (the engine's model is refreshed), with or without gradients
    with torch.no_grad():
        ex[1] = 4.2
    oex2 = oex.detach().clone()
    oex2[1] = 4.2
    ll2 = param_oracle.log_likelihood(tree.post_triples, w.patterns.partials(), w.patterns.weights, obl.detach(),
                                      pg.q_matrix(oex2, f.detach()), f.detach(), orates.detach(), weights)
    assert rel_err(float(m().detach()), float(ll2)) < 1e-10
    with torch.no_grad():
        assert rel_err(float(m()), float(ll2)) < 1e-10
    # only the branch lengths ask for a gradient: the ordinary route, same numbers
    for t in (ex, fr, shape):
        t.requires_grad_(False)
    bl.grad = None
    m().backward()
    assert bl.grad is not None and bl.grad.shape == obl.shape
    m.close()


def test_gamma_shape_through_the_model_object():
    """``DifferentiableGammaSiteModel``: ``model().backward()`` leaves dlogL/dalpha on the shape —
    against a central difference of the model's own log-likelihood in alpha."""
    from b200phylo.diffmodels import DifferentiableGammaSiteModel
    from b200phylo.treelikelihood import Parameter, TreeLikelihoodModel, UnrootedTreeModel
    w, _, _ = _case("hky", 30, 200, True)
    alpha = torch.tensor([0.6], dtype=torch.float64, requires_grad=True)
    bl = torch.tensor(w.tree.branch_lengths, dtype=torch.float64, device="cuda")
    m = TreeLikelihoodModel("like", w.patterns, UnrootedTreeModel("tree", w.tree, bl), w.subst_model,
                            DifferentiableGammaSiteModel("site", Parameter("alpha", alpha), 4))
    m().backward()
    got = float(alpha.grad)

    def logl(a):
        with torch.no_grad():
            alpha[0] = a
            return float(m())

    fd = (logl(0.6 + 1e-4) - logl(0.6 - 1e-4)) / 2e-4
    assert got == pytest.approx(fd, rel=1e-5)
    m.close()

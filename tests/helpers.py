"""Shared helpers for the parity tests (test infrastructure)."""
import numpy as np

from b200phylo.workloads import Workload, make_workload
from oracle import c_oracle, oracle


def oracle_c(w: Workload, bl=None, want_grad=True, n_threads=0):
    bl = w.tree.branch_lengths if bl is None else bl
    return c_oracle.loglik_and_gradient(w.tree.post_triples, w.patterns.masks, w.patterns.weights,
                                        bl, *w.eigen_args(), n_threads=n_threads, want_grad=want_grad)


def oracle_np(w: Workload, bl=None):
    bl = w.tree.branch_lengths if bl is None else bl
    return oracle.loglik_and_gradient(w.tree.post_triples, w.patterns.partials(), w.patterns.weights,
                                      bl, *w.eigen_args())


def oracle_ad(w: Workload, bl=None):
    bl = w.tree.branch_lengths if bl is None else bl
    return oracle.loglik_autograd(w.tree.post_triples, w.patterns.partials(), w.patterns.weights,
                                  bl, *w.eigen_args())


def rel_err(a, b, floor=1e-300):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor)))


def grad_err(g, g_ref):
    """Gradient error relative to the gradient's own scale: entries that are tiny only
    through cancellation are measured against the vector's infinity norm / 1e4."""
    g, g_ref = np.asarray(g), np.asarray(g_ref)
    scale = np.maximum(np.abs(g_ref), np.max(np.abs(g_ref)) * 1e-4)
    return float(np.max(np.abs(g - g_ref) / scale))


def make_engine(w: Workload, precision=None, **kw):
    from b200phylo.engine import Engine
    eng = Engine(w.n_tips, w.n_patterns, w.n_categories, precision=precision or w.precision,
                 n_states=getattr(w.patterns, "n_states", 4), **kw)
    eng.set_topology(w.tree.post_triples)
    eng.set_tip_states(w.patterns.masks)
    eng.set_pattern_weights(w.patterns.weights)
    eng.set_model(w.subst_model, w.site_model)
    return eng

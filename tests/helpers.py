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


def grad_report(g, g_ref):
    """The gradient error both ways: ``floored`` is ``grad_err`` (denominator floored at 1e-4 of the
    largest entry), ``plain`` is the entry-wise relative error |g - g_ref| / |g_ref| with no floor,
    ``n_floored`` counts the entries whose denominator the floor replaced, ``worst_plain_entry``
    is the reference value of the entry with the largest plain error."""
    g, g_ref = np.asarray(g, dtype=np.float64), np.asarray(g_ref, dtype=np.float64)
    big = np.max(np.abs(g_ref))
    floor = big * 1e-4
    nz = np.abs(g_ref) > 0
    plain = np.zeros_like(g_ref)
    plain[nz] = np.abs(g - g_ref)[nz] / np.abs(g_ref)[nz]
    i = int(np.argmax(plain))
    return {"floored": grad_err(g, g_ref), "plain": float(plain.max()), "n_floored": int((np.abs(g_ref) < floor).sum()),
            "n": int(g_ref.size), "worst_plain_entry": float(g_ref[i]), "max_abs": float(big)}

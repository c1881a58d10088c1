"""Generate the golden vectors under tests/golden/ (run once, here, on CPU):

    python tests/golden/make_golden.py

The reference ships no expected values for this path, and none of its
contenders can be imported here, so these vectors come from the numpy oracle
(formulation B), cross-checked at generation time against autograd-of-pruning
(formulation A) to 1e-10.  They pin the oracle against drift and give the GPU parity
tests fixed inputs/outputs that travel to the GPU box.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "gradient-benchmark_b200"))
sys.path.insert(0, ROOT)

from b200phylo.workloads import make_workload  # noqa: E402
from oracle import oracle  # noqa: E402

CASES = [  # (file stem, workload, n_tips, n_patterns, ambiguity fraction)
    ("jc69_n20", "jc69_20x1000", 20, 1000, 0.0),
    ("hky_g4_n40", "hky_g4_flu", 40, 400, 0.0),
    ("gtr_g4_n64", "gtr_g4_5000x10k", 64, 333, 0.0),
    ("hky_g4_n24_ambig", "hky_g4_flu", 24, 200, 0.15),
    ("jc69_n300_deep", "jc69_1000x10k", 300, 150, 0.0),
]


def main():
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for stem, name, n, p, amb in CASES:
        w = make_workload(name, n_tips=n, n_patterns=p)
        masks = w.patterns.masks.copy()
        if amb > 0:
            rng = np.random.default_rng(42)
            idx = rng.random(masks.shape) < amb
            masks[idx] = rng.choice([15, 5, 10, 3, 12, 7, 14], size=int(idx.sum()))
        tp = oracle.tip_partials_from_masks(masks)
        args = (w.tree.post_triples, tp, w.patterns.weights, w.tree.branch_lengths, *w.eigen_args())
        logl, grad = oracle.loglik_and_gradient(*args)
        logl_a, grad_a = oracle.loglik_autograd(*args)
        assert abs(logl - logl_a) <= 1e-12 * abs(logl), (stem, logl, logl_a)
        assert np.max(np.abs(grad - grad_a)) <= 1e-10 * np.max(np.abs(grad)), stem
        evec, ivec, evals, freqs, rates, cat_w = w.eigen_args()
        meta = {"workload": name, "n_tips": w.n_tips, "n_patterns": w.n_patterns,
                "n_categories": w.n_categories, "ambiguity": amb}
        np.savez_compressed(os.path.join(out_dir, stem + ".npz"), triples=w.tree.post_triples,
                            masks=masks, weights=w.patterns.weights, bl=w.tree.branch_lengths,
                            evec=evec, ivec=ivec, evals=evals, freqs=freqs, rates=np.asarray(rates),
                            cat_w=np.asarray(cat_w), logl=np.float64(logl), grad=grad,
                            meta=json.dumps(meta))
        print(stem, meta, logl)


if __name__ == "__main__":
    main()

"""Seeded synthetic workloads for the BASELINE.json configs.

Trees: ``simulate_tree`` (seed 20261104).  Sequences: evolved down the tree with the
config's own model (seed 20261105), ACGT only.  When a config names a *pattern* count
the simulation runs until that many unique columns exist.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import numpy as np

from .alignment import SitePattern, compress_alignment, simulate_alignment, simulate_patterns
from .sitemodel import ConstantSiteModel, GammaSiteModel
from .substitution import GTR, HKY, JC69, CodonGY94, SubstitutionModel
from .tree import Tree, simulate_tree

__all__ = ["Workload", "make_workload", "WORKLOADS", "random_patterns"]

TREE_SEED = 20261104
SEQ_SEED = 20261105


@dataclass
class Workload:
    name: str
    tree: Tree
    patterns: SitePattern
    subst_model: SubstitutionModel
    site_model: object
    precision: str = "float64"

    @property
    def n_tips(self) -> int:
        return self.tree.n_tips

    @property
    def n_patterns(self) -> int:
        return self.patterns.n_patterns

    @property
    def n_categories(self) -> int:
        return len(self.site_model.rates)

    def eigen_args(self):
        e = self.subst_model.eigen()
        return (e.evec, e.ivec, e.evals, e.freqs, self.site_model.rates, self.site_model.weights)


def random_patterns(n_tips: int, n_patterns: int, seed: int = SEQ_SEED, p_change: float = 0.08) -> SitePattern:
    """Cheap surrogate for very large configs: each column is a constant site with a
    sprinkling of substitutions, weights ~ geometric.  Not simulated on the tree, so
    likelihoods are poor, but the kernel's data movement is identical; used only when a
    full simulation would take minutes (flagged in the workload name)."""
    rng = np.random.default_rng(seed)
    base = rng.integers(0, 4, size=n_patterns, dtype=np.uint8)
    states = np.broadcast_to(base, (n_tips, n_patterns)).copy()
    flip = rng.random((n_tips, n_patterns)) < p_change
    states[flip] = rng.integers(0, 4, size=int(flip.sum()), dtype=np.uint8)
    masks = (1 << states).astype(np.uint8)
    weights = rng.geometric(0.5, size=n_patterns).astype(np.float64)
    return SitePattern([f"t{i}_{2000 + i % 14}" for i in range(n_tips)], masks, weights)


def _simulated(name, n_tips, model, site_model, n_patterns: Optional[int] = None,
               n_sites: Optional[int] = None, shape="coalescent", precision="float64",
               mean_branch=0.02, seq_seed: int = SEQ_SEED) -> Workload:
    tree = simulate_tree(n_tips, shape=shape, seed=TREE_SEED, mean_branch=mean_branch)
    rates = np.asarray(site_model.rates)
    pm = model.p_t(tree.branch_lengths[:, None] * rates[None, :])
    if n_patterns is not None:
        sp = simulate_patterns(tree, pm, model.freqs, n_patterns, seed=seq_seed,
                               cat_weights=site_model.weights)
    else:
        aln = simulate_alignment(tree, pm, model.freqs, n_sites, seed=seq_seed,
                                 cat_weights=site_model.weights)
        sp = compress_alignment(aln)
    return Workload(name, tree, sp, model, site_model, precision)


def make_workload(name: str, precision: Optional[str] = None, n_tips: Optional[int] = None,
                  n_patterns: Optional[int] = None, seq_seed: int = SEQ_SEED) -> Workload:
    """Build one of the named workloads; ``n_tips`` / ``n_patterns`` shrink it for tests;
    ``seq_seed`` draws a different alignment on the same tree (per-rank shards)."""
    pi = (0.3, 0.2, 0.2, 0.3)
    if name == "jc69_20x1000":  # BASELINE.json configs[0]
        w = _simulated(name, n_tips or 20, JC69(), ConstantSiteModel(), n_sites=n_patterns or 1000, seq_seed=seq_seed)
    elif name == "jc69_1000x10k":  # configs[1] — the headline metric
        w = _simulated(name, n_tips or 1000, JC69(), ConstantSiteModel(), n_patterns=n_patterns or 10000, seq_seed=seq_seed)
    elif name == "hky_g4_flu":  # configs[2]
        w = _simulated(name, n_tips or 1704, HKY(3.0, pi), GammaSiteModel(0.5, 4),
                       n_sites=n_patterns or 1701, seq_seed=seq_seed)
    elif name == "gtr_g4_5000x100k":  # configs[3] (per-rank slices are cut by the caller)
        w = _simulated(name, n_tips or 5000, GTR((1, 2, 1, 1, 2, 1), pi), GammaSiteModel(0.5, 4),
                       n_patterns=n_patterns or 100000, seq_seed=seq_seed)
    elif name == "gtr_g4_5000x10k":  # north-star gate shape
        w = _simulated(name, n_tips or 5000, GTR((1, 2, 1, 1, 2, 1), pi), GammaSiteModel(0.5, 4),
                       n_patterns=n_patterns or 10000, seq_seed=seq_seed)
    elif name == "jc69_5000x10k":  # north-star gate shape, one rate category
        w = _simulated(name, n_tips or 5000, JC69(), ConstantSiteModel(), n_patterns=n_patterns or 10000, seq_seed=seq_seed)
    elif name == "codon_500x2000":  # configs[4]: 61-state GY94-style, dense state contraction
        w = _simulated(name, n_tips or 500, CodonGY94(2.0, 0.5, pi), ConstantSiteModel(),
                       n_patterns=n_patterns or 2000, seq_seed=seq_seed, mean_branch=0.05)
    else:
        raise ValueError(f"unknown workload {name!r}; known: {sorted(WORKLOADS)}")
    if precision:
        w.precision = precision
    return w


WORKLOADS = {
    "jc69_20x1000": "JC69, 20 taxa x 1000 sites, fp64 (micro.nf shape)",
    "jc69_1000x10k": "JC69, 1000 taxa x 10k patterns, fp64",
    "hky_g4_flu": "HKY+G4, 1704 taxa x 1701 sites",
    "gtr_g4_5000x100k": "GTR+G4, 5000 taxa x 100k patterns",
    "gtr_g4_5000x10k": "GTR+G4, 5000 taxa x 10k patterns",
    "jc69_5000x10k": "JC69, 5000 taxa x 10k patterns, fp64",
    "codon_500x2000": "GY94 codon model (61 states), 500 taxa x 2000 codon patterns",
}

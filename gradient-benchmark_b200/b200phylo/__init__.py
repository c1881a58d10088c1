"""b200phylo — B200-native tree log-likelihood + branch-length gradient engine.

Host side (Python) over ``libb200phylo.so`` (C ABI, ``include/b200phylo.h``; hand-written
sm_100a kernels in ``../csrc``).  No CPU fallback.
"""
from .alignment import Alignment, SitePattern, compress_alignment, read_fasta, write_fasta  # noqa: F401
from .sitemodel import ConstantSiteModel, GammaSiteModel, WeibullSiteModel  # noqa: F401
from .substitution import GTR, HKY, JC69, GeneralReversible  # noqa: F401
from .tree import Tree, parse_newick, simulate_tree  # noqa: F401

__version__ = "0.1.0"


def __getattr__(name):  # lazy: these import torch / load the CUDA library
    if name in ("Engine", "EngineError"):
        from . import engine
        return getattr(engine, name)
    if name in ("TreeLikelihoodModel", "UnrootedTreeModel", "StrictClockModel"):
        from . import treelikelihood
        return getattr(treelikelihood, name)
    raise AttributeError(name)

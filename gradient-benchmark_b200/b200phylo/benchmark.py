"""``b200phylo-benchmark`` — the micro-benchmark CLI contract.

Honours the flags the pipeline passes to every ``<program>-benchmark``
(/root/reference/modules/micro.nf:47-53; physher form :20-24):

    -i <fasta>  -t <newick>  -r <replicates>  -s <branch-length scaler>  -o <csv>
    [--gtr]  [-d float32]

and writes the CSV whose header, once ``program,size,rep,precision,`` is prepended by
the pipeline's awk (/root/reference/modules/micro.nf:54-56), yields the columns
``plot.Rmd`` reads: ``function,mode,JIT,time`` (/root/reference/plot.Rmd:257-266).
Rows: ``treelikelihoodJC69,{evaluation,gradient},off,<seconds for all replicates>``
and, with ``--gtr``, ``treelikelihoodGTR,...`` (/root/reference/plot.Rmd:262,305) — whose
gradient row differentiates w.r.t. the branch lengths AND the GTR exchangeabilities and
frequencies (held as parameters, as in torchtree), from one outer-products evaluation.

When the tip labels carry sampling dates (``name_date``, /root/reference/bin/helper.py:28-47)
the time-tree functions of the micro benchmark are timed too
(/root/reference/plot.Rmd:49-60,301-303): ``ratio_transform``, ``ratio_transform_jacobian``
and ``coalescent`` x ``evaluation|gradient``.

Like the other contenders each replicate is one call of the public model object: branch
lengths in, log-likelihood (and ``backward()`` for the gradient rows) out.  Branch
lengths live on the GPU across replicates, as torchtree's parameters live in torch
tensors across replicates.  An extra, separate file ``<csv>.b200.json`` carries
evaluations/s and achieved GB/s (never mixed into the contract CSV).
"""
from __future__ import annotations

import argparse
import json
import sys
import time
from typing import List, Optional

import numpy as np
import torch

from .alignment import compress_alignment, read_fasta
from .sitemodel import ConstantSiteModel
from .diffmodels import DifferentiableGTR
from .substitution import JC69
from .timetree import (ConstantCoalescentModel, ReparameterizedTimeTreeModel, tip_heights_from_names)
from .tree import Tree
from .treelikelihood import Parameter, TreeLikelihoodModel, UnrootedTreeModel

__all__ = ["main", "run_benchmark"]


def _time_loop(fn, replicates: int) -> float:
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(replicates):
        fn()
    torch.cuda.synchronize()
    return time.perf_counter() - t0


def run_benchmark(fasta: str, newick: str, replicates: int, scaler: float, gtr: bool,
                  dtype: str = "float64", device: int = 0, warmup: int = 3) -> List[dict]:
    aln = read_fasta(fasta)
    with open(newick) as fh:
        tree = Tree.from_newick(fh.read(), taxa=aln.names)
    time_tree = tree
    try:
        tip_heights_from_names(tree.names)
        dated = True
    except ValueError:
        dated = False
    tree = tree.scaled(scaler)
    sp = compress_alignment(aln)
    dev = torch.device(f"cuda:{device}")
    torch.cuda.set_device(dev)
    rows = []
    models = [("treelikelihoodJC69", JC69())]
    gtr_params = []
    if gtr:
        # torchtree's GTR holds its exchangeabilities and frequencies as parameters and the --gtr
        # gradient rows differentiate w.r.t. them as well as the branch lengths (by autograd of the
        # pruning recursion there; here: one outer-products evaluation — in fp64 also under
        # -d float32: the outer-products instance has no fp32 form).
        gtr_params = [Parameter("gtr.rates", torch.tensor([1.0, 2.0, 1.0, 1.0, 2.0, 1.0], dtype=torch.float64, device=dev)),
                      Parameter("gtr.frequencies", torch.tensor([0.3, 0.2, 0.2, 0.3], dtype=torch.float64, device=dev))]
        models.append(("treelikelihoodGTR", DifferentiableGTR("gtr", *gtr_params)))
    for fname, subst in models:
        own = gtr_params if fname == "treelikelihoodGTR" else []
        bl = torch.tensor(tree.branch_lengths, dtype=torch.float64, device=dev)
        tm = UnrootedTreeModel("tree", tree, bl)
        model = TreeLikelihoodModel("like", sp, tm, subst, ConstantSiteModel(), precision=dtype,
                                    device=device)

        def evaluation():
            with torch.no_grad():
                return model()

        def gradient():
            bl.grad = None
            for p in own:
                p.tensor.grad = None
            model().backward()

        bl.requires_grad_(False)
        for _ in range(warmup):
            evaluation()
        t_eval = _time_loop(evaluation, replicates)
        logl = float(evaluation())
        bl.requires_grad_(True)
        for p in own:
            p.requires_grad = True
        for _ in range(warmup):
            gradient()
        t_grad = _time_loop(gradient, replicates)
        info = model.engine.info()
        rows.append({"function": fname, "mode": "evaluation", "JIT": "off", "time": t_eval,
                     "logl": logl, "alg_bytes": info["alg_bytes_loglik"]})
        rows.append({"function": fname, "mode": "gradient", "JIT": "off", "time": t_grad,
                     "logl": logl, "alg_bytes": info["alg_bytes_grad"],
                     "grad_norm": float(bl.grad.norm()),
                     **({"grad_wrt": "branch lengths, GTR rates, frequencies",
                         "grad_rates": own[0].tensor.grad.cpu().tolist(), "grad_frequencies": own[1].tensor.grad.cpu().tolist()}
                        if own else {})})
        model.close()
    if dated:
        rows += run_time_functions(time_tree, replicates, device, warmup)
    for r in rows:
        r.update(n_taxa=tree.n_tips, n_patterns=sp.n_patterns, replicates=replicates, dtype=dtype)
    return rows


def run_time_functions(tree: Tree, replicates: int, device: int = 0, warmup: int = 3) -> List[dict]:
    """ratio_transform (ratios/root height -> internal heights, derivatives w.r.t. ratios),
    ratio_transform_jacobian (heights -> log|det J|, derivatives w.r.t. heights) and coalescent
    (heights + population size -> log density, derivatives w.r.t. both)
    (/root/reference/plot.Rmd:49-60).  One kernel launch per evaluation, one more per gradient."""
    tm = ReparameterizedTimeTreeModel("tree", tree, None, device=device)
    dev = torch.device(f"cuda:{device}")
    ratios, root = tm.ratios, tm.root_height
    heights = tm.heights().detach()
    theta = torch.tensor([4.0], dtype=torch.float64, device=dev)
    coal = ConstantCoalescentModel("coalescent", tm, theta)
    ones = torch.ones_like(heights)
    rows = []

    def both(fname, evaluation, gradient, leaves):
        for x in leaves:
            x.requires_grad_(False)
        for _ in range(warmup):
            evaluation()
        t_eval = _time_loop(evaluation, replicates)
        for x in leaves:
            x.requires_grad_(True)
        for _ in range(warmup):
            gradient()
        t_grad = _time_loop(gradient, replicates)
        for x in leaves:
            x.requires_grad_(False)
            x.grad = None
        rows.append({"function": fname, "mode": "evaluation", "JIT": "off", "time": t_eval, "alg_bytes": 0})
        rows.append({"function": fname, "mode": "gradient", "JIT": "off", "time": t_grad, "alg_bytes": 0})

    def transform_grad():
        ratios.grad = None
        root.grad = None
        tm.heights().backward(ones)

    def jacobian_grad():
        heights.grad = None
        tm.log_jacobian(heights).backward()

    def coalescent_grad():
        heights.grad = None
        theta.grad = None
        coal(heights).backward()

    both("ratio_transform", lambda: tm.heights(), transform_grad, [ratios, root])
    both("ratio_transform_jacobian", lambda: tm.log_jacobian(heights), jacobian_grad, [heights])
    both("coalescent", lambda: coal(heights), coalescent_grad, [heights, theta])
    tm.close()
    return rows


def main(argv: Optional[List[str]] = None) -> int:
    ap = argparse.ArgumentParser(prog="b200phylo-benchmark", description=__doc__.split("\n")[0])
    ap.add_argument("-i", "--input", required=True, help="alignment (FASTA)")
    ap.add_argument("-t", "--tree", required=True, help="tree (Newick, name_date tips)")
    ap.add_argument("-r", "--replicates", type=int, required=True)
    ap.add_argument("-s", "--scaler", type=float, default=1.0, help="branch-length scaler")
    ap.add_argument("-o", "--output", required=True, help="output CSV")
    ap.add_argument("-d", "--dtype", choices=["float64", "float32"], default="float64")
    ap.add_argument("--gtr", action="store_true", help="also time a GTR tree likelihood")
    ap.add_argument("--device", type=int, default=0)
    a = ap.parse_args(argv)
    if not torch.cuda.is_available():
        print("b200phylo-benchmark: no CUDA device; this engine has no CPU fallback", file=sys.stderr)
        return 2
    rows = run_benchmark(a.input, a.tree, a.replicates, a.scaler, a.gtr, a.dtype, a.device)
    with open(a.output, "w") as fh:
        fh.write("function,mode,JIT,time\n")
        for r in rows:
            fh.write(f"{r['function']},{r['mode']},{r['JIT']},{r['time']}\n")
    extra = []
    for r in rows:
        e = dict(r)
        e["evals_per_s"] = r["replicates"] / r["time"]
        e["achieved_gbs"] = r["alg_bytes"] * r["replicates"] / r["time"] / 1e9
        extra.append(e)
    with open(a.output + ".b200.json", "w") as fh:
        json.dump(extra, fh, indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())

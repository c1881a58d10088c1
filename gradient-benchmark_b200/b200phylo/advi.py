"""``b200phylo-advi`` — the macro benchmark's run as one program.

The reference times, per contender, 5000 iterations of mean-field ADVI on a dated flu tree:
JC69, strict clock, constant-size coalescent, one ELBO sample per iteration
(/root/reference/modules/macro_flu.nf:6,108-157; model: flu_H3N2/phylotorch-JC69.template).
For torchtree that is ``torchtree-cli advi ... > model.json`` then ``torchtree model.json``;
this CLI takes the same flags (/root/reference/modules/macro_flu.nf:124-137) and runs the
loop directly.  The variational family, optimiser and priors follow the template
(Normal mean-field on the unconstrained parameters :125-230, Adam ``lr`` :104-110, OneOnX on
theta :86-91, Exponential(1000) on the rate :92-100).  Progress lines are written in the
format ``bin/time-parser.py:60-61`` parses for torchtree (``^\\s+(\\d+)\\s+(-\\d+\\.\\d+).+``), so
``macro.csv`` can carry this engine as one more program.

Only the model density and its gradient are this package's business: every iteration is one
``b2p_joint_density_grad`` call (nine kernel launches, no host round trip) plus a handful of
elementwise torch ops for the reparameterisation and Adam.  No CPU fallback.
"""
from __future__ import annotations

import argparse
import math
import sys
import time
from typing import List, Optional

import numpy as np
import torch

from .alignment import compress_alignment, read_fasta
from .sitemodel import ConstantSiteModel
from .substitution import JC69
from .timetree import (ConstantCoalescentModel, JointTimeTreeDensity, ReparameterizedTimeTreeModel,
                       initial_params_from_tree, tip_heights_from_names)
from .tree import Tree
from .treelikelihood import Parameter, StrictClockModel, TreeLikelihoodModel

__all__ = ["main", "run_advi"]


class _Holder:
    """A tensor slot the model objects read through ``.tensor`` (rebound every iteration)."""

    def __init__(self, t: torch.Tensor):
        self.tensor = t

    def parameters(self):
        return []


def run_advi(fasta: str, newick: str, iterations: int, lr: float = 1e-4, rate_init: float = 0.001,
             theta_init: float = 3.0, seed: int = 1, print_every: int = 100, device: int = 0,
             out=sys.stdout, precision: str = "float64", cuda_graph: bool = True) -> dict:
    aln = read_fasta(fasta)
    with open(newick) as fh:
        tree = Tree.from_newick(fh.read(), taxa=aln.names)
    sp = compress_alignment(aln)
    n = tree.n_tips
    dev = torch.device(f"cuda:{device}")
    torch.cuda.set_device(dev)
    torch.manual_seed(seed)
    tips = tip_heights_from_names(tree.names)
    x0 = initial_params_from_tree(tree, tips)  # --heights_init tree
    shift = float(tips.max())
    # unconstrained starting point: logit(ratios), log(root - shift), log(rate), log(theta)
    r0 = np.clip(x0[: n - 2], 1e-6, 1 - 1e-6)
    z0 = np.concatenate([np.log(r0) - np.log1p(-r0), [math.log(max(x0[n - 2] - shift, 1e-6))],
                         [math.log(rate_init)], [math.log(theta_init)]])
    dim = z0.shape[0]
    mu = torch.tensor(z0, dtype=torch.float64, device=dev, requires_grad=True)
    log_sigma = torch.full((dim,), math.log(1e-3), dtype=torch.float64, device=dev, requires_grad=True)

    ratios, root, rate, theta = (_Holder(torch.empty(0, device=dev)) for _ in range(4))
    tm = ReparameterizedTimeTreeModel("tree", tree, tips, ratios, root, device=device)
    like = TreeLikelihoodModel("like", sp, tm, JC69(), ConstantSiteModel(), StrictClockModel("clock", torch.empty(0)),
                               precision=precision, device=device)
    coal = ConstantCoalescentModel("coalescent", tm, theta)
    joint = JointTimeTreeDensity(like, tm, coal, include_jacobian=True)
    opt = torch.optim.Adam([mu, log_sigma], lr=lr, maximize=True, capturable=cuda_graph)
    const_entropy = 0.5 * dim * (1.0 + math.log(2.0 * math.pi))

    def elbo_sample() -> torch.Tensor:
        z = mu + torch.exp(log_sigma) * torch.randn(dim, dtype=torch.float64, device=dev)
        zr, zh, zc, zt = z[: n - 2], z[n - 2: n - 1], z[n - 1: n], z[n: n + 1]
        ratios.tensor = torch.sigmoid(zr)
        root.tensor = shift + torch.exp(zh)
        like.clock_model._rate = Parameter("rate", torch.exp(zc))
        theta.tensor = torch.exp(zt)
        # log|d constrained / d unconstrained|: sigmoid, exp, exp, exp
        log_det = (torch.nn.functional.logsigmoid(zr) + torch.nn.functional.logsigmoid(-zr)).sum() + zh.sum() + zc.sum() + zt.sum()
        log_prior = -zt.sum() + (math.log(1000.0) - 1000.0 * torch.exp(zc)).sum()  # OneOnX(theta), Exponential(1000)(rate)
        return joint() + log_det + log_prior + log_sigma.sum() + const_entropy

    def iteration() -> torch.Tensor:
        opt.zero_grad(set_to_none=True)
        elbo = elbo_sample()
        elbo.backward()
        opt.step()
        return elbo.detach()

    out.write(f"  {'iter':>6}  {'ELBO':>16}  {'seconds':>10}\n")
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    elbo_val = float("nan")
    history = []

    def report(it: int, elbo: torch.Tensor) -> None:
        nonlocal elbo_val
        if it % print_every == 0 or it == 1 or it == iterations:
            elbo_val = float(elbo)
            history.append((it, elbo_val))
            out.write(f"  {it:6d}  {elbo_val:16.3f}  {time.perf_counter() - t0:10.3f}\n")
            out.flush()

    # One iteration = sample, density + gradient (b2p_joint_density_grad), Adam.  Its GPU work is
    # a few hundred microseconds; issued call by call the Python interpreter is the bottleneck.
    # So after three eager iterations the whole iteration is captured once into a CUDA graph
    # (the engine launches on torch's current stream, so its kernels are captured like torch's
    # own) and the remaining iterations are graph replays.
    it = 0
    graph = None
    if cuda_graph and iterations > 4:
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(3):
                it += 1
                report(it, iteration())
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        it += 1
        with torch.cuda.graph(graph):
            static_elbo = iteration()
        graph.replay()  # capture does not execute: this is iteration `it`
        report(it, static_elbo)
    while it < iterations:
        it += 1
        if graph is not None:
            graph.replay()
            report(it, static_elbo)
        else:
            report(it, iteration())
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    res = {"iterations": iterations, "seconds": wall, "cuda_graph": graph is not None, "iters_per_s": iterations / wall, "elbo": elbo_val,
           "history": history, "n_taxa": n, "n_patterns": sp.n_patterns,
           "rate": float(torch.exp(mu[n - 1]).detach()), "theta": float(torch.exp(mu[n]).detach()),
           "root_height": shift + float(torch.exp(mu[n - 2]).detach())}
    like.close()
    tm.close()
    return res


def main(argv: Optional[List[str]] = None) -> int:
    ap = argparse.ArgumentParser(prog="b200phylo-advi", description=__doc__.split("\n")[0])
    ap.add_argument("-i", "--input", required=True, help="alignment (FASTA, name_date taxa)")
    ap.add_argument("-t", "--tree", required=True, help="dated tree (Newick)")
    ap.add_argument("--clock", choices=["strict"], default="strict")
    ap.add_argument("--coalescent", choices=["constant"], default="constant")
    ap.add_argument("--heights_init", choices=["tree"], default="tree")
    ap.add_argument("--rate_init", type=float, default=0.001)
    ap.add_argument("--clockpr", choices=["exponential"], default="exponential")
    ap.add_argument("--lr", type=float, default=1e-4)
    ap.add_argument("--elbo_samples", type=int, default=1)
    ap.add_argument("--tol_rel_obj", type=float, default=0.0)
    ap.add_argument("--iter", type=int, default=5000)
    ap.add_argument("--samples", type=int, default=0)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--print_every", type=int, default=100)
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--no_cuda_graph", action="store_true", help="issue every iteration call by call")
    ap.add_argument("-d", "--dtype", choices=["float64", "float32"], default="float64")
    a = ap.parse_args(argv)
    if a.elbo_samples != 1:
        ap.error("one ELBO sample per iteration (the reference's setting, macro_flu.nf:133)")
    if not torch.cuda.is_available():
        print("b200phylo-advi: no CUDA device; this engine has no CPU fallback", file=sys.stderr)
        return 2
    res = run_advi(a.input, a.tree, a.iter, a.lr, a.rate_init, seed=a.seed, print_every=a.print_every,
                   device=a.device, precision=a.dtype, cuda_graph=not a.no_cuda_graph)
    sys.stdout.write(f"Ran inference for {res['iterations']} iterations in {res['seconds']:.3f} s "
                     f"({res['iters_per_s']:.1f} iterations/s)\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())

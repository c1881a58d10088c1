"""Under torchrun (N GPUs): TreeLikelihoodModel with tensor-valued GTR / Weibull parameters on
pattern SHARDS — gradients w.r.t. exchangeabilities, frequencies, shape and branch lengths after
the all-reduce of logL, branch gradient and outer products — against the same model on the whole
alignment on one GPU.    torchrun --nproc-per-node N tests/tools/paramgrad_multi_check.py"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "gradient-benchmark_b200"), ROOT, os.path.join(ROOT, "tests")]
from b200phylo.diffmodels import DifferentiableGTR, DifferentiableWeibullSiteModel  # noqa: E402
from b200phylo.treelikelihood import Parameter, TreeLikelihoodModel, UnrootedTreeModel  # noqa: E402
from b200phylo.workloads import make_workload  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
dev = torch.device("cuda", torch.cuda.current_device())
w = make_workload("hky_g4_flu", n_tips=120, n_patterns=3000)


def build(**kw):
    mk = lambda v: torch.tensor(v, dtype=torch.float64, device=dev, requires_grad=True)  # noqa: E731
    ex, fr, shape, bl = mk([1.3, 2.9, 0.7, 1.1, 3.4, 1.0]), mk([0.31, 0.19, 0.22, 0.28]), mk([0.8]), mk(w.tree.branch_lengths)
    m = TreeLikelihoodModel("like", w.patterns, UnrootedTreeModel("tree", w.tree, bl),
                            DifferentiableGTR("gtr", Parameter("r", ex), Parameter("f", fr)),
                            DifferentiableWeibullSiteModel("site", Parameter("shape", shape), 4), **kw)
    logl = m()
    logl.backward()
    out = (float(logl.detach()), [x.grad.detach().cpu().numpy().copy() for x in (ex, fr, shape, bl)])
    m.close()
    return out


sharded = build()                      # the process group's shards
whole = build(rank=0, world_size=1)    # every rank: the whole alignment on its own GPU
ok = abs(sharded[0] - whole[0]) <= 1e-11 * abs(whole[0])
worst = 0.0
for a, b in zip(sharded[1], whole[1]):
    worst = max(worst, float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300)))
ok = ok and worst < 1e-9
print(f"rank {rank}/{world}: logL {sharded[0]:.10f} vs {whole[0]:.10f}, worst gradient difference {worst:.2e} -> {'OK' if ok else 'FAIL'}", flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)

"""Dense-state path: how much of an evaluation is the tree's depth (levels with fewer tasks than
the GPU has CTA slots run at one task latency per level)?  Same alignment size on trees of
different shape.    python tests/tools/dense_depth_probe.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "gradient-benchmark_b200"), ROOT, os.path.join(ROOT, "tests")]
from b200phylo.tree import simulate_tree  # noqa: E402
from b200phylo.workloads import Workload, make_workload  # noqa: E402
from helpers import make_engine  # noqa: E402

base = make_workload("codon_500x2000")
for shape in ("as benchmarked", "balanced", "coalescent", "caterpillar"):
    if shape == "as benchmarked":
        tree = base.tree
    else:
        tree = simulate_tree(base.n_tips, shape=shape, seed=5)
        tree.branch_lengths[:] = np.resize(base.tree.branch_lengths, tree.branch_lengths.shape)
    w = Workload("codon", tree, base.patterns, base.subst_model, base.site_model)
    eng = make_engine(w)
    bl = torch.tensor(tree.branch_lengths, dtype=torch.float64, device="cuda")
    out = torch.empty(1 + eng.n_branches, dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        eng.loglik_grad_device(bl.data_ptr(), out.data_ptr(), out.data_ptr() + 8, st)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        eng.loglik_grad_device(bl.data_ptr(), out.data_ptr(), out.data_ptr() + 8, st)
    e1.record()
    torch.cuda.synchronize()
    i = eng.info()
    print(f"{shape:15s} tree depth {i['tree_depth']:4d}: {e0.elapsed_time(e1) / 20:.3f} ms per evaluation", flush=True)
    eng.close()

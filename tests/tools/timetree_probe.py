"""Launch the height kernels a few times on one tree (for ncu): python timetree_probe.py N shape"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "gradient-benchmark_b200"), ROOT]
from b200phylo.timetree import TimeTreeEngine  # noqa: E402
from b200phylo.tree import simulate_tree  # noqa: E402

n, shape = int(sys.argv[1]), sys.argv[2]
tree = simulate_tree(n, shape=shape, seed=11)
rng = np.random.default_rng(3)
tips = rng.uniform(0.0, 40.0, n)
x = rng.uniform(0.05, 0.95, n - 1)
x[n - 2] = 50.0
eng = TimeTreeEngine(tree.post_triples, tips)
px = torch.tensor(x, device="cuda")
g = torch.ones(n - 1, dtype=torch.float64, device="cuda")
for _ in range(3):
    h = eng.heights(px)
    eng.heights_vjp(px, h, g)
    eng.constant_coalescent(h, torch.tensor([4.0], dtype=torch.float64, device="cuda"))
torch.cuda.synchronize()
print("ok", float(h.sum()))

"""Timing of b2p_branch_outer_products next to the plain gradient on the same (outer-products)
instance — used with the debug build's B2P_DBG switches to see where the extra time goes.
    python tests/tools/outer_probe.py gtr_1000x10k|hky_g4_flu|gtr_g4_5000x10k ..."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "gradient-benchmark_b200"), ROOT, os.path.join(ROOT, "tests")]
from b200phylo.sitemodel import ConstantSiteModel  # noqa: E402
from b200phylo.substitution import GTR  # noqa: E402
from b200phylo.workloads import Workload, make_workload  # noqa: E402
from helpers import make_engine  # noqa: E402


def timed(fn, n=30, warm=5):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for name in sys.argv[1:]:
    if name == "gtr_1000x10k":
        b = make_workload("jc69_1000x10k")
        w = Workload(name, b.tree, b.patterns, GTR((1.3, 2.9, 0.7, 1.1, 3.4, 1.0), (0.31, 0.19, 0.22, 0.28)), ConstantSiteModel())
    else:
        w = make_workload(name)
    eng = make_engine(w, outer_products=True)
    bl = torch.tensor(w.tree.branch_lengths, dtype=torch.float64, device="cuda")
    out = torch.empty(1 + eng.n_branches, dtype=torch.float64, device="cuda")
    g = torch.empty(eng.n_categories, eng.n_branches, 4, 4, dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    plain = timed(lambda: eng.loglik_grad_device(bl.data_ptr(), out.data_ptr(), out.data_ptr() + 8, st))
    outer = timed(lambda: eng.branch_outer_products(bl.data_ptr(), out.data_ptr(), out.data_ptr() + 8, g.data_ptr(), st))
    i = eng.info()
    print(f"{name}: plain {plain:.4f} ms, outer products {outer:.4f} ms (x{outer / plain:.2f}); tile {i['cols_per_tile']} slots {i['tasks_per_round']} "
          f"rounds {i['post_rounds']}+{i['pre_rounds']} pool {i['pool_blocks']} depth {i['copy_depth']} regs {i['regs_per_thread']}", flush=True)
    eng.close()

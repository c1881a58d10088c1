"""Cost of the all-parameter gradient at full sizes: the outer products G in one pass
(b2p_branch_outer_products) and by 16 K evaluations (b2p_set_gradient_operator), the 4x4
algebra on top, and one plain branch-gradient evaluation as the yardstick (device events).
    python tests/tools/paramgrad_bench.py [out.json]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "gradient-benchmark_b200"), ROOT, os.path.join(ROOT, "tests")]
from b200phylo import paramgrad as pg  # noqa: E402
from b200phylo.sitemodel import ConstantSiteModel  # noqa: E402
from b200phylo.substitution import GTR  # noqa: E402
from b200phylo.workloads import Workload, make_workload  # noqa: E402
from helpers import make_engine  # noqa: E402

rows = []
for name in ("gtr_1000x10k", "hky_g4_flu", "gtr_g4_5000x10k"):
    if name == "gtr_1000x10k":
        b = make_workload("jc69_1000x10k")
        w = Workload(name, b.tree, b.patterns, GTR((1.3, 2.9, 0.7, 1.1, 3.4, 1.0), (0.31, 0.19, 0.22, 0.28)), ConstantSiteModel())
    else:
        w = make_workload(name)
    eng = make_engine(w)
    dev = torch.device("cuda")
    e = w.subst_model.eigen()
    ex = torch.tensor(w.subst_model.rates, dtype=torch.float64, device=dev, requires_grad=True)
    fr = torch.tensor(e.freqs, dtype=torch.float64, device=dev, requires_grad=True)
    rt = torch.tensor(np.asarray(w.site_model.rates), dtype=torch.float64, device=dev, requires_grad=True)
    bl = torch.tensor(w.tree.branch_lengths, dtype=torch.float64, device=dev)

    def model():
        f = fr / fr.sum()
        return {"q": pg.q_matrix(ex, f), "freqs": f, "rates": rt}

    def run():
        for t in (ex, fr, rt):
            t.grad = None
        pg.parameter_gradients(eng, bl, w.tree.children[w.tree.root], model).backward()

    def timed(fn, n=10, warm=3):
        for _ in range(warm):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n, (time.perf_counter() - t0) / n * 1e3

    out = torch.empty(1 + eng.n_branches, dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    ms_eval, _ = timed(lambda: eng.loglik_grad_device(bl.data_ptr(), out.data_ptr(), out.data_ptr() + 8, stream), n=50)
    g_buf = torch.empty(eng.n_categories, eng.n_branches, 4, 4, dtype=torch.float64, device=dev)
    eng.branch_outer_products(bl.data_ptr(), out.data_ptr(), None, g_buf.data_ptr(), stream)  # sets the twin up
    ms_g1_dev, _ = timed(lambda: eng.branch_outer_products(bl.data_ptr(), out.data_ptr(), out.data_ptr() + 8, g_buf.data_ptr(), stream), n=50)
    twin = eng._outer_twin()
    ms_twin_eval, _ = timed(lambda: twin.loglik_grad_device(bl.data_ptr(), out.data_ptr(), out.data_ptr() + 8, stream), n=50)
    ms_g16_dev, ms_g16 = timed(lambda: pg.branch_outer_products(eng, bl, one_pass=False), n=3, warm=1)
    ms_dev, ms = timed(run, n=20)
    mg = torch.empty(20 + eng.n_categories, dtype=torch.float64, device=dev)
    root_child = int(w.tree.children[w.tree.root][0])

    def device_part():
        eng.branch_outer_products(bl.data_ptr(), out.data_ptr(), out.data_ptr() + 8, g_buf.data_ptr(), stream)
        eng.model_gradients(bl.data_ptr(), g_buf.data_ptr(), root_child, mg.data_ptr(), stream)

    ms_all_dev, _ = timed(device_part, n=50)
    g1 = pg.branch_outer_products(eng, bl)
    g16 = pg.branch_outer_products(eng, bl, one_pass=False)
    err = float(((g1 - g16).abs() / g16.abs().amax(dim=(2, 3), keepdim=True).clamp_min(1e-300)).max())
    rows.append({"workload": name, "n_taxa": w.n_tips, "n_patterns": w.n_patterns, "categories": w.n_categories,
                 "ms_one_branch_gradient_evaluation": ms_eval,
                 "ms_branch_gradient_on_the_outer_products_instance": ms_twin_eval,
                 "ms_outer_products_one_pass": ms_g1_dev, "one_pass_over_one_evaluation": ms_g1_dev / ms_eval,
                 "ms_outer_products_16K_evaluations": ms_g16_dev, "evaluations_16K": 16 * w.n_categories,
                 "ms_outer_products_and_model_gradients": ms_all_dev, "all_gradients_over_one_evaluation": ms_all_dev / ms_eval,
                 "ms_all_parameter_gradients_wall": ms, "ms_all_parameter_gradients_device_events": ms_dev, "one_pass_vs_16K_max_rel_diff": err,
                 "twin_geometry": {k: eng._outer_twin().info()[k] for k in ("warps_per_block", "cols_per_tile", "tasks_per_round", "grid_blocks", "n_tiles",
                                                                             "regs_per_thread", "copy_depth", "pool_blocks")},
                 "d_exchangeabilities": ex.grad.cpu().tolist(), "d_rates": rt.grad.cpu().tolist()})
    print(json.dumps(rows[-1]), flush=True)
    eng.close()
if len(sys.argv) > 1:
    json.dump(rows, open(sys.argv[1], "w"), indent=1)

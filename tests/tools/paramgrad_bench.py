"""Cost of the all-parameter gradient (16 K evaluations + 4x4 algebra) at full sizes.
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
for name in ("gtr_1000x10k", "hky_g4_flu"):
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

    for _ in range(2):
        run()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = 5
    for _ in range(n):
        run()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / n * 1e3
    t0 = time.perf_counter()
    for _ in range(n):
        pg.branch_outer_products(eng, bl)
    torch.cuda.synchronize()
    ms_g = (time.perf_counter() - t0) / n * 1e3
    rows.append({"workload": name, "n_taxa": w.n_tips, "n_patterns": w.n_patterns, "categories": w.n_categories,
                 "evaluations": 16 * w.n_categories, "ms_all_parameter_gradients": ms, "ms_outer_products": ms_g,
                 "d_exchangeabilities": ex.grad.cpu().tolist(), "d_rates": rt.grad.cpu().tolist()})
    print(json.dumps(rows[-1]), flush=True)
    eng.close()
if len(sys.argv) > 1:
    json.dump(rows, open(sys.argv[1], "w"), indent=1)

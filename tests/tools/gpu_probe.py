"""Developer probe (test tooling): parity vs the C oracle + kernel timing on one B200.

usage: python tests/tools/gpu_probe.py [workload[:n_tips[:n_patterns]] ...] [--reps R] [--f32]
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "gradient-benchmark_b200"))
sys.path.insert(0, ROOT)

import numpy as np
import torch

from b200phylo.engine import Engine
from b200phylo.workloads import make_workload
from oracle import c_oracle


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    reps = 20
    if "--reps" in sys.argv:
        reps = int(sys.argv[sys.argv.index("--reps") + 1])
        args = [a for a in args if a != str(reps)]
    prec = "float32" if "--f32" in sys.argv else "float64"
    check = "--nocheck" not in sys.argv
    for spec in args or ["jc69_20x1000", "jc69_1000x10k"]:
        parts = spec.split(":")
        name = parts[0]
        nt = int(parts[1]) if len(parts) > 1 and parts[1] else None
        npat = int(parts[2]) if len(parts) > 2 and parts[2] else None
        t0 = time.time()
        w = make_workload(name, n_tips=nt, n_patterns=npat)
        t_gen = time.time() - t0
        ns = getattr(w.patterns, 'n_states', 4)
        eng = Engine(w.n_tips, w.n_patterns, w.n_categories, precision=prec, n_states=ns)
        eng.set_topology(w.tree.post_triples)
        eng.set_tip_states(w.patterns.masks)
        eng.set_pattern_weights(w.patterns.weights)
        eng.set_model(w.subst_model, w.site_model)
        info = eng.info()
        bl = w.tree.branch_lengths
        logl, grad = eng.loglik_grad_host(bl)
        logl_e, _ = eng.loglik_grad_host(bl, want_grad=False)
        line = {"workload": spec, "prec": prec, "N": w.n_tips, "P": w.n_patterns, "K": w.n_categories,
                "gen_s": round(t_gen, 2), "logl": logl, "logl_evalonly": logl_e}
        if check:
            t0 = time.time()
            if ns == 4:
                ol, og = c_oracle.loglik_and_gradient(w.tree.post_triples, w.patterns.masks, w.patterns.weights,
                                                      bl, *w.eigen_args())
            else:
                from oracle import oracle as np_oracle
                ol, og = np_oracle.loglik_and_gradient(w.tree.post_triples, w.patterns.partials(), w.patterns.weights,
                                                       bl, *w.eigen_args())
            line["oracle_s"] = round(time.time() - t0, 3)
            line["oracle_logl"] = ol
            line["rel_logl"] = abs(logl - ol) / abs(ol)
            line["max_rel_grad"] = float(np.max(np.abs(grad - og) / np.maximum(np.abs(og), 1e-300)))
            line["max_abs_grad"] = float(np.max(np.abs(grad - og)))
            line["grad_inf_norm"] = float(np.max(np.abs(og)))
        # device-resident timing
        d_bl = torch.tensor(bl, dtype=torch.float64, device="cuda")
        d_out = torch.zeros(1 + eng.n_branches, dtype=torch.float64, device="cuda")
        st = torch.cuda.current_stream().cuda_stream
        eng.set_profiling(True)
        for _ in range(5):
            eng.loglik_grad_device(d_bl.data_ptr(), d_out.data_ptr(), d_out.data_ptr() + 8, st)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        kms = []
        e0.record()
        for _ in range(reps):
            eng.loglik_grad_device(d_bl.data_ptr(), d_out.data_ptr(), d_out.data_ptr() + 8, st)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        kms = eng.last_kernel_ms()
        # eval-only timing
        e0.record()
        for _ in range(reps):
            eng.loglik_grad_device(d_bl.data_ptr(), d_out.data_ptr(), None, st)
        e1.record()
        torch.cuda.synchronize()
        ms_eval = e0.elapsed_time(e1) / reps
        t0 = time.perf_counter()
        for _ in range(reps):
            eng.loglik_grad_host(bl)
        ms_host = (time.perf_counter() - t0) / reps * 1e3
        line.update(ms_grad=ms, ms_traverse_kernel=kms, ms_eval=ms_eval, ms_host_e2e=ms_host,
                    gbs=info["alg_bytes_grad"] / (kms * 1e-3) / 1e9,
                    gbs_eval=info["alg_bytes_loglik"] / (ms_eval * 1e-3) / 1e9, info=info)
        print(line, flush=True)
        eng.close()


if __name__ == "__main__":
    main()

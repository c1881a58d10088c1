"""Per-launch latency of the time-tree kernels (CUDA events over a loop of launches, on the
launching stream) next to the CPU restatement timed on this box.  Writes one JSON document.

    python tests/tools/timetree_bench.py [out.json]
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "gradient-benchmark_b200"), ROOT]

from b200phylo.timetree import TimeTreeEngine  # noqa: E402
from b200phylo.tree import simulate_tree  # noqa: E402
from oracle import timetree_oracle as tto  # noqa: E402


def gpu_us(fn, reps=300, warm=20):
    for _ in range(warm):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / reps


def cpu_us(fn, budget=0.3):
    fn()
    n, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < budget:
        fn()
        n += 1
    return (time.perf_counter() - t0) * 1e6 / n


def joint_rows():
    """One ADVI-style model evaluation (joint log density + gradient w.r.t. ratios, root height,
    rate, theta) at the headline size: the single C call, the same through autograd, and the
    function-by-function autograd chain."""
    from b200phylo.timetree import ConstantCoalescentModel, JointTimeTreeDensity, ReparameterizedTimeTreeModel
    from b200phylo.treelikelihood import StrictClockModel, TreeLikelihoodModel
    from b200phylo.workloads import make_workload
    w = make_workload("jc69_1000x10k")
    n = w.n_tips
    rng = np.random.default_rng(5)
    tips = rng.uniform(0.0, 30.0, n)
    tips[0] = 0.0
    x = rng.uniform(0.1, 0.9, n - 1)
    x[n - 2] = 45.0
    cu = lambda a: torch.tensor(np.asarray(a, dtype=np.float64), device="cuda")  # noqa: E731
    ratios, root = cu(x[: n - 2]).requires_grad_(True), cu(x[n - 2:]).requires_grad_(True)
    rate, theta = cu([0.004]).requires_grad_(True), cu([6.0]).requires_grad_(True)
    tm = ReparameterizedTimeTreeModel("tree", w.tree, tips, ratios, root)
    like = TreeLikelihoodModel("like", w.patterns, tm, w.subst_model, w.site_model, StrictClockModel("clock", rate))
    coal = ConstantCoalescentModel("coalescent", tm, theta)
    joint = JointTimeTreeDensity(like, tm, coal)
    leaves = (ratios, root, rate, theta)

    def fused():
        for t in leaves:
            t.grad = None
        joint().backward()

    def chain():
        for t in leaves:
            t.grad = None
        h = tm.heights()
        (like() + coal(h) + tm.log_jacobian(h)).backward()

    params = tm.params().detach()
    eng = tm.engine

    def c_call():
        eng.joint_density_grad(like.engine, params, rate.detach(), theta.detach(), True)

    row = {"workload": "jc69_1000x10k + strict clock + constant coalescent + log-Jacobian", "unit": "us per value+gradient",
           "joint_c_call": gpu_us(c_call, reps=200), "joint_autograd": gpu_us(fused, reps=200),
           "function_by_function_autograd": gpu_us(chain, reps=200)}
    like.close()
    tm.close()
    return row


def main():
    out = []
    for n, shape in [(1000, "coalescent"), (2000, "coalescent"), (5000, "coalescent"), (2000, "caterpillar"),
                     (4096, "balanced")]:
        tree = simulate_tree(n, shape=shape, seed=11)
        rng = np.random.default_rng(3)
        tips = rng.uniform(0.0, 40.0, n)
        tips[0] = 0.0
        x = rng.uniform(0.05, 0.95, n - 1)
        x[n - 2] = 50.0
        tr = tree.post_triples
        eng = TimeTreeEngine(tr, tips)
        cu = lambda a: torch.tensor(np.asarray(a, dtype=np.float64), device="cuda")  # noqa: E731
        px, g, gb, th, rate = cu(x), cu(rng.normal(size=n - 1)), cu(rng.normal(size=2 * n - 2)), cu([4.0]), cu([0.001])
        h = eng.heights(px)
        hn = h.cpu().numpy()
        lib, hd = eng._lib, eng._h
        m = n - 1
        o1, o2, o3 = torch.empty(m, dtype=torch.float64, device="cuda"), torch.empty(2 * n, dtype=torch.float64, device="cuda"), \
            torch.empty(2, dtype=torch.float64, device="cuda")
        s = torch.cuda.current_stream().cuda_stream
        row = {"n_tips": n, "shape": shape, **eng.info(), "gpu_us": {}, "cpu_us": {}}
        row["gpu_us"]["ratio_transform"] = gpu_us(lambda: lib.b2p_timetree_heights(hd, px.data_ptr(), o1.data_ptr(), s))
        row["gpu_us"]["ratio_transform_vjp"] = gpu_us(
            lambda: lib.b2p_timetree_heights_vjp(hd, px.data_ptr(), h.data_ptr(), g.data_ptr(), o1.data_ptr(), s))
        row["gpu_us"]["ratio_transform_jacobian+grad"] = gpu_us(
            lambda: lib.b2p_timetree_log_jacobian(hd, h.data_ptr(), o3.data_ptr(), o1.data_ptr(), s))
        row["gpu_us"]["branch_lengths"] = gpu_us(
            lambda: lib.b2p_timetree_branch_lengths(hd, h.data_ptr(), rate.data_ptr(), o2.data_ptr(), s))
        row["gpu_us"]["branch_lengths_vjp"] = gpu_us(
            lambda: lib.b2p_timetree_branch_lengths_vjp(hd, h.data_ptr(), rate.data_ptr(), gb.data_ptr(), o1.data_ptr(),
                                                        o3.data_ptr(), s))
        row["gpu_us"]["coalescent+grad"] = gpu_us(
            lambda: lib.b2p_coalescent_constant(hd, h.data_ptr(), th.data_ptr(), o3.data_ptr(), o1.data_ptr(),
                                                o3.data_ptr() + 8, s))
        gn, gbn = g.cpu().numpy(), gb.cpu().numpy()
        row["cpu_us"]["ratio_transform"] = cpu_us(lambda: tto.heights_from_ratios(n, tr, tips, x))
        row["cpu_us"]["ratio_transform_vjp"] = cpu_us(lambda: tto.heights_vjp(n, tr, tips, x, gn))
        row["cpu_us"]["ratio_transform_jacobian+grad"] = cpu_us(
            lambda: (tto.log_jacobian(n, tr, tips, hn), tto.log_jacobian_grad(n, tr, tips, hn)))
        row["cpu_us"]["branch_lengths"] = cpu_us(lambda: tto.branch_lengths(n, tr, tips, hn, 0.001))
        row["cpu_us"]["branch_lengths_vjp"] = cpu_us(lambda: tto.branch_lengths_vjp(n, tr, tips, hn, 0.001, gbn))
        row["cpu_us"]["coalescent+grad"] = cpu_us(lambda: tto.constant_coalescent(n, tips, hn, 4.0))
        row["cpu_kind"] = "port (numpy/Python restatement, 1 core)"
        out.append(row)
        eng.close()
        print(json.dumps(row), flush=True)
    out.append(joint_rows())
    print(json.dumps(out[-1]), flush=True)
    if len(sys.argv) > 1:
        with open(sys.argv[1], "w") as fh:
            json.dump(out, fh, indent=1)


if __name__ == "__main__":
    main()

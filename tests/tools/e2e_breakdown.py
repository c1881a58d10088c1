"""Where the end-to-end microseconds go (test tooling): the same evaluation timed at four depths.
usage: python tests/tools/e2e_breakdown.py [workload] [reps]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "gradient-benchmark_b200"), ROOT]
import numpy as np
import torch

from b200phylo.treelikelihood import TreeLikelihoodModel, UnrootedTreeModel
from b200phylo.workloads import make_workload

name = sys.argv[1] if len(sys.argv) > 1 else "jc69_1000x10k"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 300
w = make_workload(name)
dev = torch.device("cuda", 0)
bl_dev = torch.tensor(w.tree.branch_lengths, dtype=torch.float64, device=dev)
model = TreeLikelihoodModel("like", w.patterns, UnrootedTreeModel("tree", w.tree, bl_dev), w.subst_model, w.site_model, device=0, rank=0, world_size=1)
eng = model.engine
n_br = eng.n_branches
out = torch.zeros(1 + n_br, dtype=torch.float64, device=dev)
st = torch.cuda.current_stream().cuda_stream


def timed(fn, n=reps):
    for _ in range(20):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e6


def dev_queued():
    eng.loglik_grad_device(bl_dev.data_ptr(), out.data_ptr(), out.data_ptr() + 8, st)


def dev_sync():
    eng.loglik_grad_device(bl_dev.data_ptr(), out.data_ptr(), out.data_ptr() + 8, st)
    torch.cuda.synchronize()


bl_host = torch.tensor(w.tree.branch_lengths, dtype=torch.float64).pin_memory()
res = torch.empty(1 + n_br, dtype=torch.float64)


def host_c():
    eng.loglik_grad_host_ptr(bl_host.data_ptr(), res.data_ptr(), res.data_ptr() + 8)


blh = bl_host.clone().requires_grad_(True)
model.tree_model = UnrootedTreeModel("tree", w.tree, blh)


def host_autograd():
    blh.grad = None
    logl = model()
    logl.backward()
    return float(logl.detach())


def host_forward_only():
    with torch.no_grad():
        return float(model())


print(f"{name}: us per evaluation")
print(f"  device pointers, launches queued back to back      {timed(dev_queued):8.1f}")
print(f"  device pointers, synchronise after every call      {timed(dev_sync):8.1f}")
print(f"  b2p_loglik_grad_host (C call, host buffers)        {timed(host_c):8.1f}")
print(f"  model() + backward() on a host tensor (autograd)   {timed(host_autograd):8.1f}")
print(f"  model() under no_grad (log-likelihood only)        {timed(host_forward_only):8.1f}")
os.environ["B2P_NO_GRAPH"] = "1"
model.close()

"""Multi-GPU check of the fused reduction + all-reduce (run under torchrun, one rank per GPU):
every rank evaluates its own pattern shard; the engine's results must equal the NCCL sum of
the per-rank results and the CPU oracle on the whole alignment, bit-identical across ranks.
    python -m torch.distributed.run --nproc-per-node N tests/tools/peer_check.py
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "gradient-benchmark_b200"), ROOT, os.path.join(ROOT, "tests")]
from b200phylo.parallel import attach_peers, shard_bounds  # noqa: E402
from b200phylo.workloads import make_workload  # noqa: E402
from helpers import grad_err, make_engine, oracle_c  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device(f"cuda:{local}")
dist.init_process_group("nccl", device_id=dev)
ok = True
for name, nt, npat in (("jc69_1000x10k", 200, 3001), ("hky_g4_flu", 150, 777), ("codon_500x2000", 40, 700)):
    w = make_workload(name, n_tips=nt, n_patterns=npat)
    lo, hi = shard_bounds(w.n_patterns, world, rank)
    sub = make_workload(name, n_tips=nt, n_patterns=npat)
    sub.patterns = w.patterns.slice(lo, hi)
    bl = torch.tensor(w.tree.branch_lengths, device=dev)
    n_br = bl.numel()
    # NCCL reference: plain engine + all_reduce
    e0 = make_engine(sub, device=local)
    ref = torch.zeros(1 + n_br, dtype=torch.float64, device=dev)
    e0.loglik_grad_device(bl.data_ptr(), ref.data_ptr(), ref.data_ptr() + 8, torch.cuda.current_stream().cuda_stream)
    dist.all_reduce(ref)
    e0.close()
    # fused
    e1 = make_engine(sub, device=local)
    assert attach_peers(e1), "peer attach failed"
    out = torch.zeros(1 + n_br, dtype=torch.float64, device=dev)
    for it in range(5):  # several epochs: both slot sets, flag reuse
        out.zero_()
        e1.loglik_grad_device(bl.data_ptr(), out.data_ptr(), out.data_ptr() + 8, torch.cuda.current_stream().cuda_stream)
    # a late rank: the others' kernels wait for its flags (up to ~30 s) instead of summing stale slots
    if rank == world - 1:
        import time
        time.sleep(1.5)
    late = torch.zeros(1 + n_br, dtype=torch.float64, device=dev)
    e1.loglik_grad_device(bl.data_ptr(), late.data_ptr(), late.data_ptr() + 8, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ok = ok and torch.equal(late, out)
    ev = torch.zeros(1, dtype=torch.float64, device=dev)
    e1.loglik_grad_device(bl.data_ptr(), ev.data_ptr(), None, torch.cuda.current_stream().cuda_stream)  # evaluation-only epoch
    torch.cuda.synchronize()
    attached, err = e1.peer_status()
    rel = float(((out - ref).abs() / ref.abs().clamp_min(1e-300)).max())
    # bit-identical across ranks
    gathered = [torch.empty_like(out) for _ in range(world)]
    dist.all_gather(gathered, out)
    same = all(torch.equal(g, gathered[0]) for g in gathered)
    line = f"rank {rank} {name}: attached={attached} err={err} rel_vs_nccl={rel:.2e} identical_across_ranks={same} evalonly_rel={abs(float(ev) - float(ref[0])) / abs(float(ref[0])):.2e}"
    if rank == 0:
        if getattr(w.patterns, "n_states", 4) == 4:
            ol, og = oracle_c(w)
        else:  # dense-state path: the run-time-S C port
            from oracle import c_oracle
            ol, og = c_oracle.dense_loglik_and_gradient(w.tree.post_triples, w.patterns.masks, w.patterns.weights, w.tree.branch_lengths,
                                                        *w.eigen_args())
        line += f" logl_rel_vs_oracle={abs(float(out[0]) - ol) / abs(ol):.2e} grad_vs_oracle={grad_err(out[1:].cpu().numpy(), og):.2e}"
        ok = ok and abs(float(out[0]) - ol) / abs(ol) < 1e-10 and grad_err(out[1:].cpu().numpy(), og) < 1e-8
    print(line, flush=True)
    ok = ok and attached and err == 0 and rel < 1e-12 and same
    e1.close()
dist.barrier()
dist.destroy_process_group()
print(f"rank {rank} {'OK' if ok else 'FAILED'}")
sys.exit(0 if ok else 1)

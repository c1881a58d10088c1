#!/usr/bin/env python3
"""bench.py — gradient evaluations/s of the tree-likelihood hot path on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

A *step* is one evaluation of log-likelihood + all 2N-2 branch-length derivatives over one
rank's site patterns (K1 transition matrices -> walk kernel (both passes) -> reduce kernel), plus,
for N>1, the single all-reduce of 2N-1 doubles that finishes a sharded evaluation.  The
default workload is BASELINE.json configs[1]: JC69, 1000 taxa x 10k patterns, fp64.  With
N>1 every rank holds its own 10k-pattern shard of one N*10k-pattern alignment on the same
tree (weak scaling; value counts one unit per rank per step).

Prints ONE JSON line (rank 0).  `value` is device-resident throughput (branch lengths
already in HBM, CUDA events on the launch stream); `e2e` goes through the public
`TreeLikelihoodModel` with HOST tensors (pinned H2D of branch lengths, D2H of logL + gradient
inside the timed region); `roofline` is the traversal kernel's algorithmic bytes (SURVEY.md
§8(d): (5N-9)U + 2NP) over its mean CUDA-event duration inside the timed region, against
MEASURED_PEAKS.json; `cpu_baseline` is the C oracle (a port — the reference ships no
likelihood code) timed on this box's host cores.

`--impl reference` times that CPU port alone, all host threads, same workload and metric.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (os.path.join(ROOT, "gradient-benchmark_b200"), ROOT):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "gradient_evals_per_s"
UNIT_FMT = "logL+branch-gradient evaluations/s ({n} taxa x {p} patterns per GPU)"
FALLBACK_HBM_GBS = 6650.0  # /opt/skills/guides/B200_PROFILING.md, used only if MEASURED_PEAKS.json is absent


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="jc69_1000x10k")
    ap.add_argument("--precision", default="float64", choices=["float64", "float32"])
    ap.add_argument("--patterns", type=int, default=0,
                    help="site patterns per GPU (0: the workload's own size); e.g. --workload gtr_g4_5000x100k "
                         "--patterns 12500 --gpus 8 is BASELINE configs[3]")
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="budget of the cpu_baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def ncu_traffic(workload, precision):
    """dram__bytes_read.sum + dram__bytes_write.sum of the traversal kernel, per launch, from the
    committed `ncu --set full` capture of this workload (profiles/ncu_traffic.json); None if absent."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as fh:
            return float(json.load(fh)[f"{workload}:{precision}"]["dram_bytes"])
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [l for t, l in self.lines if (t0 is None or t >= t0 - 0.05) and (t1 is None or t <= t1 + 0.15)]
        rows = rows or [l for _, l in self.lines]
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in rows:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower() == "active":
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def build_workload(name, precision, rank, n_patterns=0):
    from b200phylo.workloads import SEQ_SEED, make_workload
    return make_workload(name, precision=precision, seq_seed=SEQ_SEED + rank, n_patterns=n_patterns or None)


def cpu_port_rate(w, seconds, n_threads=0, min_evals=3):
    """Timed C oracle (CPU port of the physher/BEAGLE-style analytic two-pass path)."""
    fn, cores = cpu_port(w, n_threads)
    fn()  # warm-up
    n, t0 = 0, time.perf_counter()
    while n < min_evals or time.perf_counter() - t0 < seconds:
        logl, _ = fn()
        n += 1
    dt = time.perf_counter() - t0
    return n / dt, cores, n, dt, logl


def cpu_port(w, n_threads=0):
    """The CPU implementation of the path for this workload and the cores it uses: the C/OpenMP
    port for 4-state models, the numpy restatement (BLAS threads) for dense-state models."""
    if getattr(w.patterns, "n_states", 4) == 4:
        from oracle import c_oracle
        args = (w.tree.post_triples, w.patterns.masks, w.patterns.weights, w.tree.branch_lengths, *w.eigen_args())
        return (lambda: c_oracle.loglik_and_gradient(*args, n_threads=n_threads)), (n_threads or c_oracle.max_threads())
    from oracle import oracle
    args = (w.tree.post_triples, w.patterns.partials(), w.patterns.weights, w.tree.branch_lengths, *w.eigen_args())
    return (lambda: oracle.loglik_and_gradient(*args)), 1  # numpy einsum contractions: one core


def run_reference(a):
    """--impl reference: the CPU implementation of the path, all host threads.  The
    reference repo itself holds no likelihood code and its contenders (physher, BEAGLE via
    bito, torchtree) are not installable here, so this arm times the C port (oracle/)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    w = build_workload(a.workload, a.precision, 0, a.patterns)
    fn, cores = cpu_port(w)
    for _ in range(max(a.warmup, 1)):
        fn()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        fn()
    dt = time.perf_counter() - t0
    value = a.steps / dt
    unit = UNIT_FMT.format(n=w.n_tips, p=w.n_patterns)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": unit, "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": dt / a.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "gpu_launches": 0,
        "config": {"workload": a.workload, "n_taxa": w.n_tips, "patterns_per_gpu": w.n_patterns,
                   "categories": w.n_categories, "note": "CPU port on host cores; one shard's worth of work per step"},
        "cpu_baseline": {"value": value, "unit": unit, "cores": cores, "kind": "port",
                         "sample": f"{a.steps} full evaluations of the workload (C oracle, OpenMP over pattern blocks)"},
        "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)
    return 0


def run_b200(a):
    import torch
    import torch.distributed as dist

    from b200phylo.treelikelihood import TreeLikelihoodModel, UnrootedTreeModel

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the engine has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    stdout_fd = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL writes its version banner to file descriptor 1 on first use: route everything but the
        # JSON line to stderr, so stdout carries exactly one line
        sys.stdout.flush()
        stdout_fd = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)
    if world != a.gpus and rank == 0:
        print(f"bench.py: --gpus {a.gpus} but WORLD_SIZE={world}; using {world}", file=sys.stderr)

    w = build_workload(a.workload, a.precision, rank, a.patterns)
    n_br = 2 * w.n_tips - 2
    # world_size=1 on the model: each rank already holds exactly its own shard
    bl_dev = torch.tensor(w.tree.branch_lengths, dtype=torch.float64, device=dev)
    model = TreeLikelihoodModel("like", w.patterns, UnrootedTreeModel("tree", w.tree, bl_dev), w.subst_model,
                                w.site_model, precision=a.precision, device=local, rank=0, world_size=1)
    eng = model.engine
    info = eng.info()
    out = torch.zeros(1 + n_br, dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream(dev)
    # N > 1: the engine's reduction kernel is the all-reduce (NVLink peer stores, b2p_peer_*);
    # NCCL only if the peer mapping cannot be set up or B2P_NCCL_ALLREDUCE=1 asks for it
    fused = False
    if world > 1 and info["n_states"] == 4:
        from b200phylo.parallel import attach_peers
        fused = attach_peers(eng)
    model.fused_allreduce = fused

    def step_device():
        eng.loglik_grad_device(bl_dev.data_ptr(), out.data_ptr(), out.data_ptr() + 8, stream.cuda_stream)
        if world > 1 and not fused:
            dist.all_reduce(out)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput (value) + live kernel timing for the roofline
    for _ in range(a.warmup):
        step_device()
    eng.set_profiling(True)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_wall0 = time.perf_counter()
    e0.record(stream)
    for _ in range(a.steps):
        step_device()
    e1.record(stream)
    barrier()
    t_wall1 = time.perf_counter()
    ms_total = e0.elapsed_time(e1)
    k_ms, k_n = eng.kernel_ms_stats()
    eng.set_profiling(False)
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    result_dev = out.cpu().numpy().copy()

    # ---- end to end through the public API with host tensors
    bl_host = torch.tensor(w.tree.branch_lengths, dtype=torch.float64).pin_memory().requires_grad_(True)
    model.tree_model = UnrootedTreeModel("tree", w.tree, bl_host)
    model.world_size = world  # host path: all-reduce inside the autograd function

    def step_host():
        bl_host.grad = None
        logl = model()
        logl.backward()
        return float(logl.detach()), bl_host.grad

    for _ in range(max(3, a.warmup // 2)):
        step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        logl_h, grad_h = step_host()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt_e2e = float(t.item())
    import numpy as np
    assert abs(logl_h - result_dev[0]) <= 1e-9 * abs(result_dev[0]), (logl_h, result_dev[0])
    assert np.allclose(grad_h.numpy(), result_dev[1:], rtol=1e-9, atol=1e-9)

    if rank == 0:
        peak, peak_src = hbm_peak()
        alg = info["alg_bytes_grad"]
        k_avg_ms = k_ms / max(k_n, 1)
        achieved = alg / (k_avg_ms * 1e-3) / 1e9
        dense = info["n_states"] != 4
        unit = UNIT_FMT.format(n=w.n_tips, p=w.n_patterns)
        U = w.n_patterns * w.n_categories * 4 * (8 if a.precision == "float64" else 4)
        line = {
            "metric": METRIC, "value": world * a.steps / (ms_total * 1e-3), "unit": unit, "n_gpus": world,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_total / a.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64" if a.precision == "float64" else "f32", "data": "synthetic",
            "config": {"workload": a.workload, "n_taxa": w.n_tips, "patterns_per_gpu": w.n_patterns,
                       "total_patterns": w.n_patterns * world, "categories": w.n_categories,
                       "model": type(w.subst_model).__name__, "tree": "coalescent, seed 20261104",
                       "l2": f"no flush: per-step working set {(2 * (w.n_tips - 1) * U * info['n_states'] // 4) / 1e6:.0f} MB of partials > 126 MB L2",
                       "collective": ("fused into the reduction kernel: NVLink peer stores + flags, sum in rank order (b2p_peer_*)" if fused
                                      else "1 NCCL all-reduce of 2N-1 doubles per step") if world > 1 else "none",
                       "engine": {k: info[k] for k in ("warps_per_block", "cols_per_tile", "n_tiles", "grid_blocks",
                                                        "post_rounds", "pre_rounds", "tasks_per_round", "tree_depth",
                                                        "regs_per_thread", "smem_bytes", "pool_blocks", "copy_depth",
                                                        "resident_reads", "block_loads", "block_stores")}},
            "e2e": {"value": world * a.steps / dt_e2e, "unit": unit, "ms_per_step": dt_e2e / a.steps * 1e3,
                    "h2d_bytes_per_step": n_br * 8, "d2h_bytes_per_step": (n_br + 1) * 8,
                    "api": "TreeLikelihoodModel()(host tensor).backward() -> b2p_loglik_grad_host"},
            "gpu_launches": 3 * a.steps,
            "roofline": {"bound": "hbm", "kernel": "walk_kernel", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": ncu_traffic(a.workload, a.precision),
                         "alg_bytes_per_launch": alg, "kernel_ms": k_avg_ms, "launches_timed": k_n,
                         "share_of_step": k_ms / (ms_total * k_n / a.steps) if k_n else None, "peak_source": peak_src},
            "clocks": clocks,
            "logl": float(result_dev[0]),
        }
        if dense:
            # compute-bound; algorithmic flops per evaluation
            # (N-1) P K (4S^2+S) + (2N-2) P K (6S^2+4S), against an fp64 matmul measured in this run
            S, N_, P_, K_ = info["n_states"], w.n_tips, w.n_patterns, w.n_categories
            flops = (N_ - 1) * P_ * K_ * (4 * S * S + S) + (2 * N_ - 2) * P_ * K_ * (6 * S * S + 4 * S)
            a_ = torch.randn(4096, 4096, dtype=torch.float64, device=dev)
            best = 1e9
            for _ in range(5):
                m0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                m0.record(); torch.matmul(a_, a_); m1.record(); torch.cuda.synchronize()
                best = min(best, m0.elapsed_time(m1))
            peak_tf = 2 * 4096 ** 3 / (best * 1e-3) / 1e12
            ach_tf = flops / (k_avg_ms * 1e-3) / 1e12
            # what the message formulation actually issues on the tensor cores (64-padded tiles): one
            # product per internal node, two per internal non-root node, two per branch in K1
            tiles = info["n_tiles"]
            issued = 2 * 64 ** 3 * (tiles * ((N_ - 1) + 2 * (N_ - 2)) + 2 * (2 * N_ - 2))
            line["dtype"] = "f64"
            # transition, one post-order launch per level, one pre-order launch per depth, tips, reduce
            line["gpu_launches"] = (1 + info["post_rounds"] + info["pre_rounds"] + 2) * a.steps
            line["roofline"] = {"bound": "tensor", "kernel": "dense_post_kernel + dense_pre_kernel (one launch per tree level) + dense_tip_kernel",
                                "achieved": ach_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": ach_tf / peak_tf, "traffic": None,
                                "alg_flops_per_eval": flops, "issued_flops_per_eval": issued,
                                "issued_tflops": issued / (k_avg_ms * 1e-3) / 1e12, "issued_frac": issued / (k_avg_ms * 1e-3) / 1e12 / peak_tf,
                                "kernel_ms": k_avg_ms, "launches_timed": k_n,
                                "peak_source": "fp64 torch.matmul 4096^3 (cuBLAS DGEMM), best of 5, measured in this run"}
        if not a.no_cpu_baseline:
            rate, cores, n, secs, logl_cpu = cpu_port_rate(w, a.cpu_seconds)
            rate1 = rate if dense else cpu_port_rate(w, min(a.cpu_seconds, 5.0), n_threads=1, min_evals=2)[0]
            line["cpu_baseline"] = {"value": rate, "unit": unit, "cores": cores, "kind": "port",
                                    "sample": f"{n} full evaluations of the same workload in {secs:.1f} s "
                                              + ("(numpy restatement, one core)" if dense else "(C oracle, OpenMP)"),
                                    "single_core_value": rate1,
                                    "logl_rel_diff_vs_gpu": abs(logl_cpu - result_dev[0]) / abs(logl_cpu) if world == 1 else None}
        if stdout_fd is not None:
            sys.stdout.flush()
            os.dup2(stdout_fd, 1)
        print(json.dumps(line), flush=True)
    model.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    a = parse_args()
    return run_reference(a) if a.impl == "reference" else run_b200(a)


if __name__ == "__main__":
    sys.exit(main())

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
    ap.add_argument("--no-secondary", action="store_true",
                    help="skip the short runs of the other BASELINE.json configs that ride in the line's `secondary` array")
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


def config_of(name, w, world):
    """The workload as both arms describe it (same keys, same values: the driver compares them)."""
    return {"workload": name, "n_taxa": w.n_tips, "patterns_per_gpu": w.n_patterns,
            "total_patterns": w.n_patterns * world, "categories": w.n_categories,
            "states": int(getattr(w.patterns, "n_states", 4)), "model": type(w.subst_model).__name__,
            "tree": "coalescent, seed 20261104"}


def cpu_port(w, n_threads=0, want_grad=True):
    """The CPU implementation of the path for this workload and the cores it uses: the C/OpenMP
    restatement of the analytic two-pass algorithm (oracle/phylo_oracle.c for 4 states,
    oracle/phylo_oracle_dense.c for any state count)."""
    from oracle import c_oracle
    args = (w.tree.post_triples, w.patterns.masks, w.patterns.weights, w.tree.branch_lengths, *w.eigen_args())
    fn = c_oracle.loglik_and_gradient if getattr(w.patterns, "n_states", 4) == 4 else c_oracle.dense_loglik_and_gradient
    return (lambda: fn(*args, n_threads=n_threads, want_grad=want_grad)), (n_threads or c_oracle.max_threads())


def cpu_port_rate(w, seconds, n_threads=0, min_evals=3):
    """Timed C oracle (CPU port of the physher/BEAGLE-style analytic two-pass path)."""
    fn, cores = cpu_port(w, n_threads)
    fn()  # warm-up
    n, t0 = 0, time.perf_counter()
    while n < min_evals or time.perf_counter() - t0 < seconds:
        logl, grad = fn()
        n += 1
    dt = time.perf_counter() - t0
    return n / dt, cores, n, dt, logl, grad


def host_cores():
    """Every core this process may use, whatever OMP_NUM_THREADS says (torchrun sets it to 1)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(a):
    """--impl reference: the CPU implementation of the path, all host threads.  The reference repo
    itself holds no likelihood code and its contenders (physher, BEAGLE via bito, torchtree) are
    not installable here, so this arm times the C port (oracle/).  Under torchrun rank 0 alone
    runs: it evaluates the whole N-shard alignment (the N shards the other arm's N GPUs hold) on
    all host cores, and `value` counts shard evaluations per second like the other arm's."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return 0
    cores = host_cores()
    shards = [build_workload(a.workload, a.precision, r, a.patterns) for r in range(world)]
    fns = [cpu_port(w, n_threads=cores)[0] for w in shards]
    w = shards[0]

    def step():
        return [fn() for fn in fns]

    for _ in range(max(min(a.warmup, 3), 1)):
        step()
    # bounded: full evaluations until --steps are done or ~60 s have passed (at least 2)
    n, t0 = 0, time.perf_counter()
    while n < a.steps and (n < 2 or time.perf_counter() - t0 < 60.0):
        step()
        n += 1
    dt = time.perf_counter() - t0
    value = world * n / dt
    unit = UNIT_FMT.format(n=w.n_tips, p=w.n_patterns)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": unit, "n_gpus": world,
        "steps": n, "warmup": a.warmup, "ms_per_step": dt / n * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "gpu_launches": 0,
        "config": config_of(a.workload, w, world),
        "cpu_baseline": {"value": value, "unit": unit, "cores": cores, "kind": "port",
                         "sample": f"{n} evaluations of all {world} shard(s) of the workload, one after the other, each on "
                                   f"{cores} threads (C oracle, OpenMP over pattern blocks)"},
        "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)
    return 0

class Dist:
    """Process-group facts of this run (one process per GPU under torchrun)."""

    def __init__(self):
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))


def measure(name, precision, patterns, steps, warmup, ctx, want_e2e=True, cpu_seconds=0.0, cpu_check=True):
    """One workload on this rank's GPU: device-resident throughput (CUDA events on the launch stream,
    max over ranks), live kernel timing for the roofline, clocks sampled during the timed region,
    optionally the end-to-end rate through the public model object with host tensors, and a check
    of the result against the CPU port.  Returns the fields of a JSON line (rank 0) or None."""
    import numpy as np
    import torch
    import torch.distributed as dist

    from b200phylo.treelikelihood import TreeLikelihoodModel, UnrootedTreeModel

    rank, world, local = ctx.rank, ctx.world, ctx.local
    dev = torch.device("cuda", local)
    w = build_workload(name, precision, rank, patterns)
    n_br = 2 * w.n_tips - 2
    # world_size=1 on the model: each rank already holds exactly its own shard
    bl_dev = torch.tensor(w.tree.branch_lengths, dtype=torch.float64, device=dev)
    model = TreeLikelihoodModel("like", w.patterns, UnrootedTreeModel("tree", w.tree, bl_dev), w.subst_model,
                                w.site_model, precision=precision, device=local, rank=0, world_size=1)
    eng = model.engine
    info = eng.info()
    dense = info["n_states"] != 4
    out = torch.zeros(1 + n_br, dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream(dev)
    # N > 1: the engine's reduction kernel is the all-reduce (NVLink peer stores, b2p_peer_*);
    # NCCL only if the peer mapping cannot be set up or B2P_NCCL_ALLREDUCE=1 asks for it
    fused = False
    if world > 1 and not dense:
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

    for _ in range(warmup):
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
    for _ in range(steps):
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

    e2e = None
    if want_e2e:
        # end to end through the public API with host tensors
        bl_host = torch.tensor(w.tree.branch_lengths, dtype=torch.float64).pin_memory().requires_grad_(True)
        model.tree_model = UnrootedTreeModel("tree", w.tree, bl_host)
        model.world_size = world  # host path: all-reduce inside the autograd function

        def step_host():
            bl_host.grad = None
            logl = model()
            logl.backward()
            return float(logl.detach()), bl_host.grad

        for _ in range(max(3, warmup // 2)):
            step_host()
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            logl_h, grad_h = step_host()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt_e2e = float(t.item())
        assert abs(logl_h - result_dev[0]) <= 1e-9 * abs(result_dev[0]), (logl_h, result_dev[0])
        assert np.allclose(grad_h.numpy(), result_dev[1:], rtol=1e-9, atol=1e-9)
        e2e = dt_e2e

    # ---- the CPU port: every rank evaluates ITS shard (cores / world threads each); summed over
    # ranks this is the whole alignment's logL and gradient, the same quantity the GPUs' all-reduce
    # produced — a check of the sharded sum as well as of the kernels
    check, cpu = None, None
    if cpu_check or cpu_seconds > 0:
        cores = host_cores()
        thr = max(1, cores // world)
        if cpu_seconds > 0 and world == 1:
            rate, used, n, secs, logl_cpu, grad_cpu = cpu_port_rate(w, cpu_seconds, n_threads=thr)
            rate1 = cpu_port_rate(w, min(cpu_seconds, 4.0), n_threads=1, min_evals=1)[0]
            cpu = {"value": rate, "cores": used, "kind": "port", "single_core_value": rate1,
                   "sample": f"{n} full evaluations of the same workload in {secs:.1f} s (C oracle, OpenMP over pattern blocks)"}
        else:
            logl_cpu, grad_cpu = cpu_port(w, n_threads=thr)[0]()
        ref = torch.tensor(np.concatenate([[logl_cpu], grad_cpu]), dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ref)
        ref = ref.cpu().numpy()
        scale = np.maximum(np.abs(ref[1:]), np.abs(ref[1:]).max() * 1e-4)
        check = {"logl_rel_diff_vs_cpu": float(abs(result_dev[0] - ref[0]) / abs(ref[0])),
                 "grad_rel_diff_vs_cpu": float(np.max(np.abs(result_dev[1:] - ref[1:]) / scale)),
                 "cpu": f"C oracle on every rank's shard ({thr} threads each), summed over {world} rank(s)"}

    fields = None
    if rank == 0:
        peak, peak_src = hbm_peak()
        alg = info["alg_bytes_grad"]
        k_avg_ms = k_ms / max(k_n, 1)
        achieved = alg / (k_avg_ms * 1e-3) / 1e9
        unit = UNIT_FMT.format(n=w.n_tips, p=w.n_patterns)
        U = w.n_patterns * w.n_categories * 4 * (8 if precision == "float64" else 4)
        traffic = ncu_traffic(name, precision)
        fields = {
            "value": world * steps / (ms_total * 1e-3), "unit": unit, "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": ms_total / steps, "dtype": "f64" if precision == "float64" else "f32",
            "config": config_of(name, w, world),
            "timing": {"l2": f"no flush: per-step working set {(2 * (w.n_tips - 1) * U * info['n_states'] // 4) / 1e6:.0f} MB "
                             "of partials > 126 MB L2" if (2 * (w.n_tips - 1) * U * info['n_states'] // 4) > 126e6 else
                             "no flush, and the working set fits the 126 MB L2 (a latency-bound shape; the roofline fraction is not meaningful)",
                       "collective": ("fused into the reduction kernel: NVLink peer stores + flags, sum in rank order (b2p_peer_*)" if fused
                                      else "1 NCCL all-reduce of 2N-1 doubles per step") if world > 1 else "none"},
            "engine": {k: info[k] for k in ("warps_per_block", "cols_per_tile", "n_tiles", "grid_blocks", "post_rounds", "pre_rounds",
                                            "tasks_per_round", "tree_depth", "regs_per_thread", "smem_bytes", "pool_blocks",
                                            "copy_depth", "resident_reads", "block_loads", "block_stores")},
            "gpu_launches": 3 * steps,
            "roofline": {"bound": "hbm", "kernel": "walk_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic,
                         "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full capture of this workload "
                                           "(profiles/ncu_traffic.json; the builder's box, not this run)" if traffic else None,
                         "alg_bytes_per_launch": alg, "kernel_ms": k_avg_ms, "launches_timed": k_n,
                         "share_of_step": k_ms / (ms_total * k_n / steps) if k_n else None, "peak_source": peak_src},
            "clocks": clocks, "logl": float(result_dev[0]),
        }
        if e2e is not None:
            fields["e2e"] = {"value": world * steps / e2e, "unit": unit, "ms_per_step": e2e / steps * 1e3,
                             "h2d_bytes_per_step": n_br * 8, "d2h_bytes_per_step": (n_br + 1) * 8,
                             "api": "TreeLikelihoodModel()(host tensor).backward() -> b2p_loglik_grad_host"}
        if check:
            fields["check"] = check
        if cpu:
            cpu["unit"] = unit
            fields["cpu_baseline"] = cpu
        if dense:
            fields.update(dense_roofline(info, w, k_avg_ms, k_n, steps, dev))
    model.close()
    return fields

def dense_roofline(info, w, k_avg_ms, k_n, steps, dev):
    """SURVEY.md §8(d) config 5 is compute-bound: algorithmic flops per evaluation
    (N-1) P K (4S^2+S) + (2N-2) P K (6S^2+4S) against an fp64 matmul measured in this run.
    `frac` is that SURVEY-formula figure; `issued_frac` — the flops the message formulation
    actually issues on the tensor cores (64-padded tiles) over the same peak — is the utilisation."""
    import torch
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
    # one product per internal node, two per internal non-root node, two per branch in K1
    tiles = info["n_tiles"]
    issued = 2 * 64 ** 3 * (tiles * ((N_ - 1) + 2 * (N_ - 2)) + 2 * (2 * N_ - 2))
    return {
        # transition, one post-order launch, one pre-order launch (dependencies by flag), tips, reduce
        "gpu_launches": (1 + info["post_rounds"] + info["pre_rounds"] + 2) * steps,
        "roofline": {"bound": "tensor", "kernel": "dense_transition_kernel + dense_post_kernel + dense_pre_kernel (one launch per pass, flag-chained) + dense_tip_kernel",
                     "achieved": ach_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": ach_tf / peak_tf,
                     "frac_is": "SURVEY.md §8(d) algorithmic flops / time / peak (the formulation issues fewer: see issued_frac)",
                     "traffic": None, "alg_flops_per_eval": flops, "issued_flops_per_eval": issued,
                     "issued_tflops": issued / (k_avg_ms * 1e-3) / 1e12, "issued_frac": issued / (k_avg_ms * 1e-3) / 1e12 / peak_tf,
                     "kernel_ms": k_avg_ms, "launches_timed": k_n,
                     "peak_source": "fp64 torch.matmul 4096^3 (cuBLAS DGEMM), best of 5, measured in this run"}}


def measure_parameter_gradients(dev, steps=30):
    """SURVEY.md §8(f) row 3 as an extra entry of the line (N = 1): every substitution- / site-model
    gradient of a GTR model at the headline shape from ONE outer-products walk + b2p_model_gradients,
    next to one branch-gradient evaluation of the same model (device events, resident inputs)."""
    import torch
    from b200phylo.engine import Engine
    from b200phylo.sitemodel import ConstantSiteModel
    from b200phylo.substitution import GTR
    from b200phylo.workloads import make_workload
    w = make_workload("jc69_1000x10k")
    eng = Engine(w.n_tips, w.n_patterns, 1, device=dev.index or 0)
    try:
        eng.set_topology(w.tree.post_triples)
        eng.set_tip_states(w.patterns.masks)
        eng.set_pattern_weights(w.patterns.weights)
        eng.set_model(GTR((1.3, 2.9, 0.7, 1.1, 3.4, 1.0), (0.31, 0.19, 0.22, 0.28)), ConstantSiteModel())
        bl = torch.tensor(w.tree.branch_lengths, dtype=torch.float64, device=dev)
        out = torch.empty(1 + eng.n_branches, dtype=torch.float64, device=dev)
        g = torch.empty(1, eng.n_branches, 4, 4, dtype=torch.float64, device=dev)
        mg = torch.empty(21, dtype=torch.float64, device=dev)
        st = torch.cuda.current_stream(dev).cuda_stream
        root_child = int(w.tree.post_triples[-1][1])

        def plain():
            eng.loglik_grad_device(bl.data_ptr(), out.data_ptr(), out.data_ptr() + 8, st)

        def everything():
            eng.branch_outer_products(bl.data_ptr(), out.data_ptr(), out.data_ptr() + 8, g.data_ptr(), st)
            eng.model_gradients(bl.data_ptr(), g.data_ptr(), root_child, mg.data_ptr(), st)

        ms = []
        for fn in (plain, everything):
            for _ in range(3):
                fn()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(dev)
            e0.record()
            for _ in range(steps):
                fn()
            e1.record()
            torch.cuda.synchronize(dev)
            ms.append(e0.elapsed_time(e1) / steps)
        return {"label": "f3: gradients w.r.t. GTR exchangeabilities, frequencies, category rates AND branch lengths in one pass",
                "config": {"workload": "gtr_1000x10k (headline tree and alignment, GTR)", "n_taxa": w.n_tips, "n_patterns": w.n_patterns},
                "ms_one_branch_gradient_evaluation": ms[0], "ms_all_gradients": ms[1], "ratio": ms[1] / ms[0], "steps": steps,
                "api": "b2p_branch_outer_products + b2p_model_gradients", "dlogl_dQ_diag": [float(x) for x in mg[[0, 5, 10, 15]].cpu()]}
    finally:
        eng.close()


# BASELINE.json configs besides the headline one, as short runs inside the same JSON line
SECONDARY_1GPU = [
    ("configs[0]", "jc69_20x1000", "float64", 0),
    ("configs[2] fp64", "hky_g4_flu", "float64", 0),
    ("configs[2] fp32", "hky_g4_flu", "float32", 0),
    ("north-star gate shape (5000 taxa x 10k patterns), GTR+G4", "gtr_g4_5000x10k", "float64", 0),
    ("north-star gate shape (5000 taxa x 10k patterns), JC69", "jc69_5000x10k", "float64", 0),
    ("configs[4]", "codon_500x2000", "float64", 0),
]


def run_b200(a):
    import torch
    import torch.distributed as dist

    ctx = Dist()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the engine has no CPU fallback")
    torch.cuda.set_device(ctx.local)
    dev = torch.device("cuda", ctx.local)
    stdout_fd = None
    if ctx.world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL writes its version banner to file descriptor 1 on first use: route everything but the
        # JSON line to stderr, so stdout carries exactly one line
        sys.stdout.flush()
        stdout_fd = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)
    if ctx.world != a.gpus and ctx.rank == 0:
        print(f"bench.py: --gpus {a.gpus} but WORLD_SIZE={ctx.world}; using {ctx.world}", file=sys.stderr)

    head = measure(a.workload, a.precision, a.patterns, a.steps, a.warmup, ctx, want_e2e=True,
                   cpu_seconds=0.0 if a.no_cpu_baseline else a.cpu_seconds, cpu_check=not a.no_cpu_baseline)
    secondary = []
    default_run = a.workload == "jc69_1000x10k" and a.precision == "float64" and not a.patterns
    if default_run and not a.no_secondary and os.environ.get("B2P_BENCH_SECONDARY", "1") != "0":
        if ctx.world == 1:
            todo = SECONDARY_1GPU
        else:
            # BASELINE.json configs[3]: GTR+G4, 5000 taxa x 100k patterns sharded over the N GPUs
            # (strong scaling: 100k / N patterns per rank; 64 GB of partials per GPU at N = 2)
            todo = [(f"configs[3]: 100k patterns over {ctx.world} GPUs", "gtr_g4_5000x100k", "float64", 100000 // ctx.world)]
        for label, name, prec, pat in todo:
            try:
                # configs[0] is the shape the reference's micro benchmark times on one CPU core
                # (/root/reference/nextflow.config:9): its entry carries the CPU port's rate too
                r = measure(name, prec, pat, min(a.steps, 5 if name == "gtr_g4_5000x100k" else 30), 3, ctx, want_e2e=ctx.world == 1,
                            cpu_seconds=1.0 if name == "jc69_20x1000" and not a.no_cpu_baseline else 0.0)
            except Exception as exc:  # a secondary run must not take the headline down with it
                r = {"error": f"{type(exc).__name__}: {exc}"} if ctx.rank == 0 else None
            if ctx.rank == 0 and r is not None:
                keep = {k: r[k] for k in ("value", "unit", "ms_per_step", "steps", "dtype", "config", "roofline", "clocks", "check",
                                          "e2e", "gpu_launches", "cpu_baseline", "error") if k in r}
                if name == "gtr_g4_5000x100k" and "value" in r:
                    keep["scaling"] = "strong"
                    keep["whole_alignment_evals_per_s"] = r["value"] / ctx.world
                secondary.append({"label": label, **keep})
        if ctx.world == 1:
            try:
                secondary.append(measure_parameter_gradients(dev))
            except Exception as exc:
                secondary.append({"label": "f3: parameter gradients", "error": f"{type(exc).__name__}: {exc}"})

    if ctx.rank == 0:
        line = {"metric": METRIC, "value": head["value"], "unit": head["unit"], "n_gpus": ctx.world, "steps": a.steps,
                "warmup": a.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": head["dtype"], "data": "synthetic"}
        for k in ("config", "timing", "engine", "e2e", "gpu_launches", "roofline", "clocks", "logl", "check"):
            if k in head:
                line[k] = head[k]
        if "cpu_baseline" in head:
            line["cpu_baseline"] = head["cpu_baseline"]
            line["cpu_baseline"]["logl_rel_diff_vs_gpu"] = head.get("check", {}).get("logl_rel_diff_vs_cpu")
        elif "check" in head and not a.no_cpu_baseline:
            # N > 1: the CPU port ran once per rank on that rank's shard (the check above); its timed
            # all-core number is the reference arm's (bench.py --impl reference), on rank 0 at N = 1 here
            line["cpu_baseline"] = {"value": None, "unit": head["unit"], "cores": host_cores(), "kind": "port",
                                    "sample": "not timed at N > 1 (rank 0 times it at N = 1; see --impl reference)"}
        if secondary:
            line["secondary"] = secondary
        if stdout_fd is not None:
            sys.stdout.flush()
            os.dup2(stdout_fd, 1)
        print(json.dumps(line), flush=True)
    if ctx.world > 1:
        dist.destroy_process_group()
    return 0


def main():
    a = parse_args()
    return run_reference(a) if a.impl == "reference" else run_b200(a)


if __name__ == "__main__":
    sys.exit(main())

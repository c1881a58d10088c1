#!/bin/bash
# multi-GPU evidence at N = $1: peer check (fused all-reduce vs NCCL vs CPU oracle), bench.py (headline weak scaling +
# BASELINE configs[3] sharded over N GPUs in `secondary`), the reference arm
N=$1
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$T --master-port 29533 tests/tools/peer_check.py > gpurun_out/r02_peer_check_${N}gpu.log 2>&1; echo "peer rc=$?"; grep -E "^rank" gpurun_out/r02_peer_check_${N}gpu.log | tail -4
$T --master-port 29534 bench.py --gpus $N --steps 100 --warmup 5 > gpurun_out/r02_bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_${N}gpu.err
$T --master-port 29535 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/r02_bench_${N}gpu_reference_arm.json 2>/dev/null; echo "ref rc=$?"

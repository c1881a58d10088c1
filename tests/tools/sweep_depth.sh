#!/bin/bash
mkdir -p gpurun_out
for d in 1 2 3; do for nt in 1024 512; do
echo "== DEPTH=$d NT=$nt"; B2P_DEPTH=$d B2P_NT=$nt timeout 120 python tests/tools/gpu_probe.py jc69_1000x10k hky_g4_flu --nocheck 2>&1 | python tests/tools/fmt_probe.py
done; done 2>&1 | tee gpurun_out/sweep_depth.log

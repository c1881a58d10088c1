#!/bin/bash
mkdir -p gpurun_out
P="python tests/tools/gpu_probe.py"
F="python tests/tools/fmt_probe.py"
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -5
for nt in 512 640 768; do
  echo "== B2P_NT=$nt"
  B2P_NT=$nt timeout 600 $P jc69_1000x10k hky_g4_flu --reps 30 2>&1 | $F
  B2P_NT=$nt timeout 600 $P gtr_g4_5000x10k jc69_1000x10k:5000:10000 --nocheck --reps 5 2>&1 | $F
  B2P_NT=$nt timeout 600 $P hky_g4_flu --f32 --reps 30 2>&1 | $F
done

#!/bin/bash
P="python tests/tools/gpu_probe.py"
F="python tests/tools/fmt_probe.py"
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_reference_data.py -m gpu -x -q 2>&1 | tail -5
for cfg in "B2P_X=0" "B2P_DEPTH=2" "B2P_DEPTH=3" "B2P_NP=1 B2P_DEPTH=2" "B2P_NP=3 B2P_DEPTH=2" "B2P_NT=640 B2P_DEPTH=2" "B2P_NT=768 B2P_DEPTH=2"; do
  echo "== $cfg"
  env $cfg timeout 600 $P jc69_1000x10k hky_g4_flu --reps 30 2>&1 | $F | cut -c1-150
  env $cfg timeout 600 $P gtr_g4_5000x10k --nocheck --reps 5 2>&1 | $F | cut -c1-150
done
export B2P_LIB=$PWD/gradient-benchmark_b200/b200phylo/libb200phylo_dbg.so
for cfg in "B2P_DEPTH=1" "B2P_DEPTH=2"; do
for dbg in 4 5 6 7; do
  echo "== $cfg B2P_DBG=$dbg"
  env $cfg B2P_DBG=$dbg timeout 600 $P jc69_1000x10k --nocheck --reps 10 2>&1 | $F | cut -c1-100
done
done

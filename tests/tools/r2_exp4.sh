#!/bin/bash
P="python tests/tools/gpu_probe.py"
F="python tests/tools/fmt_probe.py"
export B2P_LIB=$PWD/gradient-benchmark_b200/b200phylo/libb200phylo_dbg.so
for dbg in 0 1 2 3 4 7; do
  echo "== B2P_DBG=$dbg"
  B2P_DBG=$dbg timeout 600 $P jc69_1000x10k hky_g4_flu --nocheck --reps 20 2>&1 | $F | cut -c1-120
  B2P_DBG=$dbg timeout 600 $P gtr_g4_5000x10k --nocheck --reps 5 2>&1 | $F | cut -c1-120
done

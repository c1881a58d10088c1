#!/bin/bash
P="python tests/tools/gpu_probe.py"
F="python tests/tools/fmt_probe.py"
export B2P_LIB=$PWD/gradient-benchmark_b200/b200phylo/libb200phylo_dbg.so
for cfg in "B2P_DEPTH=1" "B2P_DEPTH=2" "B2P_DEPTH=3" "B2P_DEPTH=2 B2P_NP=3"; do
for dbg in 0 4 5 6 7; do
  echo "== $cfg B2P_DBG=$dbg"
  env $cfg B2P_DBG=$dbg timeout 600 $P jc69_1000x10k --nocheck --reps 10 2>&1 | $F | cut -c1-100
done
done

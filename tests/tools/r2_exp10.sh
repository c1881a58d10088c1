#!/bin/bash
P="python tests/tools/gpu_probe.py"
F="python tests/tools/fmt_probe.py"
for cfg in "B2P_NP=2" "B2P_NP=2 B2P_CT=136" "B2P_NP=3 B2P_CT=136" "B2P_NP=2 B2P_CT=136 B2P_DEPTH=2" "B2P_NP=2 B2P_CT=104" "B2P_NP=2 B2P_CT=112" "B2P_NP=2 B2P_CT=72" "B2P_NP=2 B2P_CT=64" "B2P_NP=2 B2P_CT=128" "B2P_NP=1" "B2P_NP=1 B2P_CT=136"; do
  echo "== gtr_g4_5000x10k $cfg"
  env $cfg timeout 600 $P gtr_g4_5000x10k --nocheck --reps 5 2>&1 | $F | cut -c1-170
done
for cfg in "B2P_NP=2" "B2P_NP=3" "B2P_NP=4" "B2P_NP=3 B2P_CT=40" "B2P_NP=3 B2P_CT=56" "B2P_NP=3 B2P_CT=64" "B2P_NP=3 B2P_MAT_ROWS=9" "B2P_NP=3 B2P_MAT_ROWS=14"; do
  echo "== hky_g4_flu $cfg"
  env $cfg timeout 600 $P hky_g4_flu --nocheck --reps 30 2>&1 | $F | cut -c1-170
  env $cfg timeout 600 $P hky_g4_flu --nocheck --f32 --reps 30 2>&1 | $F | cut -c1-170
done
for cfg in "B2P_DEPTH=2 B2P_NP=3" "B2P_DEPTH=2 B2P_NP=3 B2P_MAT_ROWS=9" "B2P_DEPTH=2 B2P_NP=3 B2P_MAT_ROWS=14" "B2P_DEPTH=2 B2P_NP=3 B2P_CT=72"; do
  echo "== jc69 $cfg"
  env $cfg timeout 600 $P jc69_1000x10k jc69_1000x10k:5000:10000 --nocheck --reps 10 2>&1 | $F | cut -c1-170
done

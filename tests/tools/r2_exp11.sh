#!/bin/bash
P="python tests/tools/gpu_probe.py"
F="python tests/tools/fmt_probe.py"
for cfg in "B2P_X=0" "B2P_NP=2" "B2P_NP=4" "B2P_DEPTH=1" "B2P_DEPTH=3" "B2P_NT=640" "B2P_NT=640 B2P_NP=2" "B2P_NT=640 B2P_NP=4" "B2P_NT=640 B2P_DEPTH=1" "B2P_NT=768" "B2P_NT=768 B2P_NP=4"; do
  echo "== $cfg"
  env $cfg timeout 600 $P jc69_1000x10k hky_g4_flu --nocheck --reps 30 2>&1 | $F | cut -c1-170
  env $cfg timeout 600 $P hky_g4_flu --nocheck --f32 --reps 30 2>&1 | $F | cut -c1-170
  env $cfg timeout 600 $P gtr_g4_5000x10k --nocheck --reps 5 2>&1 | $F | cut -c1-170
done

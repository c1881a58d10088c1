#!/bin/bash
# round-2 experiment 1: baseline vs kind-sorted plan; per-round clock trace.
mkdir -p gpurun_out
P="python tests/tools/gpu_probe.py"
F="python tests/tools/fmt_probe.py"
for srt in 0 1; do
  echo "== B2P_PLAN_SORT=$srt"
  B2P_PLAN_SORT=$srt timeout 600 $P jc69_1000x10k hky_g4_flu --reps 30 2>&1 | $F
  B2P_PLAN_SORT=$srt timeout 600 $P gtr_g4_5000x10k jc69_1000x10k:5000:10000 --nocheck --reps 5 2>&1 | $F
  B2P_PLAN_SORT=$srt timeout 600 $P hky_g4_flu --f32 --reps 30 2>&1 | $F
done
export B2P_LIB=$PWD/gradient-benchmark_b200/b200phylo/libb200phylo_trace.so
for w in jc69_1000x10k hky_g4_flu; do
  B2P_TRACE=gpurun_out/trace_$w.txt timeout 600 $P $w --nocheck --reps 2 > gpurun_out/trace_$w.probe 2>&1
  RP=$(python -c "
import ast,sys
d=ast.literal_eval(open('gpurun_out/trace_$w.probe').read().strip().splitlines()[-1]); print(d['info']['post_rounds'])")
  echo "== trace $w (post rounds $RP)"
  python tests/tools/trace_summary.py gpurun_out/trace_$w.txt $RP | head -12
done

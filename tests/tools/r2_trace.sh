#!/bin/bash
# per-round clock trace (debug build) of the workloads given as arguments; env passes through
mkdir -p gpurun_out
export B2P_LIB=$PWD/gradient-benchmark_b200/b200phylo/libb200phylo_trace.so
for w in "$@"; do
  tag=$(echo "$w" | tr ':' '_')
  B2P_TRACE=gpurun_out/trace_$tag.txt timeout 600 python tests/tools/gpu_probe.py $w --nocheck --reps 2 > gpurun_out/trace_$tag.probe 2>&1
  RP=$(python -c "
import ast
d=ast.literal_eval(open('gpurun_out/trace_$tag.probe').read().strip().splitlines()[-1]); print(d['info']['post_rounds'])")
  echo "== trace $w (post rounds $RP) $B2P_DEPTH"
  python tests/tools/trace2.py gpurun_out/trace_$tag.txt $RP
done

#!/bin/bash
# ncu --set full of one walk_kernel launch per workload spec given as arguments (name[:N[:P]][@f32])
mkdir -p gpurun_out
for spec in "$@"; do
  w=${spec%@*}; f32=""; [[ "$spec" == *@f32 ]] && f32="--f32"
  tag=$(echo "$spec" | tr ':@' '__')
  CMD="python tests/tools/gpu_probe.py $w --nocheck --reps 2 $f32"
  timeout 600 $CMD > gpurun_out/plain_$tag.log 2>&1 || { echo "plain run failed for $spec"; tail -5 gpurun_out/plain_$tag.log; continue; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:walk_kernel -s 4 -c 1 -f -o gpurun_out/ncu_$tag $CMD > gpurun_out/ncu_$tag.log 2>&1
  echo "ncu $spec rc=$?"
done

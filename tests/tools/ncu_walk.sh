#!/bin/bash
# ncu full capture of the walk kernel on the headline workload (after a plain run exits 0)
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:walk_kernel -s 3 -c 1 -f -o gpurun_out/prof_walk $CMD > gpurun_out/ncu_walk.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/plain.log | cut -c1-400

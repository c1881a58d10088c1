#!/bin/bash
# ncu full capture of the wave kernel on the headline workload (after a plain run exits 0)
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:wave_kernel -s 3 -c 1 -f -o gpurun_out/prof_wave $CMD > gpurun_out/ncu_wave.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/plain.log | cut -c1-600

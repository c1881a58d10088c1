#!/bin/bash
# `ncu --set full` of the outer-product walk (b2p_branch_outer_products) and its reduction kernel,
# each after its plain command exited 0; summaries only travel back (see r2_ncu_all.sh).
mkdir -p gpurun_out
P="python tests/tools/outer_probe.py"
cap() {  # tag, kernel regex, launches to skip, command...
  local tag=$1 rx=$2 skip=$3; shift 3
  timeout 600 "$@" > gpurun_out/plain_$tag.log 2>&1 || { echo "plain run failed: $tag"; tail -3 gpurun_out/plain_$tag.log; return; }
  cat gpurun_out/plain_$tag.log
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c 1 -f -o gpurun_out/ncu_$tag "$@" > gpurun_out/ncu_$tag.log 2>&1
  echo "ncu $tag rc=$?"
  python tests/tools/ncu_summary.py gpurun_out/ncu_$tag.ncu-rep gpurun_out/r02_ncu_${tag}_summary.txt > /dev/null 2>&1
  case $tag in walk_*) ncu -i gpurun_out/ncu_$tag.ncu-rep --page source --csv --print-source sass 2>/dev/null | gzip > gpurun_out/r02_ncu_${tag}_sass.csv.gz;; esac
  rm -f gpurun_out/ncu_$tag.ncu-rep gpurun_out/plain_$tag.log
}
# outer_probe.py: 35 plain walk_kernel launches, then the outer-product ones
cap walk_outer_f64_k1_gtr_1000x10k '^walk_kernel'         37 $P gtr_1000x10k
cap walk_outer_f64_k4_hky_g4_flu   '^walk_kernel'         37 $P hky_g4_flu
cap reduce_outer_hky_g4_flu        '^reduce_outer_kernel' 2  $P hky_g4_flu
cap model_grad_hky_g4_flu          '^model_grad_kernel'   2  python tests/tools/paramgrad_bench.py

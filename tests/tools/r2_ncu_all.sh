#!/bin/bash
# One `ncu --set full` capture per kernel of the path (each after its plain command exited 0).
# Reports land in gpurun_out/ncu_<tag>.ncu-rep; summarise here with tests/tools/ncu_summary.py.
mkdir -p gpurun_out
P="python tests/tools/gpu_probe.py"
cap() {  # tag, kernel regex, launches to skip, command...
  local tag=$1 rx=$2 skip=$3; shift 3
  timeout 600 "$@" > gpurun_out/plain_$tag.log 2>&1 || { echo "plain run failed: $tag"; tail -3 gpurun_out/plain_$tag.log; return; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c 1 -f -o gpurun_out/ncu_$tag "$@" > gpurun_out/ncu_$tag.log 2>&1
  echo "ncu $tag rc=$?"
  # the reports are ~10 MB each and gpurun_out/ travels back only up to 64 MiB: summarise here
  # (same ncu, same tool as at home) and keep only the text
  python tests/tools/ncu_summary.py gpurun_out/ncu_$tag.ncu-rep gpurun_out/r02_ncu_${tag}_summary.txt > /dev/null 2>&1
  case $tag in walk_*) ncu -i gpurun_out/ncu_$tag.ncu-rep --page source --csv --print-source sass 2>/dev/null | gzip > gpurun_out/r02_ncu_${tag}_sass.csv.gz;; esac
  rm -f gpurun_out/ncu_$tag.ncu-rep gpurun_out/plain_$tag.log
}
cap walk_f64_k1_jc69_1000x10k      '^walk_kernel'            4 $P jc69_1000x10k --nocheck --reps 2
cap walk_f64_k4_gtr_g4_5000x10k    '^walk_kernel'            4 $P gtr_g4_5000x10k --nocheck --reps 2
cap walk_f64_k4_hky_g4_flu         '^walk_kernel'            4 $P hky_g4_flu --nocheck --reps 2
cap walk_f32_k4_hky_g4_flu         '^walk_kernel'            4 $P hky_g4_flu --nocheck --reps 2 --f32
cap transition_gtr_g4_5000x10k     '^transition_kernel'      4 $P gtr_g4_5000x10k --nocheck --reps 2
cap reduce_gtr_g4_5000x10k         '^reduce_kernel'          4 $P gtr_g4_5000x10k --nocheck --reps 2
cap dense_transition_codon         '^dense_transition_kernel' 2 $P codon_500x2000 --nocheck --reps 2
cap dense_post_codon               '^dense_post_kernel'      38 $P codon_500x2000 --nocheck --reps 2
cap dense_pre_codon                '^dense_pre_kernel'       40 $P codon_500x2000 --nocheck --reps 2
cap dense_tip_codon                '^dense_tip_kernel'       2 $P codon_500x2000 --nocheck --reps 2
cap tt_heights                     '^tt_heights_kernel'      3 python tests/tools/timetree_bench.py
cap tt_heights_vjp                 '^tt_heights_vjp_kernel'  3 python tests/tools/timetree_bench.py
cap tt_coalescent                  '^tt_coalescent_kernel'   3 python tests/tools/timetree_bench.py
cap tt_log_jacobian                '^tt_log_jacobian_kernel' 3 python tests/tools/timetree_bench.py

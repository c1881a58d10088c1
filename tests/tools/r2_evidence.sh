#!/bin/bash
# Round-2 evidence in one go: TMA probe, per-round clock traces, debug-switch decomposition, ncu of the walk kernel.
mkdir -p gpurun_out
P="python tests/tools/gpu_probe.py"
F="python tests/tools/fmt_probe.py"
timeout 300 tests/tools/tma_probe > gpurun_out/r02_tma_probe.txt 2>&1
{ for w in jc69_1000x10k hky_g4_flu gtr_g4_5000x10k; do bash tests/tools/r2_trace.sh $w; done; } > gpurun_out/r02_trace_rounds.txt 2>&1
{ export B2P_LIB=$PWD/gradient-benchmark_b200/b200phylo/libb200phylo_dbg.so
  echo "# B2P_DBG bits: 1 skip the fold, 2 skip the stores, 4 skip the tasks (results are wrong by design; timing only)"
  for dbg in 0 1 2 3 4 7; do echo "== B2P_DBG=$dbg"; B2P_DBG=$dbg timeout 600 $P jc69_1000x10k gtr_g4_5000x10k --nocheck --reps 5 2>&1 | $F | cut -c1-110; done
  unset B2P_LIB; } > gpurun_out/r02_debug_switches.txt 2>&1
grep -vE "^cap (transition|reduce|dense|tt_)" tests/tools/r2_ncu_all.sh > /tmp/ncu_walk.sh; bash /tmp/ncu_walk.sh
python bench.py --workload gtr_g4_5000x100k --steps 5 --warmup 2 --no-secondary --cpu-seconds 1 > gpurun_out/r02_bench_gtr_g4_5000x100k_1gpu.json 2> gpurun_out/cfg4_1gpu.err; echo "cfg4 rc=$?"

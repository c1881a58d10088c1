#!/bin/bash
for W in jc69_1000x10k gtr_g4_5000x10k; do
for cfg in "B2P_X=0" "B2P_DEPTH=2" "B2P_NP=2" "B2P_NP=4" "B2P_NP=2 B2P_CW_LOADS=14" "B2P_NP=3 B2P_CW_LOADS=6" "B2P_DEPTH=2 B2P_NP=2"; do
  echo "== $W $cfg"; env $cfg timeout 300 python tests/tools/gpu_probe.py $W --nocheck --reps 10 2>&1 | tail -1 | python -c "
import sys,ast
d=ast.literal_eval(sys.stdin.read()); i=d['info']
print('ms_grad %.4f ms_eval %.4f gbs %.0f  ct %d wn %d rounds %d+%d pool %d' % (d['ms_traverse_kernel'], d['ms_eval'], d['gbs'], i['cols_per_tile'], i['tasks_per_round'], i['post_rounds'], i['pre_rounds'], i['pool_blocks']))"
done; done

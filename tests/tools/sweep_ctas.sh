#!/bin/bash
for W in jc69_1000x10k hky_g4_flu gtr_g4_5000x10k; do
for cfg in "B2P_X=0" "B2P_CTAS=2" "B2P_CTAS=2 B2P_NP=1" "B2P_CTAS=2 B2P_NP=3"; do
  echo "== $W $cfg"; env $cfg timeout 300 python tests/tools/gpu_probe.py $W --reps 10 2>&1 | tail -1 | python -c "
import sys,ast
d=ast.literal_eval(sys.stdin.read()); i=d['info']
print('ms_grad %.4f ms_eval %.4f gbs %.0f rel_logl %.1e  ct %d wn %d tiles %d grid %d rounds %d+%d pool %d regs %d' % (d['ms_traverse_kernel'], d['ms_eval'], d['gbs'], d.get('rel_logl',-1), i['cols_per_tile'], i['tasks_per_round'], i['n_tiles'], i['grid_blocks'], i['post_rounds'], i['pre_rounds'], i['pool_blocks'], i['regs_per_thread']))"
done; done

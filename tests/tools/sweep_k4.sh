#!/bin/bash
# Geometry sweep on the north-star gate shape (GTR+G4, 5000 taxa x 10k patterns).
W=${1:-gtr_g4_5000x10k}
run() { echo "== $*"; env "$@" timeout 300 python tests/tools/gpu_probe.py $W --nocheck --reps 5 2>&1 | tail -1 | python -c "
import sys,ast
d=ast.literal_eval(sys.stdin.read()); i=d['info']
print('ms_grad %.3f ms_eval %.3f gbs %.0f  ct %d wn %d tiles %d rounds %d+%d pool %d loads %d stores %d resident %d regs %d' % (d['ms_traverse_kernel'], d['ms_eval'], d['gbs'], i['cols_per_tile'], i['tasks_per_round'], i['n_tiles'], i['post_rounds'], i['pre_rounds'], i['pool_blocks'], i['block_loads'], i['block_stores'], i['resident_reads'], i['regs_per_thread']))"; }
run B2P_X=0
run B2P_CT=136
run B2P_CT=72
run B2P_CT=64
run B2P_CT=128
run B2P_CT=192
run B2P_NP=1
run B2P_NP=3
run B2P_CPT=1
run B2P_CT=136 B2P_NP=3

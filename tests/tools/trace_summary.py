"""Summarise a B2P_TRACE file (per-round clock stamps of CTA 0).  usage: trace_summary.py file rounds_post"""
import sys
import numpy as np
a = np.loadtxt(sys.argv[1]); rp = int(sys.argv[2])
for w in (0, 1):
    x = a[a[:, 0] == w][:, 2:]
    d = np.column_stack([x[:, 1] - x[:, 0], x[:, 2] - x[:, 1], x[:, 3] - x[:, 2], x[:, 4] - x[:, 3]])
    rt = np.diff(x[:, 0])  # round-to-round
    print("warp", "0" if w == 0 else "last", "cycles per round (start to start): post %.0f pre %.0f" % (rt[:rp].mean(), rt[rp:].mean()))
    print("  mean post: wait %.0f issue %.0f compute %.0f bar %.0f" % tuple(d[:rp].mean(0)))
    print("  mean pre : wait %.0f issue %.0f compute %.0f bar %.0f" % tuple(d[rp:].mean(0)))
    if w == 1 and x.shape[1] >= 7:
        c = np.column_stack([x[:, 5] - x[:, 1], x[:, 2] - x[:, 5], x[:, 6] - x[:, 2], x[:, 3] - x[:, 6]])
        print("  copy warp post: stores %.0f loads %.0f fold %.0f wait_group %.0f" % tuple(c[1:rp].mean(0)))
        print("  copy warp pre : stores %.0f loads %.0f fold %.0f wait_group %.0f" % tuple(c[rp:].mean(0)))
    if w == 0 and x.shape[1] >= 8:
        print("  compute warp 0, post rounds: desc+own load issue %.0f, operand reads %.0f" % ((x[:rp, 7] - x[:rp, 1]).mean(), (x[:rp, 2] - x[:rp, 7]).mean()))
        print("  compute warp 0, pre rounds : desc %.0f, operand reads + own load issue %.0f" % ((x[rp:, 7] - x[rp:, 1]).mean(), (x[rp:, 2] - x[rp:, 7]).mean()))
    if w == 0:
        for q in range(0, len(x), 10):
            print("  round %3d: wait %5d issue %5d compute %5d bar %5d  next-start %6d" % (q, *d[q], rt[q] if q < len(rt) else 0))

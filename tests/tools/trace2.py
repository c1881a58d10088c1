"""usage: trace2.py file rounds_post — per-round timeline of compute warp 0 and the last copy warp (CTA 0, tile 0).
stamps: 0 round start, 1 after the mbarrier wait, 2 compute: task done / copy: stores issued, 5 copy: loads issued,
6 copy: fold done, 3 before the barrier, 4 after the barrier."""
import sys
import numpy as np
a = np.loadtxt(sys.argv[1]); rp = int(sys.argv[2])
c = a[a[:, 0] == 0][:, 2:]   # compute warp 0
k = a[a[:, 0] == 1][:, 2:]   # last copy warp
n = min(len(c), len(k))
c, k = c[:n], k[:n]
for name, sl in (("post", slice(2, rp - 1)), ("pre", slice(rp, n - 1))):
    rt = (c[1:, 0] - c[:-1, 0])[sl]
    m = lambda x: x[sl].mean()
    print("%s rounds: %.0f cycles per round" % (name, rt.mean()))
    print("  compute w0: mbar wait %.0f | task %.0f | fence %.0f | barrier %.0f | to next round start %.0f" % (
        m(c[:, 1] - c[:, 0]), m(c[:, 2] - c[:, 1]), m(c[:, 3] - c[:, 2]), m(c[:, 4] - c[:, 3]), (c[1:, 0] - c[:-1, 4])[sl].mean()))
    print("  copy warp : mbar wait %.0f | loads issued %.0f | stores issued %.0f | fold %.0f | group waits + fence %.0f | barrier %.0f | to next round start %.0f" % (
        m(k[:, 1] - k[:, 0]), m(k[:, 5] - k[:, 1]), m(k[:, 2] - k[:, 5]), m(k[:, 6] - k[:, 2]), m(k[:, 3] - k[:, 6]), m(k[:, 4] - k[:, 3]),
        (k[1:, 0] - k[:-1, 4])[sl].mean()))
    if name == "pre" and k.shape[1] > 7:
        print("  copy warp fold: first group's loads + adds %.0f | rest (shuffles, stores, further groups) %.0f" % (m(k[:, 7] - k[:, 2]), m(k[:, 6] - k[:, 7])))
    print("  copy warp reaches the barrier %.0f cycles after compute warp 0" % m(k[:, 3] - c[:, 3]))

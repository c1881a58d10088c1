"""The macro benchmark's run at one of its sizes on synthetic data (no network for the flu set):
N dated taxa, `sites` JC69-simulated sites, 5000 ADVI iterations (/root/reference/modules/macro_flu.nf:6).
    python tests/tools/advi_macro.py [n_taxa] [sites] [iterations] [out.json]
"""
import io
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "gradient-benchmark_b200"), ROOT]
from b200phylo.advi import run_advi  # noqa: E402
from b200phylo.alignment import simulate_alignment, write_fasta  # noqa: E402
from b200phylo.substitution import JC69  # noqa: E402
from b200phylo.tree import simulate_tree  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
sites = int(sys.argv[2]) if len(sys.argv) > 2 else 1700
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 5000
tree = simulate_tree(n, shape="coalescent", seed=7, mean_branch=0.004)
time_tree = tree.scaled(1000.0)  # rate 0.001 subs/site/year
h = time_tree.node_heights()
tree.names = [f"t{i}_{2015.0 - h[i]:.6f}" for i in range(n)]
time_tree.names = list(tree.names)
m = JC69()
pm = m.p_t(tree.branch_lengths[:, None] * np.ones(1)[None, :])
aln = simulate_alignment(tree, pm, m.freqs, sites, seed=3)
with tempfile.TemporaryDirectory() as d:
    fa, nw = os.path.join(d, "a.fasta"), os.path.join(d, "t.nwk")
    write_fasta(fa, aln)
    open(nw, "w").write(time_tree.to_newick(17))
    buf = io.StringIO()
    res = run_advi(fa, nw, iters, lr=1e-4, rate_init=0.001, print_every=max(1, iters // 10), out=buf,
                   cuda_graph=os.environ.get("B2P_ADVI_EAGER") != "1")
print(buf.getvalue())
res["history"] = res["history"][-3:]
res["sites"] = sites
print(json.dumps(res))
if len(sys.argv) > 4:
    json.dump(res, open(sys.argv[4], "w"), indent=1)

"""Per-launch table of one evaluation from an `ncu --csv` launch list (dense-state path).

usage: python tests/tools/launch_table.py gpurun_out/launches.csv
"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
iK, iM, iV, iID, iG = (hdr.index(n) for n in ("Kernel Name", "Metric Name", "Metric Value", "ID", "Grid Size"))
d = collections.OrderedDict()
for r in rows[1:]:
    d.setdefault(r[iID], {"k": r[iK], "g": r[iG]})[r[iM]] = float(r[iV].replace(",", ""))
ids = list(d)
tr = [i for i in ids if "transition" in d[i]["k"]]
s, e = ids.index(tr[-2]), ids.index(tr[-1])
tot = collections.Counter()
for i in ids[s:e]:
    x = d[i]
    us = x["gpu__time_duration.sum"] / 1e3
    cyc = x.get("sm__cycles_elapsed.max", 0.0)
    wf = x.get("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", 0.0)
    print(x["k"][:26].ljust(26), x["g"].ljust(13),
          "%7.1f us  fp64 %5.1f%%  tensor %5.1f%%  dram %6.1f MB  l2 %6.1f MB  warps %4.1f%%  smem-wavefronts/cycle/SM %.2f" % (
              us, x.get("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", 0.0),
              x.get("sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active", 0.0),
              x["dram__bytes.sum"] / 1e6, x["lts__t_bytes.sum"] / 1e6,
              x["sm__warps_active.avg.pct_of_peak_sustained_active"], wf / max(cyc, 1.0) / 148))
    tot[x["k"].split("(")[0]] += us
print({k: round(v, 1) for k, v in tot.items()}, "total %.1f us" % sum(tot.values()))

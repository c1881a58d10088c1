"""Summarise an .ncu-rep (test tooling): key raw metrics, opcode mix, hottest source lines.
usage: python tests/tools/ncu_summary.py gpurun_out/prof.ncu-rep [out.txt]"""
import collections, csv, io, re, subprocess, sys

rep = sys.argv[1]
out = open(sys.argv[2], "w") if len(sys.argv) > 2 else sys.stdout
def run(*a):
    return subprocess.run(["ncu", "-i", rep, *a], capture_output=True, text=True).stdout
raw = list(csv.reader(io.StringIO(run("--page", "raw", "--csv"))))
hdr, units, vals = raw[0], raw[1], raw[2]
want = re.compile(r'^(Kernel Name|Block Size|Grid Size)$|dram__bytes_(read|write)\.sum$|gpu__time_duration.sum|lts__t_sector_hit_rate.pct|lts__throughput.avg.pct|sm__warps_active.avg.pct|launch__registers_per_thread$|launch__shared_mem_per_block_dynamic|smsp__average_warps_issue_stalled.*_per_issue_active|sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active|smsp__inst_executed.sum$|smsp__issue_active.avg.pct|sm__cycles_elapsed.max$|l1tex__data_pipe_lsu_wavefronts_mem_shared.sum$|sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active|l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum$|dram__throughput.avg.pct_of_peak_sustained_elapsed')
print(f"# {rep}", file=out)
for h, u, v in zip(hdr, units, vals):
    if want.search(h) and v != "":
        print(f"{h} [{u}] = {v}", file=out)
def I(x):
    try: return int(x)
    except ValueError: return 0
rows = list(csv.reader(io.StringIO(run("--page", "source", "--csv", "--print-source", "sass"))))
h = rows[1]; iS, iN, iP = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
op, ops = collections.Counter(), collections.Counter()
for r in rows[2:]:
    t = r[iS].split()
    o = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
    op[o] += I(r[iN]); ops[o] += I(r[iP])
tot, tots = sum(op.values()) or 1, sum(ops.values()) or 1
print(f"\n# opcode mix: total warp instructions {tot}, samples {tots}", file=out)
for o, c in op.most_common(24):
    print(f"{o:10s} {c / tot * 100:6.2f}% inst {ops[o] / tots * 100:6.2f}% samples", file=out)
rows = list(csv.reader(io.StringIO(run("--page", "source", "--csv", "--print-source", "cuda,sass"))))
cur, hd, lines = None, None, []
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hd = r; iN, iP = hd.index("Instructions Executed"), hd.index("# Samples"); continue
    if hd and r[0] != "" and len(r) > iN and r[iN].isdigit():
        lines.append((I(r[iN]), I(r[iP]), cur, r[0], r[1][:110]))
tot, tots = sum(l[0] for l in lines) or 1, sum(l[1] for l in lines) or 1
lines.sort(reverse=True)
print("\n# hottest source lines (by instructions executed; %samples = stall samples)", file=out)
for n, s_, f, l, src in lines[:45]:
    print(f"{n / tot * 100:5.2f}% {s_ / tots * 100:5.2f}%s {f}:{l} {src}", file=out)

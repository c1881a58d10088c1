"""Print the essentials of bench.py JSON lines (test tooling).  usage: show_bench.py file.json ..."""
import json
import sys
for f in sys.argv[1:]:
    d = json.load(open(f))
    if d.get("impl") == "reference":
        print(f, "reference arm:", round(d["value"], 2), "evals/s on", d["cpu_baseline"]["cores"], "cores,", d["config"]["total_patterns"], "patterns")
        continue
    r = d["roofline"]
    print(f"{f}: N={d['n_gpus']} value {d['value']:.1f}/s ({d['ms_per_step']:.4f} ms/step) e2e {d['e2e']['value']:.1f}/s ({d['e2e']['ms_per_step']:.4f} ms) "
          f"frac {r['frac']:.3f} kernel {r['kernel_ms']:.4f} ms share {r.get('share_of_step') or 0:.3f} clocks {d['clocks']['sm_mhz']} {d['clocks']['reasons']} check {d.get('check', {}).get('logl_rel_diff_vs_cpu')} {d.get('check', {}).get('grad_rel_diff_vs_cpu')}")
    for s in d.get("secondary", []):
        if "error" in s:
            print("   ", s["label"], "ERROR", s["error"]); continue
        if "ms_all_gradients" in s:
            print(f"    f3 all parameter gradients: {s['ms_all_gradients']:.4f} ms = {s['ratio']:.2f} x one branch-gradient evaluation ({s['ms_one_branch_gradient_evaluation']:.4f} ms)")
            continue
        rr = s["roofline"]
        print(f"    {s['label']}: {s['ms_per_step']:.4f} ms/step frac {rr['frac']:.3f}" + (f" issued {rr['issued_frac']:.3f}" if "issued_frac" in rr else "")
              + (f" whole-alignment {s['whole_alignment_evals_per_s']:.1f}/s" if "whole_alignment_evals_per_s" in s else "")
              + f" check {s['check']['logl_rel_diff_vs_cpu']:.1e} {s['check']['grad_rel_diff_vs_cpu']:.1e} clocks {s['clocks']['sm_mhz']} {s['clocks']['reasons']}")

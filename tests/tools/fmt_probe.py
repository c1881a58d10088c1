import sys, ast
for l in sys.stdin:
    try:
        d = ast.literal_eval(l)
    except Exception:
        print(l.rstrip()); continue
    i = d["info"]
    chk = "rel_logl=%.1e relg=%.1e " % (d["rel_logl"], d["max_rel_grad"]) if "rel_logl" in d else ""
    print(d["workload"], d["prec"], chk + "ms_k=%.4f ms_grad=%.4f ms_eval=%.4f gbs=%.0f gbs_eval=%.0f host=%.3f" % (
        d["ms_traverse_kernel"], d["ms_grad"], d["ms_eval"], d["gbs"], d["gbs_eval"], d["ms_host_e2e"]),
        "W", i["warps_per_block"], "cpt", i["cols_per_tile"], "bps", i["blocks_per_sm"], "grid", i["grid_blocks"],
        "tiles", i["n_tiles"], "rounds", i["post_rounds"], "regs", i["regs_per_thread"])

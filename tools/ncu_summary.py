"""Prints the key metrics of an `ncu --page raw --csv` export (one column per captured kernel)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}
want = ["Kernel Name", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__sass_inst_executed_op_shared_ld.sum", "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum",
        "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum.per_cycle_elapsed", "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum.per_cycle_elapsed",
        "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum.per_cycle_elapsed", "sm__cycles_elapsed.max"]
want += [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio")]
for w in want:
    if w not in idx:
        continue
    vals = [d[idx[w]] for d in data]
    if "stalled" in w:
        try:
            if max(float(v) for v in vals) < 0.15:
                continue
        except ValueError:
            pass
    name = w.replace("smsp__average_warps_issue_stalled_", "stall_").replace("_per_issue_active.ratio", "")
    print(f"{name[:70]:70s} [{units[idx[w]][:10]:>10s}] " + " | ".join(f"{v[:22]:>22s}" for v in vals))


def kernel_counters(path, B, N):
    """Per-knot DRAM traffic and executed fp64 work of the three solver kernels, from an `ncu --set full --page raw --csv`
    export of `tools/gpu_check.py prof` (sub-batches of B instances, horizon N).  Executed FLOP = thread-level DFMA x 2 +
    DADD + DMUL (+ 512 per DMMA m8n8k4 warp instruction, taken from the tensor pipe's active cycles at 16 per instruction).
    Written to profiles/kernel_counters.json by `python tools/ncu_summary.py <csv> --json <B> <N> <rev>`; bench.py reads it
    for `roofline.traffic` / `roofline.frac_executed` and records the file's revision."""
    rows = list(csv.reader(open(path)))
    hdr, data = rows[0], rows[2:]
    col = lambda name: hdr.index(name)
    fam = {"k_node_diff_coop": "node_diff", "k_backward": "backward", "k_forward_coop": "forward"}
    units = {"node_diff": B * (N + 1), "backward": B * N, "forward": B * (N + 1)}
    best = {}
    for r in data:
        name = r[col("Kernel Name")]
        f = next((v for k, v in fam.items() if k in name), None)
        t = float(r[col("gpu__time_duration.sum")])
        if f is None or (f in best and best[f][0] >= t):
            continue   # keep the longest launch of each family (the contact-mode instantiation; forward: pass 0)
        best[f] = (t, r, name)
    out = {}
    for f, (t, r, name) in best.items():
        cyc = float(r[col("sm__cycles_elapsed.max")])
        g = lambda m: float(r[col(m)]) if m in hdr else 0.0
        flop = (2 * g("smsp__sass_thread_inst_executed_op_dfma_pred_on.sum.per_cycle_elapsed") +
                g("smsp__sass_thread_inst_executed_op_dadd_pred_on.sum.per_cycle_elapsed") +
                g("smsp__sass_thread_inst_executed_op_dmul_pred_on.sum.per_cycle_elapsed")) * cyc
        nsmsp = 4 * float(r[col("launch__sm_count")]) if "launch__sm_count" in hdr else 592.0
        dmma = g("SM_C.TriageCompute.smsp__pipe_tensor_subpipe_dmma_cycles_active.avg") / 16.0 * nsmsp * 512.0
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
        rd = float(r[col("dram__bytes_read.sum")]) * scale[rows[1][col("dram__bytes_read.sum")]]
        wr = float(r[col("dram__bytes_write.sum")]) * scale[rows[1][col("dram__bytes_write.sum")]]
        out[f] = {"kernel": name, "us": t, "units_per_launch": units[f],
                  "dram_bytes_per_unit": (rd + wr) / units[f], "dram_read_per_unit": rd / units[f], "dram_write_per_unit": wr / units[f],
                  "flop_executed_per_unit": (flop + dmma) / units[f], "flop_dmma_per_unit": dmma / units[f],
                  "fp64_pipe_pct": g("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"),
                  "dmma_pipe_pct": g("sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active"),
                  "shared_wavefront_pct": g("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed"),
                  "issue_active_pct": g("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                  "warps_active_pct": g("sm__warps_active.avg.pct_of_peak_sustained_active"),
                  "registers": int(float(r[col("launch__registers_per_thread")]))}
    return out


if "--json" in sys.argv:
    import json, os
    i = sys.argv.index("--json")
    B, N, rev = int(sys.argv[i + 1]), int(sys.argv[i + 2]), sys.argv[i + 3]
    res = {"rev": rev, "source": os.path.basename(sys.argv[1]), "B_per_launch": B, "N": N,
           "unit": "knot (node_diff, forward: one rollout knot; backward: one knot step of the sweep)",
           "kernels": kernel_counters(sys.argv[1], B, N)}
    dst = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "kernel_counters.json")
    json.dump(res, open(dst, "w"), indent=1)
    print(json.dumps(res["kernels"], indent=1))

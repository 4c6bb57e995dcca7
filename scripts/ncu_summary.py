"""Summarises an `ncu --set full` report (one kernel launch per ID) into a small JSON for profiles/.
usage: python scripts/ncu_summary.py <report.ncu-rep> <out.json>"""
import csv, json, subprocess, sys

KEEP = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size", "launch__block_size",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
        "launch__occupancy_limit_registers", "launch__waves_per_multiprocessor", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.per_cycle_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed.sum", "smsp__inst_executed.sum",
        "sm__inst_issued.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__cycles_active.avg", "sm__cycles_elapsed.max",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum", "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum",
        "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum", "sm__sass_inst_executed_op_shared_ld.sum",
        "sm__sass_inst_executed_op_shared_st.sum", "sm__sass_inst_executed_op_global_ld.sum", "sm__sass_inst_executed_op_global_st.sum")


def main():
    rep, out = sys.argv[1:3]
    rows = list(csv.reader(subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"]).decode().split("\n")))
    hdr, units = rows[0], rows[1]
    res = {}
    for r in rows[2:]:
        if len(r) != len(hdr):
            continue
        d = dict(zip(hdr, r))
        u = dict(zip(hdr, units))
        name = "%s#%s" % (d.get("Kernel Name", "?").split("(")[0], d.get("ID", "?"))
        sel = {k: {"value": d[k], "unit": u[k]} for k in hdr if k in KEEP or k.startswith("smsp__average_warps_issue_stalled") or
               k.startswith("smsp__average_warp_latency_issue_stalled") or k.startswith("smsp__pcsamp_warps_issue_stalled")}
        res[name] = sel
    json.dump(res, open(out, "w"), indent=1, sort_keys=True)
    for n, s in res.items():
        print(n, {k: v["value"] for k, v in s.items() if k in ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
                                                                "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
                                                                "sm__warps_active.avg.pct_of_peak_sustained_active",
                                                                "smsp__issue_active.avg.pct_of_peak_sustained_active")})


if __name__ == "__main__":
    main()

"""Summarise an `ncu --set full` report (one kernel launch) into the text committed beside it
and, for the line-scan groundToImage kernel, into profiles/ncu_g2i.json (the DRAM traffic
bench.py reports as roofline.traffic).  Usage:
    python profiles/summarize_ncu.py gpurun_out/r1_g2i_full.ncu-rep POINTS "capture note" [--json]
(POINTS = 0 for a report holding launches of different sizes: no per-point line.)
Runs on the CPU box: `ncu -i` only reads the report."""
import csv
import io
import json
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
    "launch__occupancy_limit_registers", "sm__maximum_warps_per_active_cycle_pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum.per_cycle_elapsed",
    "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum.per_cycle_elapsed",
    "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum.per_cycle_elapsed",
    "sm__sass_thread_inst_executed_op_dfma_pred_on.sum.peak_sustained",
    "sm__icc_request_hit_rate.pct", "smsp__thread_inst_executed_per_inst_executed.ratio",
]


def main():
    rep, points, note = sys.argv[1], int(sys.argv[2]), sys.argv[3]
    raw = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"], text=True)
    rows = list(csv.reader(io.StringIO(raw)))
    head, unit = rows[0], rows[1]
    for vals in rows[2:]:
        v = dict(zip(head, vals))
        u = dict(zip(head, unit))
        print("## %s" % v["Kernel Name"])
        print("# %s; %d points in the launch" % (note, points))
        for k in KEYS:
            if k in v:
                print("%-90s %s %s" % (k, v[k], u[k]))
        for k in head:
            if "issue_stalled" in k and k.endswith("per_issue_active.ratio") and float(v[k] or 0) >= 0.1:
                print("%-90s %s %s" % (k, v[k], u[k]))
        dadd = float(v["smsp__sass_thread_inst_executed_op_dadd_pred_on.sum.per_cycle_elapsed"])
        dmul = float(v["smsp__sass_thread_inst_executed_op_dmul_pred_on.sum.per_cycle_elapsed"])
        dfma = float(v["smsp__sass_thread_inst_executed_op_dfma_pred_on.sum.per_cycle_elapsed"])
        peak = float(v["sm__sass_thread_inst_executed_op_dfma_pred_on.sum.peak_sustained"])
        print("FP64 thread-instructions per cycle (dadd+dmul+dfma) = %.1f of %.0f issue peak = %.1f %%"
              % (dadd + dmul + dfma, peak, 100 * (dadd + dmul + dfma) / peak))
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
        rd = float(v["dram__bytes_read.sum"]) * scale[u["dram__bytes_read.sum"]]
        wr = float(v["dram__bytes_write.sum"]) * scale[u["dram__bytes_write.sum"]]
        if points > 0:
            print("DRAM bytes per point: %.2f read + %.2f written = %.2f" % (rd / points, wr / points, (rd + wr) / points))
        if "--json" in sys.argv:
            with open("profiles/ncu_g2i.json", "w") as f:
                json.dump({"kernel": v["Kernel Name"], "points": points, "dram_bytes_read": rd,
                           "dram_bytes_write": wr, "capture": note,
                           "fp64_pipe_active_pct": float(v["sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"]),
                           "fp64_thread_inst_per_cycle": dadd + dmul + dfma, "fp64_issue_peak_per_cycle": peak,
                           "duration_ms": float(v["gpu__time_duration.sum"])}, f, indent=1)
        print()


if __name__ == "__main__":
    main()

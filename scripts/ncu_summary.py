"""Summarise an `ncu --page raw --csv` export (made here with `ncu -i x.ncu-rep --page raw --csv > raw.csv`) into the
few counters DESIGN.md and bench.py quote: usage  python scripts/ncu_summary.py raw.csv out.json [note]"""
import csv, json, sys
rows = list(csv.reader(open(sys.argv[1])))
v = dict(zip(rows[0], rows[2]))
u = dict(zip(rows[0], rows[1]))
KEYS = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct", "sm__inst_issued.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "TPC.TriageCompute.sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg",
        "TPC.TriageCompute.sm__inst_executed_pipe_alu_realtime.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_active.avg"]
out = {}
for k in KEYS:
    if k in v and v[k] != "":
        try:
            out[k] = {"value": float(v[k].replace(",", "")), "unit": u[k]}
        except ValueError:
            out[k] = v[k]
for k in v:  # every stall reason, TMA byte counter and tensor sub-pipe the capture has
    if ("issue_stalled" in k and k.endswith("per_issue_active.ratio")) or ("tma" in k and k.endswith(".sum")) or \
            ("pipe_tensor" in k and "pct" in k):
        try:
            out[k] = {"value": float(v[k].replace(",", "")), "unit": u[k]}
        except ValueError:
            pass
if len(sys.argv) > 3:
    out["note"] = sys.argv[3]
json.dump(out, open(sys.argv[2], "w"), indent=1)

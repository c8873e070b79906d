"""key rows of `ncu --set full` reports -> one CSV (profiles/r02_ncu_full_raw.csv) + the constants bench.py quotes
(profiles/roofline_traffic.json).   python tools/ncu_extract.py gpurun_out/*.ncu-rep"""
import csv, io, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
        "lts__t_sectors_lookup_miss.sum", "lts__t_sectors.sum", "l1tex__t_sector_hit_rate.pct",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "sm__cycles_active.avg", "sm__cycles_elapsed.max", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
rows_out = []
for rep in sys.argv[1:]:
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    if len(rows) < 3:
        continue
    hdr, units = rows[0], rows[1]
    ik = hdr.index("Kernel Name")
    for r in rows[2:]:
        d = {"report": os.path.basename(rep), "kernel": r[ik]}
        for w in WANT:
            if w in hdr:
                d[w] = r[hdr.index(w)] + " " + units[hdr.index(w)]
        rows_out.append(d)
out = os.path.join(ROOT, "profiles", "r02_ncu_full_raw.csv")
with open(out, "w", newline="") as f:
    wr = csv.writer(f)
    wr.writerow(["report", "kernel"] + WANT)
    for d in rows_out:
        wr.writerow([d["report"], d["kernel"]] + [d.get(w, "") for w in WANT])
print("wrote", out, len(rows_out), "rows")
for d in rows_out:
    print(d["report"], d["kernel"][:60], d.get("gpu__time_duration.sum"), d.get("dram__bytes_read.sum"), d.get("dram__bytes_write.sum"))

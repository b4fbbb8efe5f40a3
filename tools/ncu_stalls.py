"""Headline metrics + warp stall reasons of the kernels in an ncu report.  usage: python tools/ncu_stalls.py REPORT.ncu-rep"""
import csv, subprocess, sys

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, u = rows[0], rows[1]
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct"]
for r in rows[2:]:
    print("==", r[h.index("Kernel Name")][:70])
    for k in KEYS:
        if k in h:
            print("  %-62s %s %s" % (k, r[h.index(k)], u[h.index(k)]))
    st = [(float(r[i]), k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")) for i, k in enumerate(h)
          if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio") and "not_issued" not in k]
    print("  stalls per issue:", ", ".join("%s %.2f" % (k, v) for v, k in sorted(st, reverse=True)[:8]))

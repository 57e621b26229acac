"""Pick the columns of the per-kernel summary tables out of `ncu -i rep --page raw --csv`.
usage: python tools/ncu_raw_pick.py raw.csv > summary.csv"""
import csv, sys
COLS = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "lts__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active"]
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]; units = rows[1]
idx = [hdr.index(c) for c in COLS if c in hdr]
w = csv.writer(sys.stdout)
print("# units: " + ", ".join(units[i] or "-" for i in idx))
w.writerow([hdr[i] for i in idx])
tot_r = tot_w = 0.0; n = 0
for r in rows[2:]:
    if len(r) <= max(idx): continue
    out = [r[i] for i in idx]
    out[0] = out[0].split("(")[0][-40:]
    w.writerow(out)
print("# launches", len(rows) - 2)

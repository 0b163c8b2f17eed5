"""Summarise `ncu -i rep --page raw --csv` (one row per launch) into a small CSV for profiles/:
duration, DRAM bytes, tensor-pipe / issue activity, occupancy, registers, launch shape.
usage: python scripts/ncu_raw_summary.py raw.csv out.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1], errors="replace")))
hdr = rows[0]
idx = {h: i for i, h in enumerate(hdr)}
cols = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "duration_us"), ("dram__bytes_read.sum", "dram_read_MB"),
        ("dram__bytes_write.sum", "dram_write_MB"), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_throughput_pct"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_throughput_pct"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_pipe_active_pct"),
        ("smsp__issue_active.avg.pct", "issue_active_pct"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved_occupancy_pct"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("launch__shared_mem_config_size", "smem_config_KB")]
with open(sys.argv[2], "w", newline="") as f:
    w = csv.writer(f)
    w.writerow([c[1] for c in cols])
    w.writerow([rows[1][idx[s]] if s in idx else "" for s, _ in cols])
    for r in rows[2:]:
        out = []
        for src, name in cols:
            v = r[idx[src]] if src in idx else ""
            if name == "kernel":
                v = v.split("(")[0].replace("void abt::tc::", "").replace("abt::", "").replace("<unnamed>::", "").replace("void at::", "").replace("void ", "")
            out.append(v)
        w.writerow(out)
print("wrote", sys.argv[2])

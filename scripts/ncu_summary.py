"""Summarise an `ncu --set full` report (raw page) into a small CSV for profiles/: per launch the
duration, tensor-pipe activity, DRAM traffic and occupancy of the conv_tc kernels.
usage: ncu -i rep.ncu-rep --page raw --csv > raw.csv; python scripts/ncu_summary.py raw.csv out.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1], errors="replace")))
hdr = rows[0]
idx = {h: i for i, h in enumerate(hdr)}
cols = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "duration_us"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_pipe_active_pct"),
        ("dram__bytes_read.sum", "dram_read_MB"), ("dram__bytes_write.sum", "dram_write_MB"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_throughput_pct"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_throughput_pct"),
        ("launch__registers_per_thread", "regs"), ("sm__cycles_elapsed.max", "sm_cycles")]
with open(sys.argv[2], "w", newline="") as f:
    w = csv.writer(f)
    w.writerow([c[1] for c in cols])
    for r in rows[2:]:
        out = []
        for src, name in cols:
            v = r[idx[src]]
            if name == "kernel":
                v = v.split("(")[0].replace("void at::tc::", "").replace("void ", "")
            out.append(v)
        w.writerow(out)
print("wrote", sys.argv[2])

"""Per-kernel summary of ONE device-resident bench step from an ncu launch list (gpu__time_duration.sum, --csv):
the first window crop_resample_lattice_kernel ... eval_accumulate_kernel.  Per-launch times under ncu are cold-cache and
serialised: read the SHARES, not the absolutes.  usage: python scripts/launch_summary.py launches.csv out.csv"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr = rows[hi]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
L = []
for r in rows[hi + 1:]:
    v = float(r[vi].replace(",", ""))
    name = re.sub(r"\(.*", "", r[ki]).replace("void ", "").replace("abt::", "").replace("<unnamed>::", "")
    L.append((name, v / 1e3 if r[ui] == "ns" else v))
start = next(i for i, (n, _) in enumerate(L) if n.startswith("crop_resample_lattice") and any(m.startswith("eval_accumulate") for m, _ in L[i:i + 70]))
end = next(i for i in range(start, len(L)) if L[i][0].startswith("eval_accumulate"))
win = L[start:end + 1]
agg = collections.OrderedDict()
for n, v in win:
    a = agg.setdefault(n, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(v for _, v in win)
with open(sys.argv[2], "w") as f:
    f.write(f"# one device-resident step (launches {start}..{end} of {len(L)}); per-launch times are cold-cache and serialised: shares, not absolutes\n")
    f.write("kernel,launches,total_us,share\n")
    for n, (c, v) in agg.items():
        f.write(f"\"{n}\",{c},{v:.1f},{v / tot:.4f}\n")
    f.write(f"\"TOTAL\",{len(win)},{tot:.1f},1.0\n")
    enc = sum(v for n, v in win if n.startswith("tc::stem_mma") or (n.startswith("tc::conv_tc") and "9, 5, 1, 0, 4" not in n))
    f.write(f"# encoder (stem_mma + backbone conv_tc launches) {enc:.1f} us = {enc / tot:.3f} of the step\n")
print(open(sys.argv[2]).read())

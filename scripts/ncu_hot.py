"""Summarise an ncu --page source --csv dump: top SASS instructions by stall samples, with the
stall-reason columns, grouped in address order so that the warp role is recognisable."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ia, isrc, iall = hdr.index("Address"), hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") or h.startswith("Stall")]
data = rows[2:]
tot = sum(int(r[iall] or 0) for r in data)
print("total samples", tot)
top = sorted(range(len(data)), key=lambda k: -int(data[k][iall] or 0))[: int(sys.argv[2]) if len(sys.argv) > 2 else 25]
for k in sorted(top):
    r = data[k]
    s = int(r[iall] or 0)
    reasons = sorted(((int(r[i] or 0), h) for i, h in stall_cols if (r[i] or "0").isdigit()), reverse=True)[:3]
    print(f"{k:5d} {100.0 * s / tot:5.1f}%  {r[isrc].strip()[:70]:70s} {[(h.replace('stall_', ''), v) for v, h in reasons if v]}")

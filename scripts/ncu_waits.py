"""For an ncu SASS source dump: list every mbarrier wait with its stall samples and the role it
belongs to (guessed from the next UTCHMMA / UBLKCP / LDTM / UTCBAR in address order)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; isrc = hdr.index("Source"); iall = hdr.index("Warp Stall Sampling (All Samples)")
data = rows[2:]
tot = sum(int(r[iall] or 0) for r in data)
def role(k):
    for j in range(k, min(k + 400, len(data))):
        s = data[j][isrc]
        for key in ("UTCHMMA", "UBLKCP", "LDTM", "UTCBAR"):
            if key in s: return f"{key}@+{j-k}"
    return "?"
for k, r in enumerate(data):
    if "TRYWAIT" in r[isrc]:
        # samples on the wait and the following 3 instructions (the spin loop)
        s = sum(int(data[j][iall] or 0) for j in range(k, min(k + 4, len(data))))
        print(f"{k:5d} {100.0*s/tot:5.1f}%  {role(k)}")

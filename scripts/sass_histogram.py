"""Per-kernel SASS opcode histogram of libabsolutetrack_b200.so (evidence that the hot kernels are tcgen05 / TMA code):
UTCHMMA (tcgen05.mma), LDTM (tcgen05.ld), UBLKCP (cp.async.bulk), UTCBAR (tcgen05.commit), HMMA (mma.sync), LDGSTS
(cp.async), F2FP (16-bit pack), SYNCS (mbarrier).  usage: python scripts/sass_histogram.py > profiles/rN_sass_opcode_histogram.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "absolutetrack_b200", "libabsolutetrack_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", txt)), capture_output=True, text=True).stdout.split("\n")
WATCH = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "UBLKCP", "UTCBAR", "HMMA", "LDGSTS", "F2FP", "SYNCS", "DFMA", "FFMA", "LDS", "STS", "LDG", "STG"]
rows = []
for name, body in zip(names, re.split(r"Function : \S+", txt)[1:]):
    ops = collections.Counter()
    n = 0
    for m in re.finditer(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", body):
        op = m.group(1)
        n += 1
        ops[op.split(".")[0]] += 1
        if op.startswith("UTCHMMA.2CTA"):
            ops["UTCHMMA.2CTA"] += 1
    short = re.sub(r"\(.*", "", name.replace("(anonymous namespace)::", "")).replace("abt::", "")
    rows.append((short, n, ops))
print(f"# {os.path.basename(so)}: {len(rows)} kernels; columns: total SASS instructions, then " + " ".join(WATCH))
for short, n, ops in sorted(rows):
    print(f"{short[:70]:70s} {n:7d} " + " ".join(f"{ops.get(w, 0):6d}" for w in WATCH))
tot = collections.Counter()
for _, _, ops in rows:
    tot.update(ops)
print(f"{'TOTAL':70s} {sum(r[1] for r in rows):7d} " + " ".join(f"{tot.get(w, 0):6d}" for w in WATCH))

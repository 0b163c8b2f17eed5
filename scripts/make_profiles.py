"""Turn the raw outputs of scripts/gpu_round1_final.sh (gpurun_out/) into the tracked summaries under
profiles/.  usage: python scripts/make_profiles.py [round-prefix, default r1]"""
import collections
import csv
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "gpurun_out") + "/"
P = os.path.join(ROOT, "profiles") + "/"
R = sys.argv[1] if len(sys.argv) > 1 else "r1"


def copy(src, dst):
    if os.path.exists(G + src) and os.path.getsize(G + src) > 0:
        shutil.copy(G + src, P + dst)
        print("copied", dst)
    else:
        print("missing", src)


subprocess.check_call([sys.executable, os.path.join(ROOT, "scripts", "ncu_summary.py"), G + "tc_full_raw.csv",
                       P + f"{R}_conv_tc_ncu_full_summary.csv"])
subprocess.check_call([sys.executable, os.path.join(ROOT, "scripts", "ncu_raw_summary.py"), G + "step_other_raw.csv",
                       P + f"{R}_step_other_kernels_ncu_full_summary.csv"])
rows = list(csv.DictReader(open(P + f"{R}_conv_tc_ncu_full_summary.csv")))
rd = sum(float(x["dram_read_MB"]) for x in rows)
wr = sum(float(x["dram_write_MB"]) for x in rows)
print(len(rows), "conv launches: DRAM read", rd, "MB write", wr, "MB total", rd + wr, "MB; sum of durations",
      sum(float(x["duration_us"]) for x in rows), "us")

copy("bench.json", f"{R}_bench_fp16_tcgen05.json")
copy("bench_reference.json", f"{R}_bench_reference_arm.json")
copy("tc_seq.log", f"{R}_encoder_timeline_2048crops.txt")
copy("kernel_roofline.log", f"{R}_hbm_kernels_roofline.jsonl")
copy("ingest_trace.log", f"{R}_ingest_overlap_timeline.txt")

# launch list: one device-resident step (between two resampler launches of the `value` loop)
rows = list(csv.reader(open(G + "launches_bench.csv")))
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr = rows[hi]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
L = []
for r in rows[hi + 1:]:
    v = float(r[vi].replace(",", ""))
    L.append((r[ki], v / 1e3 if r[ui] == "ns" else v))
# the launch list itself, compacted (id, kernel, grid, block, duration)
gi, bi = hdr.index("Grid Size"), hdr.index("Block Size")
with open(P + f"{R}_bench_launch_list.csv", "w") as f:
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none -c 2500, python bench.py --steps 2 --warmup 3 "
            "--no-cpu-baseline: every launch of the command (kernel names shortened)\n")
    f.write("id,kernel,grid,block,duration_us\n")
    for n_, r in enumerate(rows[hi + 1:]):
        k = r[ki].split("(")[0].replace("void ", "").replace("at::<unnamed>::", "").replace("<unnamed>::", "").replace("at::", "")
        if "native::" in k:
            k = "torch:" + k.split("native::")[1].split("<")[0]
        f.write(f'{n_},"{k}","{r[gi]}","{r[bi]}",{L[n_][1]:.2f}\n')
idx = [i for i, (k, v) in enumerate(L) if "crop_resample_" in k]
a, b = idx[3], idx[4]


def short(k):
    k = k.split("(")[0].replace("void ", "").replace("at::<unnamed>::", "").replace("<unnamed>::", "").replace("at::", "")
    if "native::" in k:
        k = "torch elementwise (host mirror: metres -> mm)"
    return k


agg = collections.OrderedDict()
for k, v in L[a:b]:
    s = short(k)
    c, t = agg.get(s, (0, 0.0))
    agg[s] = (c + 1, t + v)
tot = sum(t for c, t in agg.values())
with open(P + f"{R}_bench_launches_summary.csv", "w") as f:
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none, python bench.py --steps 2 --warmup 3 "
            "--no-cpu-baseline (fp16, 2048 crops/step)\n")
    f.write(f"# one device-resident step (launches {a}..{b - 1} of {len(L)}); per-launch times are cold-cache and "
            "serialised: shares, not absolutes\n")
    f.write("kernel,launches,total_us,share\n")
    enc = 0.0
    for s, (c, t) in agg.items():
        f.write(f'"{s}",{c},{t:.1f},{t / tot:.4f}\n')
        if (s.startswith("tc::conv_tc_kernel") and "<9, 5" not in s) or "stem_pack" in s:
            enc += t
    f.write(f'"TOTAL",{sum(c for c, t in agg.values())},{tot:.1f},1.0\n')
    f.write(f"# encoder (stem_pack + backbone conv_tc launches) {enc:.1f} us = {enc / tot:.3f} of the step; live "
            "CUDA-event share in the same bench: encoder_ms / ms_per_step\n")
    gen = [v for k, v in L if "crop_cameras_" in k]
    pull = [v for k, v in L if "crop_pull_boxes" in k]
    if gen:
        f.write(f"# crop-camera generator (e2e_from_poses arm, side stream): {sum(gen) / (len(gen) / 3):.1f} us per call "
                f"(3 launches), {len(gen) // 3} calls seen\n")
    if pull:
        f.write(f"# source-box pull kernel (e2e arms, side stream): {sum(pull) / len(pull):.1f} us per launch when "
                f"serialised by ncu ({len(pull)} launches seen)\n")
print(open(P + f"{R}_bench_launches_summary.csv").read())

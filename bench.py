#!/usr/bin/env python
"""bench.py - hand-frames/s of the UmeTrack multi-view inference hot path on B200.

Contract: `python bench.py --gpus N --steps K --warmup W` (for N>1 launched under torchrun, one
rank per GPU).  One "step" advances S independent synthetic 4-view sequences by one frame:
crop resample (4S crops) -> encoder -> FTL/fusion/temporal/head (2S hand samples) -> FK/skinning
(2S x 21 landmarks).  Prints ONE JSON line on rank 0.

  value     : hand-frames/s, inputs resident in HBM, CUDA-event timed, max over ranks
  e2e       : same metric through BatchedHandTracker.step with pinned HOST buffers (H2D of the
              frames + parameters and D2H of poses + landmarks inside the timed region)
  roofline  : the per-view encoder (the dominant stage), FLOPs / CUDA-event time vs the measured
              bf16 tensor peak of MEASURED_PEAKS.json
  cpu_baseline : the oracle port of the reference's CPU path (numpy fp64 warp + fixed-point
              remap, torch-CPU fp32 network, numpy FK) on a bounded sample, rank 0, N=1 only
  --impl reference : that CPU arm alone, as its own JSON line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "hand_frames_per_sec"
UNIT = "hand-frames/s"
ENCODER_FLOP_PER_CROP = 969_228_288           # SURVEY.md 8(d), counted on the reference modules
TAIL_FLOP_PER_HAND = 19_150_000
FK_BYTES_PER_HAND = 404                        # shared hand model
CROP_SRC_BYTES = 32_160                        # mean source bounding box per crop
# dram__bytes_read.sum + dram__bytes_write.sum over the 25 conv_tc launches of ONE encoder pass for 2048 crops in the ncu --set full
# capture profiles/r2_conv_tc_ncu_full_summary.csv (7614 MB = the activations read + written once) + the stem kernel's
# algorithmic bytes (crops read, level-0 planes written: 334 MB; it was not part of that capture)
ENCODER_DRAM_BYTES_PER_CROP = 7961068032 / 2048   # dram read + write of the stem + 25 conv launches of one 2048-crop pass


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sust=d["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def mark(self):
        return time.time()

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                pass

    def summary(self, windows):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            if not any(a - 0.05 <= ts <= b + 0.15 for a, b in windows):
                continue
            p = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(p[0])); mx = max(mx, float(p[1]))
            except (ValueError, IndexError):
                continue
            for nm, v in zip(names, p[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------ CPU arm
def cpu_reference_arm(n_frames, threads, warm=1, with_crop_cameras=False):
    """Oracle port of the reference CPU path, frame by frame like run_eval_known_skeleton.py:68-90:
    [gen_crop_cameras when with_crop_cameras] + 4x _warp_image + regress_pose_use_skeleton (torch CPU
    fp32) + 2x landmarks_from_hand_pose.  Returns (hand_frames, seconds)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import torch
    import model_oracle as MO
    import np_oracle as O
    import scenario
    from absolutetrack_b200.lib.models.model_loader import make_random_state_dict

    torch.set_num_threads(threads)
    sd = make_random_state_dict(0)
    hm = scenario.hand_model()
    skel = (torch.from_numpy(hm["joint_rotation_axes"]).float(), torch.from_numpy(hm["joint_rest_positions"] * 0.001).float())
    st = MO.TemporalState()
    K = scenario.geometry()["crop_cam"].shape[0]
    frames = [scenario.synth_frame("noise", i) for i in range(4)]
    hands = 0
    valid = np.zeros(2, bool)
    G = scenario.geometry()
    hm_cc = dict(hm, joint_limits=G["hm_joint_limits"])
    t0 = None
    for t in range(warm + n_frames):
        if t == warm:
            t0, hands = time.perf_counter(), 0
        k = t % K
        fi = scenario.frame_inputs(k)
        if not fi["crops"]:
            continue
        frame = frames[t % 4]
        if with_crop_cameras:  # tracker.gen_crop_cameras (tracker.py:222-260) on the labelled poses of the frame
            cams = [dict(c2w=G["camera_to_world"][k, v], f=G["cam_f"][v], c=G["cam_c"][v], coeffs=G["cam_coeffs"][v],
                         width=int(G["cam_wh"][v, 0]), height=int(G["cam_wh"][v, 1])) for v in range(4)]
            for h in range(2):
                if G["hand_confidences"][k, h] >= 0.5:
                    O.gen_crop_cameras_from_pose(cams, G["camera_angles"], hm_cc, G["joint_angles"][k, h],
                                                 G["wrist_transforms"][k, h], h)
        imgs = [O.warp_image(scenario.src_camera(k, cam), scenario.dst_camera(k, h, s), np.ascontiguousarray(frame[:, cam, :]))
                for (h, s, cam) in fi["crops"]]
        images = torch.from_numpy(np.stack(imgs).astype(np.float32) / 255.0)
        use_mem = torch.from_numpy(valid[fi["hand_idx"]].copy())
        out = MO.forward(sd, st, images, torch.from_numpy(fi["intrinsics"]), torch.from_numpy(fi["extrinsics"]),
                         torch.from_numpy(fi["sample_range"]), torch.from_numpy(fi["hand_idx"]), use_mem,
                         torch.from_numpy(fi["hand_idx"]), skel)
        valid[:] = False
        valid[fi["hand_idx"]] = True
        for i, h in enumerate(fi["hand_idx"]):
            w = out["wrist_xfs"][i].numpy().copy()
            w[:3, 3] *= 1000.0
            O.landmarks_from_hand_pose(hm, out["joint_angles"][i].numpy(), w, int(h))
        hands += len(fi["hand_idx"])
    return hands, time.perf_counter() - t0


def run_reference_impl(args, rank):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    per_step = args.ref_frames
    warm_h, warm_s = cpu_reference_arm(max(2, per_step // 8), cores)
    total_h, total_s = 0, 0.0
    for _ in range(max(1, args.steps)):
        h, s = cpu_reference_arm(per_step, cores, warm=0)
        total_h += h; total_s += s
    v = total_h / total_s
    sample = f"{per_step} consecutive 2-hand frames per step (4 crops each), oracle port, torch threads={cores}"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total_s / max(1, args.steps), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "frames_per_step": per_step},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


WORKLOAD = ("synthetic 4-view fisheye frame stream (4x 636x480 u8 per frame, recording_00 geometry), both hands, "
            "known-skeleton head (BASELINE.json configs[3])")


class PcieRxSampler:
    """NVML's PCIe RX counter of one GPU (KB/s over the driver's 20 ms window), polled from a thread while the e2e
    region runs; best effort - returns None where NVML does not expose it (some VMs)."""

    def __init__(self, torch_index):
        self.samples, self.stop_flag, self.t, self.h = [], False, None, None
        try:
            import pynvml
            import torch

            pynvml.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(torch_index).uuid)
            self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid if uuid.startswith("GPU-") else "GPU-" + uuid)
            self.nv = pynvml
        except Exception:
            self.h = None

    def _run(self):
        while not self.stop_flag:
            try:
                self.samples.append(self.nv.nvmlDeviceGetPcieThroughput(self.h, self.nv.NVML_PCIE_UTIL_RX_BYTES) * 1e3 / 1e9)
            except Exception:
                return

    def __enter__(self):
        if self.h is not None:
            self.t = threading.Thread(target=self._run, daemon=True)
            self.t.start()
        return self

    def __exit__(self, *a):
        self.stop_flag = True
        if self.t is not None:
            self.t.join(timeout=1)

    def summary(self):
        v = sorted(x for x in self.samples if x > 0)
        return {"median": v[len(v) // 2], "max": v[-1], "samples": len(v)} if v else None


def bind_to_gpu_numa_node(index):
    """One process per GPU: run (and therefore allocate the pinned frame buffers) on the CPU cores next to
    that GPU, so that the source-box pull crosses PCIe without also crossing the socket interconnect.
    Returns a short description, or None when NVML / sched_setaffinity is unavailable."""
    try:
        import pynvml

        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        n = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (n + 63) // 64)
        cpus = [64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1 and 64 * i + b < n]
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if not allowed:
            return None
        os.sched_setaffinity(0, allowed)
        return f"{len(allowed)} cores near GPU {index} ({allowed[0]}-{allowed[-1]})"
    except Exception:
        return None


# ------------------------------------------------------------------------------------ GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--frames", type=int, default=512, help="sequences (= frames) per step per GPU")
    ap.add_argument("--precision", default=os.environ.get("AT_BENCH_PRECISION", "fp16"), choices=["fp32", "bf16", "fp16"],
                    help="encoder operand format; fp16 and bf16 run the same tcgen05 kind::f16 kernels with fp32 accumulation "
                         "(fp16 meets the 0.1 mm end-to-end tolerance, bf16 does not: DESIGN.md)")
    ap.add_argument("--map-mode", default="fp32", choices=["fp32", "fp64"])
    ap.add_argument("--ingest", default="pull", choices=["pull", "zero_copy", "copy"], help="e2e arm: how host frames reach the GPU")
    ap.add_argument("--pull-ctas", type=int, default=0, help="CTAs of the source-box pull kernel (0 = 2 per SM)")
    ap.add_argument("--ref-frames", type=int, default=24, help="frames per step of the CPU reference arm")
    ap.add_argument("--cpu-frames", type=int, default=96, help="bounded CPU-baseline sample (frames)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity block (timed path vs the CPU oracle, rank 0, N=1)")
    ap.add_argument("--parity-seqs", type=int, default=256, help="sequences of the parity block (x 8 steps x 2 hands)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference_impl(args, rank)
        return

    import numpy as np
    import torch
    import torch.distributed as dist

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback on the product arm)"
    # stdout carries exactly ONE line (the JSON): libraries that print to fd 1 (NCCL's version banner) go to stderr
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    numa = bind_to_gpu_numa_node(local_rank) if world > 1 else None
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    from absolutetrack_b200 import _cabi, synthetic
    from absolutetrack_b200.lib.common.hand_skinning import skin_landmarks_cuda
    from absolutetrack_b200.lib.models.model_loader import make_random_state_dict, model_from_state_dict
    from absolutetrack_b200.lib.tracker.batch_tracker import BatchedHandTracker
    from absolutetrack_b200.lib.tracker.tracker import crop_resample_cuda

    lib = _cabi.lib()
    peaks = load_peaks()
    S = args.frames
    model = model_from_state_dict(make_random_state_dict(0), precision=args.precision)
    hm = synthetic.hand_model()
    tracker = BatchedHandTracker(model, hm, max_slots=2 * S, map_mode=args.map_mode, ingest=args.ingest)
    tracker.pull_ctas = args.pull_ctas

    # sequences are sharded by rank (r::world, the reference Sampler's rule, async_dataset.py:558-559)
    seq_ids = np.arange(S) * world + rank
    NB = 2  # alternating input batches, each larger than L2 (S*1.22 MB)
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    host, devb, gts = [], [], []
    for i in range(NB):
        frames_dev = torch.randint(0, 256, (S, synthetic.H, synthetic.V, synthetic.W), dtype=torch.uint8, device=dev, generator=g)
        hb = synthetic.make_batch(seq_ids, i, frames=frames_dev.cpu(), first_step=False).pin()
        host.append(hb)
        db = hb.to(dev)
        db.frames = frames_dev
        devb.append(db)
        ja, wx = synthetic.gt_poses(seq_ids, i, hb)
        gts.append(skin_landmarks_cuda(tracker._hm_dev, ja.to(dev), wx.to(dev)))
    n = int(host[0].sample_range.shape[0])
    c = int(host[0].crop_params.shape[0])
    in_bytes = S * synthetic.H * synthetic.V * synthetic.W
    flush_note = f"inputs {in_bytes / 1e6:.0f} MB per step > 126 MB L2, {NB} alternating batches"

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    windows = []

    def timed(fn, steps, warmup):
        for i in range(warmup):
            fn(i)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = lib.at_launch_count()
        w0 = time.time()
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        barrier()
        windows.append((w0, time.time()))
        ms = e0.elapsed_time(e1)
        launches = lib.at_launch_count() - l0
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches

    # ---- whole pipeline, device-resident inputs
    from absolutetrack_b200.distributed import accumulate_metric_sums_cuda

    err_sum = torch.zeros(3, dtype=torch.float64, device=dev)  # {sum of per-hand MPJPE (mm), sum |dtheta|, hands}

    def step_dev(i):
        r = tracker.step_device(devb[i % NB], inputs_ready=None)   # resident inputs: the front stream need not wait for the main one
        accumulate_metric_sums_cuda(r.landmarks, gts[i % NB], err_sum)  # per-hand MPJPE vs GT, one launch

    tracker.encoder_events = []
    ms_total, launches = timed(step_dev, args.steps, args.warmup)
    tracker.check()
    # the encoder's share of the TIMED steps: CUDA events around at_encoder_forward in every step of the region above
    # (the roofline's `achieved`); the stand-alone stage loop further down is kept as stages.encoder_ms_alone
    enc_pairs, tracker.encoder_events = tracker.encoder_events[-args.steps:], None
    ms_enc_in_step = sum(a.elapsed_time(b) for a, b in enc_pairs) / max(1, len(enc_pairs))
    ms_step = ms_total / args.steps
    value = world * n / (ms_step * 1e-3)

    # ---- end to end from pinned host buffers
    out_host = [None]

    issue = [0.0]

    def step_host(i):
        t0 = time.perf_counter()
        out_host[0] = tracker.step(host[i % NB], out_host[0])
        issue[0] += time.perf_counter() - t0

    with PcieRxSampler(local_rank) as pcie:
        ms_e2e, _ = timed(step_host, args.steps, args.warmup)
    issue_ms = issue[0] / (args.steps + args.warmup) * 1e3  # host time to enqueue one step (must stay below the GPU time)
    e2e_value = world * n / (ms_e2e / args.steps * 1e-3)
    small = host[0].nbytes() - host[0].frames.numel()
    pulled0 = int(tracker.pulled_bytes.item())
    for i in range(2):
        step_host(i)
    torch.cuda.synchronize()
    pulled = (int(tracker.pulled_bytes.item()) - pulled0) // 2   # bytes the resampler pulled from the pinned frames per step
    ingest = {"pull": "pull (side-stream kernel copies each crop's source box from the pinned frames, overlapped with the previous step)",
              "zero_copy": "zero_copy (source boxes read from pinned frames by the resampler itself)",
              "copy": "copy (whole frames, side stream)"}[args.ingest]
    h2d = small + (pulled if pulled > 0 else host[0].frames.numel())
    d2h = sum(v.numel() * v.element_size() for v in out_host[0].values())

    # ---- end to end from what a recording provides: frames + camera poses + the pose each crop is centred on.
    # Crop cameras and every model input are generated on the device inside the timed region
    # (at_gen_crop_cameras = the reference's HandTracker.gen_crop_cameras + _make_inputs).
    e2e_poses = None
    if args.ingest == "pull":
        from absolutetrack_b200.lib.tracker.crop_cameras import gen_crop_inputs_cuda

        tracker.attach_rig(*synthetic.rig_cameras())
        pbs = [synthetic.make_pose_batch(seq_ids, i, frames=host[i].frames).pin() for i in range(NB)]
        out_p = [None]
        issue_p = [0.0]

        def step_poses(i):
            t0 = time.perf_counter()
            out_p[0] = tracker.step_poses(pbs[i % NB], out_p[0])
            issue_p[0] += time.perf_counter() - t0

        ms_p, launches_p = timed(step_poses, args.steps, args.warmup)
        tracker.check()
        assert out_p[0]["n"] == n and out_p[0]["c"] == c, "device crop cameras selected other views than the reference"
        pulled0 = int(tracker.pulled_bytes.item())
        step_poses(0); step_poses(1)
        torch.cuda.synchronize()
        pulled_p = (int(tracker.pulled_bytes.item()) - pulled0) // 2
        pdev = [getattr(pbs[0], f).to(dev) for f in ("joint_angles", "wrist_xforms", "pose_frame", "pose_hand", "camera_to_world")]
        pvalid = pbs[0].pose_valid.to(dev)
        gen_out = [None]

        def st_gen(i):
            gen_out[0] = gen_crop_inputs_cuda(tracker._rig, *pdev, tracker._cc_opts, pose_valid=pvalid, out=gen_out[0],
                                              workspace=getattr(gen_out[0], "_workspace", None))

        ms_gen, _ = timed(st_gen, args.steps, 3)
        e2e_poses = {"value": world * n / (ms_p / args.steps * 1e-3), "unit": UNIT,
                     "h2d_bytes_per_step": pbs[0].small_nbytes() + pulled_p,
                     "d2h_bytes_per_step": sum(v.numel() * v.element_size() for k, v in out_p[0].items() if hasattr(v, "numel")) + 16,
                     "api": "BatchedHandTracker.step_poses (pinned frames + poses + camera poses in; crop cameras and model inputs "
                            "generated on the device inside the timed region)",
                     "host_issue_ms_per_step": issue_p[0] / (args.steps + args.warmup + 2) * 1e3,
                     "crop_cameras_ms": ms_gen / args.steps, "gpu_launches": int(launches_p)}

    # ---- per-stage timings (device resident) for the roofline
    b0 = devb[0]
    crops = [None]
    feats = [None]

    box_bytes = torch.zeros(1, dtype=torch.int64, device=dev)  # source-box bytes the resampler staged (counted by the kernel)

    def st_crop(i):
        b = devb[i % NB]
        crops[0] = crop_resample_cuda(
            b.frames, synthetic.H * synthetic.V * synthetic.W, synthetic.W, synthetic.V * synthetic.W, b.crop_params,
            (96, 96), torch.uint8, args.map_mode, staged_bytes=box_bytes)

    def st_enc(i):
        feats[0] = model.encode(crops[0])

    ja_d = torch.zeros(n, 22, device=dev)
    wx_d = torch.eye(4, device=dev).repeat(n, 1, 1)

    def st_fk(i):
        skin_landmarks_cuda(tracker._hm_dev, ja_d, wx_d, hand_idx=b0.hand_idx)

    from absolutetrack_b200.lib.models.umetrack_model import InputFrameData, InputFrameDesc

    def st_tail(i):  # everything between the encoder and FK, timed directly on the encoder's features
        b = devb[i % NB]
        model.tail(feats[0], InputFrameData(None, b.intrinsics, b.extrinsics),
                   InputFrameDesc(b.sample_range, b.memory_idx, b.use_memory, b.hand_idx), tracker._skel)

    ms_crop, _ = timed(st_crop, args.steps, 3)
    box0 = int(box_bytes.item())
    st_crop(0)
    torch.cuda.synchronize()
    crop_src_bytes = int(box_bytes.item()) - box0       # of one launch over batch 0
    ms_enc, enc_launches = timed(st_enc, args.steps, 3)
    ms_tail, tail_launches = timed(st_tail, args.steps, 3)
    ms_fk, _ = timed(st_fk, args.steps, 3)
    ms_crop /= args.steps; ms_enc /= args.steps; ms_fk /= args.steps; ms_tail /= args.steps

    ms_enc_alone, ms_enc = ms_enc, ms_enc_in_step
    enc_tflops = ENCODER_FLOP_PER_CROP * c / (ms_enc * 1e-3) / 1e12
    tf_peak = peaks["tf_burst"]  # the encoder is timed alone, in short bursts -> burst denominator
    crop_bytes = crop_src_bytes + c * (96 * 96 * 1 + 256)   # source boxes (summed by the kernel) + u8 crops + parameter records
    crop_gbs = crop_bytes / (ms_crop * 1e-3) / 1e9
    fk_gbs = n * FK_BYTES_PER_HAND / (ms_fk * 1e-3) / 1e9

    if world > 1:
        dist.all_reduce(err_sum)  # the only collective of the path: metric sums over shards
    mpjpe = float((err_sum[0] / err_sum[2]).item())

    clocks = None
    if rank == 0:
        time.sleep(0.15)
        sampler.stop()
        clocks = sampler.summary(windows)

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        hnd, sec = cpu_reference_arm(args.cpu_frames, cores)
        hnd_cc, sec_cc = cpu_reference_arm(args.cpu_frames, cores, with_crop_cameras=True)
        cpu_baseline = {"value": hnd / sec, "value_with_crop_cameras": hnd_cc / sec_cc, "unit": UNIT, "cores": cores, "kind": "port",
                        "sample": f"{args.cpu_frames} consecutive 2-hand frames of the same stream, frame by frame "
                                  f"(numpy fp64 warp + fixed-point remap, torch-CPU fp32 network with {cores} threads, numpy FK); "
                                  "value_with_crop_cameras adds the reference's per-frame gen_crop_cameras (the baseline of e2e_from_poses)"}

    parity_block = None
    if rank == 0 and world == 1 and not args.no_parity:
        # the timed configuration against the CPU oracle (oracle/ is test infrastructure: here only as the checker)
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import parity

        parity_block = parity.timed_path_parity(S=args.parity_seqs, T=8, precision=args.precision, map_mode=args.map_mode,
                                                ingest=args.ingest)

    if rank == 0:
        sys.stdout.flush()
        os.dup2(json_fd, 1)
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": args.precision, "data": "synthetic",
            "config": {"workload": WORKLOAD, "frames_per_step_per_gpu": S, "hand_frames_per_step_per_gpu": n,
                       "crops_per_step_per_gpu": c, "map_mode": args.map_mode, "l2": flush_note,
                       "weights": "seeded random state dict (252 keys), BN folded", "sharding": "sequences r::world",
                       "host_numa_binding": numa},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": "BatchedHandTracker.step (pinned host buffers)", "ingest": ingest, "host_issue_ms_per_step": issue_ms,
                    "host_frame_bytes_per_step": int(host[0].frames.numel()),
                    # the limiter of this arm at N = 8: bytes each GPU pulls over PCIe per second of the timed region
                    # (counted by the pull kernel), and NVML's RX counter of rank 0's GPU over the same region
                    "pcie_rx_gbs_per_gpu": h2d / (ms_e2e / args.steps * 1e-3) / 1e9, "pcie_rx_gbs_nvml": pcie.summary()},
            "e2e_from_poses": e2e_poses,
            "gpu_launches": int(launches),
            "roofline": {"bound": "tensor", "kernel": "per-view encoder (all conv launches of at_encoder_forward)",
                         "achieved": enc_tflops, "peak": tf_peak, "unit": "TFLOP/s", "frac": enc_tflops / tf_peak,
                         "traffic": ENCODER_DRAM_BYTES_PER_CROP * c,
                         "traffic_source": "bytes per encoder pass (all conv launches), scaled per crop from the ncu --set full "
                                           "capture in profiles/r2_conv_tc_s2d_ncu_full_summary.csv (stem_mma + the 25 backbone conv launches)",
                         "peak_source": peaks["src"] + " cuBLAS bf16 burst (MEASURED_PEAKS.json bf16_tflops)",
                         "frac_of_sustained_peak": enc_tflops / peaks["tf_sust"],
                         "launches_per_step": enc_launches // max(1, args.steps)},
            "stages": {"crop_ms": ms_crop, "encoder_ms": ms_enc, "encoder_ms_alone": ms_enc_alone, "tail_ms": ms_tail, "fk_ms": ms_fk,
                       "tail_launches": tail_launches // max(1, args.steps),
                       "note": "encoder_ms: CUDA events around the encoder call inside every step of the timed region (mean over the steps); "
                               "the other stages and encoder_ms_alone: each stage looped on its own (the tail directly on the encoder's "
                               "features, not by subtraction)",
                       "crop_gbs": crop_gbs, "crop_hbm_frac": crop_gbs / peaks["hbm"],
                       "crop_source_box_bytes_per_crop": crop_src_bytes / max(1, c),
                       "fk_gbs": fk_gbs, "fk_hbm_frac": fk_gbs / peaks["hbm"]},
            "parity": parity_block,
            "mpjpe_vs_gt_mm_random_weights": mpjpe,
            "cpu_baseline": cpu_baseline,
            "clocks": clocks,
        }))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

"""BASELINE.json configs[4]: the synthetic 1M-frame hand-tracking sweep - 4,096 sequences x 244 frames (999,424 frames,
1,998,848 hand-frames) with carried temporal memory, sequences sharded r::world over the ranks (the reference Sampler's
rule, async_dataset.py:558-559), one NCCL all-reduce of the metric sums at the end.

  python scripts/sweep_1m.py                                  (1 GPU)
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P scripts/sweep_1m.py

Each rank advances its sequences in lock-step groups of 512 (one BatchedHandTracker step per group and frame; memory slot
= (group, sequence, hand)).  Geometry cycles the 123 labelled frames of tests/golden/geometry.npz per sequence; images are
two alternating device-resident noise frame sets (625 MB each > L2).  `--e2e` adds a second pass from pinned HOST frames
(source-box pull inside the timed region).  Prints one JSON line on rank 0; times are CUDA events, max over ranks.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sequences", type=int, default=4096)
    ap.add_argument("--frames", type=int, default=244)
    ap.add_argument("--group", type=int, default=512)
    ap.add_argument("--e2e", action="store_true")
    args = ap.parse_args()
    import numpy as np
    import torch
    import torch.distributed as dist

    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from absolutetrack_b200 import synthetic
    from absolutetrack_b200.distributed import accumulate_metric_sums_cuda
    from absolutetrack_b200.lib.common.hand_skinning import skin_landmarks_cuda
    from absolutetrack_b200.lib.models.model_loader import make_random_state_dict, model_from_state_dict
    from absolutetrack_b200.lib.tracker.batch_tracker import BatchedHandTracker, FrameBatch

    my_seqs = np.arange(rank, args.sequences, world)
    S = min(args.group, len(my_seqs))
    groups = [my_seqs[i:i + S] for i in range(0, len(my_seqs), S)]
    K = synthetic.geometry()["crop_cam"].shape[0]
    model = model_from_state_dict(make_random_state_dict(0), precision="fp16")
    tracker = BatchedHandTracker(model, synthetic.hand_model(), max_slots=2 * len(my_seqs), map_mode="fp32", ingest="pull")
    g = torch.Generator(device=dev).manual_seed(99 + rank)
    dev_frames = [torch.randint(0, 256, (S, synthetic.H, synthetic.V, synthetic.W), dtype=torch.uint8, device=dev, generator=g) for _ in range(2)]
    host_frames = [f.cpu().pin_memory() for f in dev_frames] if args.e2e else None
    # per (group, geometry phase): the small tensors of a step (crop records, camera matrices, ranges, memory slots) + GT landmarks
    small, gts = [], []
    placeholder = torch.zeros(1, dtype=torch.uint8)
    for gi, seqs in enumerate(groups):
        row, grow = [], []
        for t in range(min(K, args.frames)):
            hb = synthetic.make_batch(seqs, t, frames=placeholder)   # (without `frames` make_batch allocates a 625 MB dummy)
            ja, wx = synthetic.gt_poses(seqs, t, hb)
            hb.memory_idx = hb.memory_idx + 2 * S * gi
            grow.append(skin_landmarks_cuda(tracker._hm_dev, ja.to(dev), wx.to(dev)))
            row.append(hb)
        small.append(row)
        gts.append(grow)
    n_per_step = [int(b.sample_range.shape[0]) for b in small[0]]
    first = [torch.zeros(int(r[0].sample_range.shape[0]), dtype=torch.bool) for r in small]

    def batches(on_device):
        out = []
        for gi, row in enumerate(small):
            o = []
            for t, hb in enumerate(row):
                fr = (dev_frames if on_device else host_frames)[t & 1][: len(groups[gi])]
                b = FrameBatch(fr, hb.crop_params, hb.intrinsics, hb.extrinsics, hb.sample_range, hb.hand_idx, hb.memory_idx, hb.use_memory)
                o.append(b.to(dev) if on_device else FrameBatch(fr, *[getattr(hb, f).pin_memory() for f in list(hb.__dataclass_fields__)[1:]]))
                if on_device:
                    o[-1].frames = fr
            out.append(o)
        return out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run(bs, on_device):
        model.reset_temporal_state()
        sums = torch.zeros(3, dtype=torch.float64, device=dev)
        firsts = [f.to(dev) if on_device else f.pin_memory() for f in first]
        hands = 0
        out_host = [None] * len(groups)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for t in range(args.frames):
            for gi in range(len(groups)):
                b = bs[gi][t % K]
                if t == 0:  # first frame of every sequence: no temporal memory (temporal.py:59-63)
                    b = FrameBatch(b.frames, b.crop_params, b.intrinsics, b.extrinsics, b.sample_range, b.hand_idx, b.memory_idx, firsts[gi])
                if on_device:
                    r = tracker.step_device(b, inputs_ready=None)
                    accumulate_metric_sums_cuda(r.landmarks, gts[gi][t % K], sums)
                else:
                    out_host[gi] = tracker.step(b, out_host[gi])
                hands += int(b.sample_range.shape[0])
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        tracker.check()
        if world > 1:
            tt = torch.tensor([ms, float(hands)], device=dev, dtype=torch.float64)
            mx = tt.clone()
            dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            dist.all_reduce(tt, op=dist.ReduceOp.SUM)
            dist.all_reduce(sums)  # the path's only collective
            ms, hands = float(mx[0]), int(tt[1])
        return ms, hands, sums

    dbs = batches(True)
    # warm-up: 3 steps of group 0 (kernels loaded, workspaces allocated), state reset by run()
    for t in range(3):
        tracker.step_device(dbs[0][t])
    ms, hands, sums = run(dbs, True)
    line = {"workload": f"{args.sequences} sequences x {args.frames} frames, carried temporal memory, groups of {S} per step, fp16 encoder, fp32 lattice crop map",
            "n_gpus": world, "frames": args.sequences * args.frames, "hand_frames": hands, "seconds": ms * 1e-3,
            "value": hands / (ms * 1e-3), "unit": "hand-frames/s", "steps_per_rank": args.frames * len(groups),
            "whole_step_tensor_tflops": hands * 1.9576e9 / (ms * 1e-3) / 1e12 / world,
            "mpjpe_vs_gt_mm_random_weights": float(sums[0] / sums[2]) if float(sums[2]) else None, "scaling": "strong (fixed 1M frames)"}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            pk = json.load(f)
        line["whole_step_frac_of_sustained_bf16_peak"] = line["whole_step_tensor_tflops"] / pk["bf16_tflops_sustained"]
    except OSError:
        pass
    if args.e2e:
        hbs = batches(False)
        for t in range(3):
            tracker.step(hbs[0][t])
        ms2, hands2, _ = run(hbs, False)
        line["e2e"] = {"value": hands2 / (ms2 * 1e-3), "seconds": ms2 * 1e-3, "api": "BatchedHandTracker.step from pinned host frames (source-box pull)"}
    if rank == 0:
        sys.stdout.flush()
        os.dup2(json_fd, 1)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

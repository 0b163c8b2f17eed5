"""Per-stage CPU baselines and the reference's deployment shape on this box's host
cores, from the oracle port of the reference path (oracle/ is test infrastructure; this script is a measurement tool).
  per stage: _warp_image per crop (numpy fp64 map + fixed-point remap), cv2.remap alone (when cv2 is installed),
             backbone forward at batch 1 / 4 / 256 (torch CPU fp32), skin_landmarks batched
  per frame: one process with all cores (bench.py's cpu_baseline shape) and Pool(8) x cores/8 torch threads
             (run_eval_known_skeleton.py:117-119)
usage: python scripts/cpu_baselines.py out.json"""
import json
import os
import sys
import time
from multiprocessing import get_context

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def _pool_worker(args):
    n_frames, threads = args
    import bench

    h, s = bench.cpu_reference_arm(n_frames, threads, warm=1)
    return h, s


def main():
    import numpy as np
    import torch

    import bench
    import model_oracle as MO
    import np_oracle as O
    import scenario
    from absolutetrack_b200.lib.models.model_loader import make_random_state_dict

    cores = os.cpu_count() or 1
    res = {"cores": cores, "torch_threads_default": torch.get_num_threads()}
    G = scenario.geometry()
    # ---- _warp_image per crop
    frame = scenario.synth_frame("noise", 0)
    jobs = [(k, h, s, int(G["crop_cam"][k, h, s])) for k in range(0, 40) for h in range(2) for s in range(2) if G["crop_cam"][k, h, s] >= 0][:64]
    t0 = time.perf_counter()
    maps = []
    for (k, h, s, cam) in jobs:
        maps.append(O.warp_map(scenario.src_camera(k, cam), scenario.dst_camera(k, h, s)))
    t_map = (time.perf_counter() - t0) / len(jobs)
    t0 = time.perf_counter()
    for (k, h, s, cam), (mx, my) in zip(jobs, maps):
        O.remap_bilinear_u8(np.ascontiguousarray(frame[:, cam, :]), mx, my)
    t_remap_np = (time.perf_counter() - t0) / len(jobs)
    res["warp_image_ms_per_crop"] = {"numpy_fp64_map": 1e3 * t_map, "numpy_fixed_point_remap": 1e3 * t_remap_np, "crops": len(jobs)}
    try:
        import cv2

        t0 = time.perf_counter()
        for _ in range(20):
            for (k, h, s, cam), (mx, my) in zip(jobs, maps):
                cv2.remap(np.ascontiguousarray(frame[:, cam, :]), mx, my, cv2.INTER_LINEAR)
        res["warp_image_ms_per_crop"]["cv2_remap_alone"] = 1e3 * (time.perf_counter() - t0) / (20 * len(jobs))
    except ImportError:
        res["warp_image_ms_per_crop"]["cv2_remap_alone"] = None
    # ---- backbone at batch 1 / 4 / 256
    sd = make_random_state_dict(0)
    torch.set_num_threads(cores)
    res["backbone_ms"] = {}
    for bs, reps in ((1, 20), (4, 10), (256, 2)):
        x = torch.rand(bs, 96, 96)
        MO.backbone(sd, x)
        t0 = time.perf_counter()
        for _ in range(reps):
            MO.backbone(sd, x)
        dt = (time.perf_counter() - t0) / reps
        res["backbone_ms"][f"batch_{bs}"] = {"ms": 1e3 * dt, "crops_per_s": bs / dt, "gflops": 0.969228288 * bs / dt}
    # ---- skin_landmarks batched
    hm = scenario.hand_model()
    n = 65536
    rng = np.random.default_rng(0)
    ja = rng.uniform(-0.5, 0.5, (n, 22)).astype(np.float32)
    wx = np.tile(np.eye(4, dtype=np.float32), (n, 1, 1))
    O.skin_landmarks(hm, ja[:1024], wx[:1024])
    t0 = time.perf_counter()
    O.skin_landmarks(hm, ja, wx)
    dt = time.perf_counter() - t0
    res["skin_landmarks_numpy"] = {"hands": n, "hands_per_s": n / dt}
    # ---- per-frame path: one process x all cores, and Pool(8) x cores/8 threads
    h, s = bench.cpu_reference_arm(48, cores)
    res["per_frame_one_process"] = {"hand_frames_per_s": h / s, "threads": cores, "frames": 48}
    per = max(1, cores // 8)
    with get_context("spawn").Pool(8) as p:
        t0 = time.perf_counter()
        outs = p.map(_pool_worker, [(24, per)] * 8)
        wall = time.perf_counter() - t0
    res["per_frame_pool8"] = {"hand_frames_per_s_sum_of_workers": sum(h / s for h, s in outs),
                              "hand_frames_per_s_wall_incl_startup": sum(h for h, _ in outs) / wall,
                              "threads_per_worker": per, "frames_per_worker": 24,
                              "shape": "Pool(8) over recordings, run_eval_known_skeleton.py:117-119"}
    with open(sys.argv[1], "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()

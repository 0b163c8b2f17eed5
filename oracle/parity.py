"""Parity of the EXACT path bench.py times against the CPU oracle - TEST INFRASTRUCTURE (the checker).

bench.py times BatchedHandTracker with the 16-bit tcgen05 encoder, the fp32 lattice crop map and
pinned-host frames pulled by the side-stream kernel.  `timed_path_parity` runs that configuration for T
time steps of S lock-step sequences (temporal memory carried, first step without memory), half of the
sequences on i.i.d. noise frames and half on smooth frames, and compares with the reference arithmetic
restated in oracle/ on the same inputs:

  crops      lattice-map crops of the timed path vs tracker._warp_image (tracker.py:61-89):
             every crop against the CUDA float64-map mode (itself pinned bit-exact to the reference by
             tests/golden/warp.npz) and a subset against the numpy restatement np_oracle.warp_image
  network    model_oracle.forward (torch CPU float32, the reference's umetrack_model.py:123-242) on the
             reference crops, with its own SimpleConvRNN state in the same memory slots
  landmarks  np_oracle.landmarks_from_hand_pose (perspective_crop.py:40-51, hand_skinning.py:189-209)

Returned: hand-frames compared, mean / max dMPJPE (mm), mean / max |dtheta| (deg, 20 joint angles),
max per-pixel crop difference (grey levels).  Called by tests/test_gpu_parity.py and by bench.py's
`parity` block; nothing of the product imports it.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)

TOL = {"d_mpjpe_mm": 0.1, "d_theta_deg": 0.05, "crop_max_abs_diff": 1}   # north_star


def synth_frames(S, step, seed):
    """uint8 [S,480,4,636]: even slots i.i.d. noise, odd slots a smooth texture that moves with the step."""
    import scenario

    g = torch.Generator().manual_seed(seed * 1000 + step)
    fr = torch.randint(0, 256, (S, 480, 4, 636), dtype=torch.uint8, generator=g)
    for s in range(1, S, 2):
        fr[s] = torch.from_numpy(scenario.synth_frame("smooth", 17 * s + step))
    return fr


def timed_path_parity(S=256, T=8, precision="fp16", map_mode="fp32", ingest="pull", seed=0, numpy_warp_every=97,
                      state_dict=None, seq_offset=0):
    import model_oracle as MO
    import np_oracle as O
    import scenario
    from absolutetrack_b200 import synthetic
    from absolutetrack_b200.lib.models.model_loader import make_random_state_dict, model_from_state_dict
    from absolutetrack_b200.lib.tracker.batch_tracker import BatchedHandTracker
    from absolutetrack_b200.lib.tracker.tracker import crop_resample_cuda

    sd = state_dict if state_dict is not None else make_random_state_dict(0)
    hm = synthetic.hand_model()
    hmn = scenario.hand_model()
    tracker = BatchedHandTracker(model_from_state_dict(sd, precision=precision).cuda(), hm, max_slots=2 * S,
                                 map_mode=map_mode, ingest=ingest)
    seq_ids = np.arange(S) + seq_offset
    skel = (torch.from_numpy(hmn["joint_rotation_axes"]).float(), torch.from_numpy(hmn["joint_rest_positions"] * 0.001).float())
    st = MO.TemporalState()
    d_lm, d_th = [], []
    crop_max, crop_diff_px, crop_px, np_checked = 0, 0, 0, 0
    G = scenario.geometry()
    for t in range(T):
        hb = synthetic.make_batch(seq_ids, t, frames=synth_frames(S, t, seed), first_step=(t == 0)).pin()
        out = tracker.step(hb)
        torch.cuda.synchronize()
        res = {k: v.clone() for k, v in out.items()}
        # ---- crops: timed mode vs the float64-map mode (== tracker._warp_image, tests/golden/warp.npz)
        dev_frames = hb.frames.cuda()
        args = (dev_frames, synthetic.H * synthetic.V * synthetic.W, synthetic.W, synthetic.V * synthetic.W,
                hb.crop_params.cuda(), (96, 96), torch.uint8)
        timed = crop_resample_cuda(*args, map_mode)
        exact = crop_resample_cuda(*args, "fp64")
        diff = (timed.to(torch.int16) - exact.to(torch.int16)).abs()
        crop_max = max(crop_max, int(diff.max().item()))
        crop_diff_px += int((diff != 0).sum().item())
        crop_px += diff.numel()
        exact_np = exact.cpu().numpy()
        # ... and a subset of the float64-mode crops against the numpy restatement of _warp_image
        recs = hb.crop_params.numpy().view(synthetic.CROP_PARAMS_DTYPE).reshape(-1)
        ks = synthetic.geometry_index(seq_ids, t)
        for ci in range(t % numpy_warp_every, len(recs), numpy_warp_every):
            slot, cam = int(recs[ci]["frame"]), int(recs[ci]["view"])
            k = int(ks[slot])
            hs = [(h, s) for h in range(2) for s in range(2) if int(G["crop_cam"][k, h, s]) == cam
                  and np.allclose(G["crop_f"][k, h, s], recs[ci]["f_dst"])]
            h, s = hs[0]
            ref = O.warp_image(scenario.src_camera(k, cam), scenario.dst_camera(k, h, s),
                               np.ascontiguousarray(hb.frames[slot, :, cam, :].numpy()))
            crop_max = max(crop_max, int(np.abs(ref.astype(np.int32) - exact_np[ci].astype(np.int32)).max()))
            np_checked += 1
        # ---- network + FK oracle on the reference crops
        ref = MO.forward(sd, st, torch.from_numpy(exact_np.astype(np.float32) / 255.0), hb.intrinsics, hb.extrinsics,
                         hb.sample_range, hb.memory_idx, hb.use_memory, hb.hand_idx, skel)
        w = ref["wrist_xfs"].numpy().astype(np.float64).copy()
        w[:, :3, 3] *= 1000.0            # tracker.py:379 (metres -> mm)
        lm_ref = O.landmarks_from_hand_pose(hmn, ref["joint_angles"].numpy(), w, hb.hand_idx.numpy())
        d_lm.append(np.linalg.norm(res["landmarks"].numpy().astype(np.float64) - lm_ref, axis=-1).mean(-1))
        d_th.append(np.abs(np.degrees(res["joint_angles"].numpy()[:, :20] - ref["joint_angles"].numpy()[:, :20])).mean(-1))
    tracker.check()
    d_lm, d_th = np.concatenate(d_lm), np.concatenate(d_th)
    return {
        "path": f"BatchedHandTracker.step(precision={precision}, map_mode={map_mode}, ingest={ingest}), S={S} sequences x T={T} steps, "
                "temporal memory carried, even slots noise frames / odd slots smooth frames",
        "against": "oracle/ (numpy float64 warp + cv2-exact remap, torch-CPU float32 network, numpy float32 FK) on the same inputs",
        "n_hand_frames": int(d_lm.size), "n_crops": int(crop_px // (96 * 96)),
        "d_mpjpe_mm": float(d_lm.mean()), "d_mpjpe_mm_max": float(d_lm.max()),
        "d_theta_deg": float(d_th.mean()), "d_theta_deg_max": float(d_th.max()),
        "crop_max_abs_diff": int(crop_max), "crop_pixels_differing": int(crop_diff_px), "crop_pixels": int(crop_px),
        "crops_checked_against_numpy_warp": int(np_checked),
        "tolerance": TOL,
        "pass": bool(d_lm.mean() <= TOL["d_mpjpe_mm"] and d_th.mean() <= TOL["d_theta_deg"] and crop_max <= TOL["crop_max_abs_diff"]),
    }

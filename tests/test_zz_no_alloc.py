"""at_gen_crop_cameras_static (the allocation-free variant of at_gen_crop_cameras: scratch in a static device
array of the library) against the existing entry point, through the C-ABI, on two calls in a row.  Collected
last on purpose; passes alone in a fresh process and after the other tests."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

FIELDS = ("crop_params", "intrinsics", "extrinsics", "crop_camera_to_world", "crop_camera_index", "sample_range",
          "hand_idx", "memory_idx", "use_memory", "pose_index")


def _call(lib, rig, opts, pdev, valid, P, F, static):
    from absolutetrack_b200 import _cabi
    from absolutetrack_b200.lib.tracker.crop_cameras import alloc_crop_inputs

    dev = rig.device
    out = alloc_crop_inputs(P, opts.max_view_num, dev)
    for f in FIELDS:  # capacity-sized buffers: rows beyond (n, c) are never written - make them comparable
        getattr(out, f).zero_()
    history = torch.zeros(2 * F, dtype=torch.uint8, device=dev)
    ja, wx, pf, ph, c2w = pdev
    cin = _cabi.AtCropCameraInputs(ja.data_ptr(), wx.data_ptr(), pf.data_ptr(), ph.data_ptr(), valid.data_ptr(),
                                   _cabi.ptr(rig.neutral), c2w.data_ptr(), rig.cameras.data_ptr(), P, F, rig.n_cams, 0)
    cout = _cabi.AtCropCameraOutputs(*[getattr(out, f).data_ptr() for f in FIELDS], history.data_ptr(), out.counts.data_ptr())
    copts = opts.to_c()
    stream = _cabi.stream_ptr(dev)
    if static:
        rc = lib.at_gen_crop_cameras_static(ctypes.byref(cin), ctypes.byref(rig.hm), ctypes.byref(copts), ctypes.byref(cout), stream)
    else:
        need = lib.at_crop_cameras_workspace_bytes(P, opts.max_view_num)
        ws = torch.empty(int(need), dtype=torch.uint8, device=dev)
        rc = lib.at_gen_crop_cameras(ctypes.byref(cin), ctypes.byref(rig.hm), ctypes.byref(copts), ctypes.byref(cout),
                                     ws.data_ptr(), ws.numel(), stream)
    torch.cuda.synchronize()
    return rc, out, history


def test_static_scratch_variant_matches_the_workspace_variant(cuda_lib):
    from absolutetrack_b200 import _cabi, synthetic
    from absolutetrack_b200.lib.tracker.crop_cameras import CropCameraOpts, DeviceRig

    lib = cuda_lib
    dev = torch.device("cuda", torch.cuda.current_device())
    cams, angles = synthetic.rig_cameras()
    rig = DeviceRig(cams, angles, synthetic.hand_model(), dev)
    # two calls in a row with DIFFERENT sizes and options: the second one must not see the first one's scratch
    for S, step, opts in ((96, 0, CropCameraOpts()), (40, 3, CropCameraOpts(max_view_num=4, sort_camera_index=False, num_crop_points=21))):
        pb = synthetic.make_pose_batch(np.arange(S) + 5 * step, step)
        P = int(pb.joint_angles.shape[0])
        pdev = [getattr(pb, f).to(dev).contiguous() for f in ("joint_angles", "wrist_xforms", "pose_frame", "pose_hand", "camera_to_world")]
        valid = pb.pose_valid.to(dev).contiguous()
        valid[3] = 0                                                   # one pose dropped by confidence
        rc_a, a, hist_a = _call(lib, rig, opts, pdev, valid, P, S, static=False)
        rc_b, b, hist_b = _call(lib, rig, opts, pdev, valid, P, S, static=True)
        _cabi.check(rc_a, "at_gen_crop_cameras")
        _cabi.check(rc_b, "at_gen_crop_cameras_static")
        assert torch.equal(a.counts, b.counts) and int(a.counts[0]) > 0 and int(a.counts[1]) >= int(a.counts[0])
        assert torch.equal(hist_a, hist_b)
        for f in FIELDS:   # same kernels, same inputs: bit-identical (the tolerance of the crop-camera tests is not needed)
            assert torch.equal(getattr(a, f), getattr(b, f)), f


def test_static_scratch_variant_rejects_sizes_beyond_its_limit(cuda_lib):
    from absolutetrack_b200 import _cabi, synthetic
    from absolutetrack_b200.lib.tracker.crop_cameras import CropCameraOpts, DeviceRig

    lib = cuda_lib
    dev = torch.device("cuda", torch.cuda.current_device())
    cams, angles = synthetic.rig_cameras()
    rig = DeviceRig(cams, angles, synthetic.hand_model(), dev)
    S = _cabi.CROP_CAMERAS_STATIC_MAX_POSES // 2 + 1                    # 2 poses per frame: one frame too many
    pb = synthetic.make_pose_batch(np.arange(S), 0)
    P = int(pb.joint_angles.shape[0])
    pdev = [getattr(pb, f).to(dev).contiguous() for f in ("joint_angles", "wrist_xforms", "pose_frame", "pose_hand", "camera_to_world")]
    rc, _, _ = _call(lib, rig, CropCameraOpts(), pdev, pb.pose_valid.to(dev).contiguous(), P, S, static=True)
    assert rc == -3                                                      # AT_ERR_WORKSPACE_TOO_SMALL
    assert b"AT_CROP_CAMERAS_STATIC_MAX_POSES" in lib.at_last_error_string()
    rc, out, _ = _call(lib, rig, CropCameraOpts(), pdev, pb.pose_valid.to(dev).contiguous(), P, S, static=False)
    assert rc == 0 and int(out.counts[0]) > 0                            # the workspace variant takes the same call

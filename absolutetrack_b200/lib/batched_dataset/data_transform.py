"""Perspective crops of config-3 sequences on the GPU (mirror of the arithmetic in the reference's
lib/batched_dataset/data_transform.py:57-277).

The reference's DataLoader workers build, per frame and view, a crop camera that looks at the hand
(`_gen_crop_matrices`, :145-209, numpy + one Python call chain per view) and then resample all views
of the sequence with a pinhole->pinhole homography (`_resample_images_batched`, :78-144).  Here both
are one launch each for the whole sequence (C-ABI at_gen_crop_matrices_f32 and
at_resample_homography_f32) and everything stays on the device.  The msgpack / .torch.idx readers in
front of this are host IO and not part of this package.  There is no CPU path.
"""
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Tuple

import numpy as np
import torch

from .. import _cabi
from ..common.hand import HandModel, mirrored_hand_model
from ..data_utils import bundles
from .sample import RawSample, parse_raw_buffers


def _dev():
    if not torch.cuda.is_available():
        raise _cabi.AbsoluteTrackError("data_transform needs a CUDA device (no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


def _f32(a, dev):
    return torch.as_tensor(a).to(device=dev, dtype=torch.float32).contiguous()


def gen_crop_matrices_cuda(orig_extrinsics, orig_intrinsics, crop_points, mirror_image, crop_size: Tuple[int, int],
                           focal_multiplier: float = 0.95):
    """`_gen_crop_matrices` (:145-209) for a whole sequence.

    orig_extrinsics [F,V,4,4] (world->eye), orig_intrinsics [F,V,3,3], crop_points [F,P,3],
    mirror_image: bool or [F] bools.  Returns CUDA float32 (extrinsics_xf [F,V,4,4], new_intrinsics
    [F,V,3,3], resample_xf [F,V,4,4]).  Raises ValueError("Unable to create crop camera") like
    crop.py:25-26 when any view fails the reference's sanity checks (one 4-byte device read)."""
    lib = _cabi.lib()
    dev = _dev()
    ext, intr, pts = _f32(orig_extrinsics, dev), _f32(orig_intrinsics, dev), _f32(crop_points, dev)
    if ext.dim() != 4 or tuple(ext.shape[2:]) != (4, 4) or tuple(intr.shape) != tuple(ext.shape[:2]) + (3, 3):
        raise ValueError(f"expected [F,V,4,4] extrinsics and [F,V,3,3] intrinsics, got {tuple(ext.shape)}, {tuple(intr.shape)}")
    F, V = int(ext.shape[0]), int(ext.shape[1])
    if pts.dim() != 3 or int(pts.shape[0]) != F or int(pts.shape[2]) != 3 or int(pts.shape[1]) < 1:
        raise ValueError(f"expected [F,P,3] crop points, got {tuple(pts.shape)}")
    if isinstance(mirror_image, (bool, int)):
        mir = torch.full((F,), int(bool(mirror_image)), dtype=torch.uint8, device=dev)
    else:
        mir = torch.as_tensor(mirror_image).to(device=dev).ne(0).to(torch.uint8).reshape(F).contiguous()
    out_e = torch.empty((F, V, 4, 4), dtype=torch.float32, device=dev)
    out_k = torch.empty((F, V, 3, 3), dtype=torch.float32, device=dev)
    out_x = torch.empty((F, V, 4, 4), dtype=torch.float32, device=dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = lib.at_gen_crop_matrices_f32(ext.data_ptr(), intr.data_ptr(), pts.data_ptr(), mir.data_ptr(), F, V,
                                          int(pts.shape[1]), int(crop_size[0]), int(crop_size[1]), float(focal_multiplier),
                                          out_e.data_ptr(), out_k.data_ptr(), out_x.data_ptr(), err.data_ptr(),
                                          _cabi.stream_ptr(dev))
    _cabi.check(rc, "at_gen_crop_matrices_f32")
    n_err = int(err.item())
    if n_err:
        raise ValueError("Unable to create crop camera", n_err)
    return out_e, out_k, out_x


def resample_images_batched_cuda(images_orig, resample_xfs, out_size: Tuple[int, int], divisor: float = 1.0) -> torch.Tensor:
    """`_resample_images_batched` (:78-144): images_orig [N,H,W] grey levels, resample_xfs [N,4,4]
    -> float32 CUDA [N, out_size[1], out_size[0]] (rows = height), divided by `divisor`."""
    lib = _cabi.lib()
    dev = _dev()
    img, xf = _f32(images_orig, dev), _f32(resample_xfs, dev)
    if img.dim() != 3 or tuple(xf.shape) != (int(img.shape[0]), 4, 4):
        raise ValueError(f"expected [N,H,W] images and [N,4,4] matrices, got {tuple(img.shape)}, {tuple(xf.shape)}")
    n, H, W = (int(v) for v in img.shape)
    ow, oh = int(out_size[0]), int(out_size[1])
    out = torch.empty((n, oh, ow), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = lib.at_resample_homography_f32(img.data_ptr(), n, H, W, xf.data_ptr(), oh, ow, out.data_ptr(), float(divisor),
                                            _cabi.stream_ptr(dev))
    _cabi.check(rc, "at_resample_homography_f32")
    return out


def _perspective_crop_images(orig_images, orig_extrinsics, orig_intrinsics, crop_points, hand_idx: int,
                             crop_size: Tuple[int, int]) -> List[torch.Tensor]:
    """The reference's `_perspective_crop_images` (:212-277), same arguments and return order:
    [cropped_left_images [F,V,h,w] in 0..1, extrinsics_xf [F,V,4,4], new_intrinsics [F,V,3,3]] - as
    CUDA float32 tensors, ready for the model (data_transform.py:348-356 wraps them with to_th)."""
    dev = _dev()
    img = _f32(orig_images, dev)
    if img.dim() != 4:
        raise ValueError(f"expected [F,V,H,W] images, got {tuple(img.shape)}")
    F, V = int(img.shape[0]), int(img.shape[1])
    ext, intr, xf = gen_crop_matrices_cuda(orig_extrinsics, orig_intrinsics, crop_points, int(hand_idx) == 1, crop_size)
    if tuple(ext.shape[:2]) != (F, V):
        raise ValueError("images and camera arrays disagree on [frames, views]")
    crops = resample_images_batched_cuda(img.reshape(F * V, int(img.shape[2]), int(img.shape[3])), xf.reshape(F * V, 4, 4),
                                         crop_size, 255.0)  # :275 normalises to 0..1
    return [crops.reshape(F, V, int(crop_size[1]), int(crop_size[0])), ext, intr]


# --------------------------------------------------------------------------------------------
# sequence -> model input / targets (data_transform.py:23-54,280-397)
# --------------------------------------------------------------------------------------------


@dataclass
class PoseData:
    joint_angles: torch.Tensor
    wrist_xfs: torch.Tensor
    left_hand_model: HandModel


@dataclass
class ModelInput:
    orig_pose_data: PoseData
    s_solved_pose_data: PoseData
    left_images: torch.Tensor
    intrinsics: torch.Tensor
    extrinsics_xf: torch.Tensor
    hand_idx: torch.Tensor


@dataclass
class PerBranchOutput:
    joint_angles: torch.Tensor
    wrist_xfs: torch.Tensor
    skel_scales: Optional[torch.Tensor] = None
    pinch_prediction: Optional[torch.Tensor] = None


@dataclass
class ModelTarget:
    gt_skel_targets: PerBranchOutput
    preds_targets: PerBranchOutput
    intrinsics: Optional[torch.Tensor] = None
    extrinsics_xf: Optional[torch.Tensor] = None


def prepare_inputs_targets(sample: RawSample, crop_size: Tuple[int, int]) -> Tuple[ModelInput, ModelTarget]:
    """`prepare_inputs_targets` of the reference (:280-385): mm -> m, left-hand skeletons, the
    perspective crops of the whole sequence (two launches on the GPU instead of a Python chain per
    view), and the targets of both branches.  Images, intrinsics and extrinsics come back as CUDA
    tensors; the small pose / skeleton tensors stay where the labels were (host)."""

    def to_th(a: np.ndarray) -> torch.Tensor:
        return torch.from_numpy(np.asarray(a)).float()

    sample.scaled(0.001)
    seq_length = sample.images.shape[0]
    is_right = to_th(sample.hand) == 1

    def left_model(hm: HandModel) -> HandModel:
        rep = bundles.map_fields(lambda t: t.unsqueeze(0).expand(seq_length, *t.shape), hm, only_type=torch.Tensor)
        return mirrored_hand_model(rep, is_right)

    s_solved = PoseData(wrist_xfs=to_th(sample.solved_wrist_xfs), joint_angles=to_th(sample.solved_joint_angles),
                        left_hand_model=left_model(sample.generic_hand_model))
    orig = PoseData(wrist_xfs=to_th(sample.wrist), joint_angles=to_th(sample.joint_angles),
                    left_hand_model=left_model(sample.hand_model))
    left_images, extrinsics_xf, intrinsics = _perspective_crop_images(
        sample.images, sample.extrinsics, sample.intrinsics, sample.enclosing_points, int(sample.hand[0]), crop_size)
    model_input = ModelInput(orig_pose_data=orig, s_solved_pose_data=s_solved, left_images=left_images,
                             intrinsics=intrinsics, extrinsics_xf=extrinsics_xf, hand_idx=to_th(sample.hand))
    pinch = to_th(sample.pinch)
    target = ModelTarget(
        gt_skel_targets=PerBranchOutput(orig.joint_angles, orig.wrist_xfs, orig.left_hand_model.hand_scale, pinch),
        preds_targets=PerBranchOutput(s_solved.joint_angles, s_solved.wrist_xfs, s_solved.left_hand_model.hand_scale, pinch),
        intrinsics=intrinsics, extrinsics_xf=extrinsics_xf,
    )
    return model_input, target


def preprocess(data: Dict[str, Any], crop_size: Tuple[int, int]) -> Tuple[ModelInput, ModelTarget]:
    """One decoded sequence {'mono': uint8 [T,V,H,W], 'labels': msgpack dict} -> (ModelInput,
    ModelTarget) (data_transform.py:388-397)."""
    return prepare_inputs_targets(parse_raw_buffers(**data), crop_size)


def iter_preprocessed_sequences(mono_idx_path: str, labels_idx_path: str, crop_size: Tuple[int, int] = (96, 96),
                                rank: int = 0, world: int = 1):
    """The sequences of one `.torch.idx/.torch.bin` pair of files (frames + labels), preprocessed on
    this rank's GPU: what the reference builds from Sampler + AsyncToIterableDataset + map_dataset(
    preprocess) (run_inference_torch_data.py:166-177), with the Sampler's rank striding
    (async_dataset.py:558-559 = distributed.shard_indices) and without worker processes - the
    per-sequence work is two launches.  Yields (sequence index, ModelInput, ModelTarget)."""
    from absolutetrack_b200.distributed import shard_indices
    from ..data_utils.idxbinfile import TorchIdx

    mono_idx, labels_idx = TorchIdx(mono_idx_path), TorchIdx(labels_idx_path)
    if len(mono_idx) != len(labels_idx):
        raise ValueError(f"{mono_idx_path} holds {len(mono_idx)} sequences, {labels_idx_path} {len(labels_idx)}")
    with open(mono_idx.bin_path, "rb") as fm, open(labels_idx.bin_path, "rb") as fl:
        for i in shard_indices(len(mono_idx), rank, world):
            lo, hi = mono_idx.byte_offset(i), mono_idx.byte_offset(i + 1)
            fm.seek(lo)
            mono = np.array(mono_idx.view_frame(i, fm.read(hi - lo)))   # writable copy: preprocess owns it
            lo, hi = labels_idx.byte_offset(i), labels_idx.byte_offset(i + 1)
            fl.seek(lo)
            labels = labels_idx.view_frame(i, fl.read(hi - lo))
            yield (i,) + preprocess({"mono": mono, "labels": labels}, crop_size)

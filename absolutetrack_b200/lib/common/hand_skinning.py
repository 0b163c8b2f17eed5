"""FK + landmark skinning on the GPU (mirror of the reference's lib/common/hand_skinning.py API).

`skin_landmarks` keeps the reference signature (hand_skinning.py:189-209) and leading-dims
behaviour ([..., 22] angles, [..., 4, 4] wrists, model with matching or no leading dims) and
runs csrc/fk_skin.cu through the C-ABI entry point at_fk_skin_f32.  float32, like the reference.
"""
import torch

from .. import _cabi
from .hand import HandModel


def _device_for(*tensors):
    for t in tensors:
        if isinstance(t, torch.Tensor) and t.is_cuda:
            return t.device
    return torch.device("cuda", torch.cuda.current_device())


def skin_landmarks_cuda(hand_model: HandModel, joint_angles, wrist_transforms, hand_idx=None, out=None,
                        translation_scale: float = 1.0, wrist_out=None):
    """Device-resident entry: all inputs are moved to the GPU if they are not there already;
    returns a CUDA tensor [..., 21, 3].  `hand_idx` (optional, [...]) mirrors right hands
    (perspective_crop.py:46-48) inside the kernel.  `translation_scale` multiplies the wrist
    translations first (tracker.py:379, metres -> mm) and `wrist_out` (CUDA float32 [n,4,4],
    contiguous, not the input) receives the scaled transforms - both inside the same launch."""
    lib = _cabi.lib()
    lead = tuple(joint_angles.shape[:-1])
    n = 1
    for s in lead:
        n *= int(s)
    dev = _device_for(joint_angles, wrist_transforms)
    f32 = dict(device=dev, dtype=torch.float32)
    ja = joint_angles.to(**f32).reshape(n, 22).contiguous()
    wx = wrist_transforms.to(**f32).reshape(n, 4, 4).contiguous()
    hm_lead = tuple(hand_model.joint_rest_positions.shape[:-2])
    per_sample = len(hm_lead) > 0
    if per_sample and hm_lead != lead:
        raise AssertionError(f"Leading dimensions do not match, got {lead} and {hm_lead}")
    K = int(hand_model.landmark_rest_bone_indices.shape[-1])
    rows = n if per_sample else 1
    axes = hand_model.joint_rotation_axes.to(**f32).reshape(rows, 22, 3).contiguous()
    rest = hand_model.joint_rest_positions.to(**f32).reshape(rows, 22, 3).contiguous()
    lms = hand_model.landmark_rest_positions.to(**f32).reshape(rows, 21, 3).contiguous()
    bw = hand_model.landmark_rest_bone_weights.to(**f32).reshape(rows, 21, K).contiguous()
    bi = hand_model.landmark_rest_bone_indices.to(device=dev, dtype=torch.int32).reshape(rows, 21, K).contiguous()
    hi = None
    if hand_idx is not None:
        hi = torch.as_tensor(hand_idx).to(device=dev, dtype=torch.int64).reshape(n).contiguous()
    if out is None:
        out = torch.empty((n, 21, 3), **f32)
    hm = _cabi.AtHandModel(
        axes.data_ptr(), rest.data_ptr(), lms.data_ptr(), bw.data_ptr(), bi.data_ptr(), K, 1 if per_sample else 0
    )
    if wrist_out is not None:
        if not (wrist_out.is_cuda and wrist_out.dtype == torch.float32 and wrist_out.is_contiguous()
                and wrist_out.numel() == n * 16 and wrist_out.data_ptr() != wx.data_ptr()):
            raise _cabi.AbsoluteTrackError("wrist_out: contiguous CUDA float32 [n,4,4], distinct from the input")
    with torch.cuda.device(dev):
        rc = lib.at_fk_skin_scaled_f32(
            ja.data_ptr(), wx.data_ptr(), _cabi.ptr(hi), hm, n, float(translation_scale), _cabi.ptr(wrist_out),
            out.data_ptr(), _cabi.stream_ptr(dev)
        )
    _cabi.check(rc, "at_fk_skin_scaled_f32")
    return out.reshape(lead + (21, 3))


def skin_landmarks(hand_model: HandModel, joint_angles: torch.Tensor, wrist_transforms: torch.Tensor) -> torch.Tensor:
    """Reference-compatible call: the result lives on the device of `joint_angles`
    (CPU in, CPU out - the computation is still done by the CUDA kernel)."""
    src_dev = joint_angles.device
    out = skin_landmarks_cuda(hand_model, joint_angles, wrist_transforms)
    return out.to(src_dev)

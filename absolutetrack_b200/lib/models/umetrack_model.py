"""UmeTrackModel on the B200 (mirror of the reference's lib/models/umetrack_model.py API).

Same dataclasses (InputFrameData / InputFrameDesc / InputSkeletonData, :21-74) and the same
public methods (getInputImageSizes :119-121, regress_pose_use_skeleton :188-219,
regress_pose_pred_skel_scale :221-242, plus nn.Module .eval()/.to()).  The arithmetic runs in
libabsolutetrack_b200.so: at_encoder_forward (per-view CNN) and at_tail_forward (FTL, fusion,
SimpleConvRNN, skeleton encoder, pose head, decoders, wrist->world).  No CPU path.
"""
from dataclasses import dataclass
from typing import Optional, Tuple

import torch
import torch.nn as nn

from .. import _cabi
from .model_opts import ModelOpts
from .regressor import RegressorOutput


@dataclass
class InputFrameData:
    """left_images [n_images,h,w], intrinsics [n_images,3,3], extrinsics_xf [n_images,4,4]."""

    left_images: torch.Tensor
    intrinsics: torch.Tensor
    extrinsics_xf: torch.Tensor


@dataclass
class InputFrameDesc:
    """sample_range [bs,2], memory_idx [bs], use_memory [bs] bool, hand_idx [bs].  memory_idx values must be
    distinct within a batch (samples sharing a temporal-memory slot race on it; the reference is last-write-wins)."""

    sample_range: torch.Tensor
    memory_idx: torch.Tensor
    use_memory: torch.Tensor
    hand_idx: torch.Tensor


@dataclass
class InputSkeletonData:
    """joint_rotation_axes / joint_rest_positions: [22,3] (one profile) or [bs,22,3]."""

    joint_rotation_axes: torch.Tensor
    joint_rest_positions: torch.Tensor


class _Workspace:
    """Grow-only device scratch buffer."""

    def __init__(self):
        self.buf = None

    def get(self, nbytes: int, device) -> torch.Tensor:
        if self.buf is None or self.buf.numel() < nbytes or self.buf.device != device:
            self.buf = torch.empty(int(nbytes), dtype=torch.uint8, device=device)
        return self.buf


class UmeTrackModel(nn.Module):
    """Holds the BN-folded weight blob and the SimpleConvRNN state (temporal.py:41-42)."""

    PRECISIONS = {"fp32": _cabi.AT_PREC_FP32, "bf16": _cabi.AT_PREC_BF16, "fp16": _cabi.AT_PREC_FP16}

    def __init__(self, blob: torch.Tensor, model_opts: Optional[ModelOpts] = None, precision: str = "fp32"):
        super().__init__()
        lib = _cabi.lib()
        if blob.dtype != torch.float32 or blob.numel() != lib.at_model_blob_floats():
            raise ValueError("weight blob has the wrong size/dtype (use model_loader.pack_state_dict)")
        self.register_buffer("_blob", blob.contiguous().clone(), persistent=False)
        self._opts = model_opts or ModelOpts()
        self._input_size: Tuple[int, int] = (96, 96)
        self.precision = precision
        self._bf16_blob: Optional[torch.Tensor] = None
        self._bf16_prec = -1
        self.tail_mode = "f16x3"  # or "tf32x3": which fp32-grade tensor-core kernel the tail convs use
        self._tail_f16_blob: Optional[torch.Tensor] = None
        # recurrent state (plain attributes, like the reference: not buffers, not moved by .to())
        self._mem_features = torch.empty(0)      # [cap, 6, 6, 18]  (NHWC)
        self._prev_extrinsics = torch.empty(0)   # [cap, 4, 4]
        self._ws_enc = _Workspace()
        self._enc_prepared = None  # ((workspace ptr, precision), chunk size the workspace is prepared for)
        self._ws_tail = _Workspace()
        self._status: Optional[torch.Tensor] = None
        self.memory_capacity_hint: Optional[int] = None  # callers that know max(memory_idx)+1 avoid a sync
        self.release_stream = None  # torch.cuda.Stream made to wait until each encoder call is under way (at_encoder_forward_release)
        self.front_stream = None    # torch.cuda.Stream the crops of the NEXT encode() call were produced on (at_encoder_forward_front)
        self.strict = True  # read the device status flag after every call (one sync, like the reference's .cpu())
        self.eval()

    # ------------------------------------------------------------------ reference API
    def getInputImageSizes(self) -> Tuple[int, int]:
        return self._input_size

    def regress_pose_use_skeleton(
        self, frame_data: InputFrameData, frame_desc: InputFrameDesc, skel_data: InputSkeletonData
    ) -> RegressorOutput:
        return self._forward(frame_data, frame_desc, skel_data, _cabi.AT_HEAD_KNOWN_SKELETON)

    def regress_pose_pred_skel_scale(self, frame_data: InputFrameData, frame_desc: InputFrameDesc) -> RegressorOutput:
        n = frame_desc.sample_range.shape[0]
        # reference :224-229 asserts every sample is multi-view; shape-only check keeps it sync-free
        assert frame_data.left_images.shape[0] == 2 * n, "Unsupported: found single-view samples when calibration scale"
        return self._forward(frame_data, frame_desc, None, _cabi.AT_HEAD_PREDICT_SCALE)

    # ------------------------------------------------------------------ internals
    def reset_temporal_state(self) -> None:
        self._mem_features = torch.empty(0)
        self._prev_extrinsics = torch.empty(0)

    @property
    def mem_features_nchw(self) -> torch.Tensor:
        """Temporal memory in the reference's layout [cap,18,6,6] (temporal.py:41)."""
        return self._mem_features.permute(0, 3, 1, 2) if self._mem_features.numel() else self._mem_features

    def _ensure_state(self, required: int, device) -> None:
        cur = self._mem_features.shape[0] if self._mem_features.numel() else 0
        if cur >= required and self._mem_features.device == device:
            return
        required = max(cur, required)  # a pure device move keeps every slot (the state may have grown before)
        mem = torch.zeros(required, 6, 6, 18, dtype=torch.float32, device=device)
        ext = torch.zeros(required, 4, 4, dtype=torch.float32, device=device)
        if cur:  # grow, keeping old entries (temporal.py:105-125)
            mem[:cur] = self._mem_features.to(device)
            ext[:cur] = self._prev_extrinsics.to(device)
        self._mem_features, self._prev_extrinsics = mem, ext

    def _precision_code(self) -> int:
        try:
            return self.PRECISIONS[self.precision]
        except KeyError:
            raise ValueError(f"precision must be one of {sorted(self.PRECISIONS)}") from None

    def _bf16_weights(self, device):
        """Prepacked 16-bit UMMA operand blob of the tcgen05 encoder (bf16 or fp16)."""
        prec = self._precision_code()
        if self._bf16_blob is None or self._bf16_blob.device != device or self._bf16_prec != prec:
            lib = _cabi.lib()
            buf = torch.zeros(lib.at_model_bf16_blob_bytes(), dtype=torch.uint8, device=device)
            with torch.cuda.device(device):
                _cabi.check(
                    lib.at_model_prepack_tc(self._blob.data_ptr(), buf.data_ptr(), prec, _cabi.stream_ptr(device)),
                    "at_model_prepack_tc",
                )
            self._bf16_blob, self._bf16_prec = buf, prec
        return self._bf16_blob

    def _tail_f16_weights(self, device):
        """Pre-split fp16 hi/lo operand blob of the tail convolutions (at_tail_prepack_f16)."""
        if self.tail_mode == "tf32x3":
            return None
        if self.tail_mode != "f16x3":
            raise ValueError("tail_mode must be 'f16x3' or 'tf32x3'")
        if self._tail_f16_blob is None or self._tail_f16_blob.device != device:
            lib = _cabi.lib()
            buf = torch.zeros(lib.at_tail_f16_blob_bytes(), dtype=torch.uint8, device=device)
            with torch.cuda.device(device):
                _cabi.check(lib.at_tail_prepack_f16(self._blob.data_ptr(), buf.data_ptr(), _cabi.stream_ptr(device)),
                            "at_tail_prepack_f16")
            self._tail_f16_blob = buf
        return self._tail_f16_blob

    def encode(self, crops: torch.Tensor) -> torch.Tensor:
        """Per-view encoder: crops [c,96,96] (uint8 grey levels or float32 in [0,1]) on the GPU ->
        features [c,6,6,72] float32 (the reference's [c,72,6,6], channels last)."""
        lib = _cabi.lib()
        dev = self._blob.device
        if dev.type != "cuda":
            raise _cabi.AbsoluteTrackError("UmeTrackModel must be on a CUDA device: there is no CPU path")
        if crops.dtype == torch.uint8:
            dt = _cabi.AT_U8
        else:
            crops = crops.float()
            dt = _cabi.AT_F32
        crops = crops.to(dev).contiguous()
        c = crops.shape[0]
        assert tuple(crops.shape[1:]) == self._input_size, f"expected [c,96,96] crops, got {tuple(crops.shape)}"
        feats = torch.empty(c, 6, 6, 72, dtype=torch.float32, device=dev)
        prec = self._precision_code()
        bf = self._bf16_weights(dev) if prec != _cabi.AT_PREC_FP32 else None
        nbytes = lib.at_encoder_workspace_bytes(c, prec)
        ws = self._ws_enc.get(nbytes, dev)
        # the workspace stays prepared (halo cells zeroed) for the chunk size of the previous call
        # as long as it is the same tensor and only the encoder wrote it
        key = (ws.data_ptr(), prec)
        prepared = self._enc_prepared[1] if self._enc_prepared is not None and self._enc_prepared[0] == key else 0
        self._enc_prepared = None
        with torch.cuda.device(dev):
            rel = self.release_stream.cuda_stream if self.release_stream is not None else None
            front = self.front_stream.cuda_stream if self.front_stream is not None else None
            rc = lib.at_encoder_forward_front(
                self._blob.data_ptr(), _cabi.ptr(bf), crops.data_ptr(), dt, c, feats.data_ptr(), ws.data_ptr(),
                ws.numel(), prec, prepared, self._status_word(dev).data_ptr(), front, rel, _cabi.stream_ptr(dev),
            )
        _cabi.check(rc, "at_encoder_forward_front")
        if c > 0:
            chunk = lib.at_encoder_chunk_crops()
            self._enc_prepared = (key, c % chunk if c % chunk else min(c, chunk))
        return feats

    def _forward(self, frame_data, frame_desc, skel_data, head_mode) -> RegressorOutput:
        return self.tail(self.encode(frame_data.left_images), frame_data, frame_desc, skel_data, head_mode)

    def tail(self, feats: torch.Tensor, frame_data, frame_desc, skel_data, head_mode=_cabi.AT_HEAD_KNOWN_SKELETON) -> RegressorOutput:
        """Everything after the per-view encoder (FTL, fusion, SimpleConvRNN step, skeleton features, pose
        head, decoders, wrist->world) on encoder features [c,6,6,72]; `frame_data.left_images` is not read."""
        lib = _cabi.lib()
        dev = self._blob.device
        c = feats.shape[0]
        f32 = dict(device=dev, dtype=torch.float32)
        i64 = dict(device=dev, dtype=torch.int64)
        intr = frame_data.intrinsics.to(**f32).contiguous()
        extr = frame_data.extrinsics_xf.to(**f32).contiguous()
        srange = frame_desc.sample_range.to(**i64).contiguous()
        midx = frame_desc.memory_idx.to(**i64).contiguous()
        usem = frame_desc.use_memory
        if usem.dtype == torch.bool and usem.device == dev and usem.is_contiguous():
            usem = usem.view(torch.uint8)  # same bytes: no conversion launch
        else:
            usem = usem.to(device=dev, dtype=torch.uint8).contiguous()
        hidx = frame_desc.hand_idx.to(**i64).contiguous()
        n = srange.shape[0]
        if self.memory_capacity_hint is not None:
            required = int(self.memory_capacity_hint)
        else:
            required = int(midx.max().item()) + 1  # the reference syncs here too (temporal.py:100)
        self._ensure_state(max(required, 1), dev)
        axes = rest = None
        n_skel = 0
        if head_mode == _cabi.AT_HEAD_KNOWN_SKELETON:
            axes = skel_data.joint_rotation_axes.to(**f32).reshape(-1, 22, 3).contiguous()
            rest = skel_data.joint_rest_positions.to(**f32).reshape(-1, 22, 3).contiguous()
            n_skel = axes.shape[0]
            if n_skel not in (1, n):
                raise ValueError(f"skeleton batch {n_skel} matches neither 1 nor the {n} samples")
        out_angles = torch.empty(n, 22, **f32)
        out_wrist = torch.empty(n, 4, 4, **f32)
        out_sig = torch.empty(n, 21, **f32)
        out_scale = torch.empty(n, **f32) if head_mode == _cabi.AT_HEAD_PREDICT_SCALE else None
        self._status_word(dev)
        tin = _cabi.AtTailInputs(
            feats.data_ptr(), intr.data_ptr(), extr.data_ptr(), srange.data_ptr(), midx.data_ptr(), usem.data_ptr(),
            hidx.data_ptr(), _cabi.ptr(axes), _cabi.ptr(rest), n, c, n_skel, head_mode,
        )
        st = _cabi.AtTemporalState(
            self._mem_features.data_ptr(), self._prev_extrinsics.data_ptr(), self._mem_features.shape[0]
        )
        tout = _cabi.AtTailOutputs(
            out_angles.data_ptr(), out_wrist.data_ptr(), _cabi.ptr(out_scale), out_sig.data_ptr(),
            self._status.data_ptr(),
        )
        ws = self._ws_tail.get(lib.at_tail_workspace_bytes(n), dev)
        f16 = self._tail_f16_weights(dev)
        with torch.cuda.device(dev):
            rc = lib.at_tail_forward_f16(
                self._blob.data_ptr(), _cabi.ptr(f16), tin, st, tout, ws.data_ptr(), ws.numel(), _cabi.stream_ptr(dev)
            )
        _cabi.check(rc, "at_tail_forward")
        if self.strict:
            self.check_status()
        return RegressorOutput(
            joint_angles=out_angles, wrist_xfs=out_wrist, skel_scales=out_scale, landmark_uncertainty_sigmas=out_sig
        )

    def _status_word(self, dev) -> torch.Tensor:
        if self._status is None or self._status.device != dev:
            self._status = torch.zeros(1, dtype=torch.int32, device=dev)
        return self._status

    def check_status(self) -> None:
        """Raise if any call since the last check saw a bad sample_range / memory_idx, or an fp16
        activation of the encoder beyond the fp16 range (syncs; clears the flags)."""
        if self._status is None:
            return
        st = int(self._status.item())
        if st == 0:
            return
        self._status.zero_()
        msgs = []
        if st & _cabi.AT_STATUS_BAD_SAMPLE_RANGE:
            msgs.append("at_tail_forward: invalid sample_range (views per sample must be 1 or 2, inside the crop "
                        "batch) or memory_idx outside the temporal-state capacity")
        if st & _cabi.AT_STATUS_FP16_OVERFLOW:
            msgs.append("at_encoder_forward: an fp16 activation saturated at +-65504 (or is inf / NaN): these weights "
                        "need precision='bf16' or 'fp32' (AT_PRECISION)")
        raise _cabi.AbsoluteTrackError("; ".join(msgs) or f"device status {st}")

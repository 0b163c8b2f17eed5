"""Batched frame-stream tracking: the throughput-shaped entry of the hot path.

The reference tracks one frame per call (HandTracker.track_frame, lib/tracker/tracker.py:262-289,
called from the per-frame loop of run_eval_known_skeleton.py:68-90).  Independent sequences
(recordings) have no data dependence on each other, so this class advances S independent
sequences by one frame per call: S multi-view frames -> 4S crops -> 2S poses -> 2S x 21
landmarks, with the SimpleConvRNN memory slot of (sequence s, hand h) at s*2+h.  It is the same
arithmetic as S HandTracker instances stepping in lock-step (tracker._make_inputs :315-368,
UmeTrackModel.regress_pose_use_skeleton umetrack_model.py:188-219, _gen_tracking_result
:370-412 and perspective_crop.landmarks_from_hand_pose :40-51), launched once per stage.

Everything runs in libabsolutetrack_b200.so; there is no CPU path.
"""
from dataclasses import dataclass
from typing import Optional

import os

import numpy as np
import torch

from .. import _cabi
from ..common.hand import HandModel
from ..common.hand_skinning import skin_landmarks_cuda
from ..models.umetrack_model import InputFrameData, InputFrameDesc, InputSkeletonData
from .crop_cameras import CropCameraOpts, DeviceRig, alloc_crop_inputs, gen_crop_inputs_cuda
from .tracker import crop_pull_source_boxes_cuda, crop_resample_cuda

M_TO_MM = 1000.0


@dataclass
class FrameBatch:
    """One time step of S sequences.  Tensors may live on the host (pinned for async copies)
    or on the device; `to(device)` moves all of them.

    frames       uint8 [S, H, V, W]  multi-view frames in the recording layout
                 (video_pose_data.py:123-125: a [H, V*W] video frame viewed as [H, V, W])
    crop_params  uint8 [c, 256]      packed AtCropParams (frame = sequence slot, view = camera)
    intrinsics   f32 [c,3,3], extrinsics f32 [c,4,4] (world->crop eye, metres)
    sample_range i64 [n,2], hand_idx i64 [n], memory_idx i64 [n], use_memory bool [n]
    """

    frames: torch.Tensor
    crop_params: torch.Tensor
    intrinsics: torch.Tensor
    extrinsics: torch.Tensor
    sample_range: torch.Tensor
    hand_idx: torch.Tensor
    memory_idx: torch.Tensor
    use_memory: torch.Tensor

    def to(self, device, non_blocking=True) -> "FrameBatch":
        return FrameBatch(*[getattr(self, f).to(device, non_blocking=non_blocking) for f in self.__dataclass_fields__])

    def pin(self) -> "FrameBatch":
        return FrameBatch(*[getattr(self, f).pin_memory() for f in self.__dataclass_fields__])

    def nbytes(self) -> int:
        return sum(getattr(self, f).numel() * getattr(self, f).element_size() for f in self.__dataclass_fields__)


@dataclass
class PoseBatch:
    """One time step of S sequences BEFORE crop-camera generation: what a recording provides per
    frame (video_pose_data.py:98-150: images, camera_to_world_transforms, and the pose the crop
    is centred on - the labelled pose in the eval scripts, run_eval_known_skeleton.py:74-82).

    frames          uint8 [S,H,V,W]   (pinned host memory for step_poses)
    joint_angles    f32 [P,22], wrist_xforms f32 [P,4,4] (mm; right hands NOT mirrored)
    pose_frame      i32 [P] sequence slot, pose_hand i32 [P], pose_valid uint8 [P] (confidence >= 0.5)
    camera_to_world f64 [S,V,4,4]
    """

    frames: torch.Tensor
    joint_angles: torch.Tensor
    wrist_xforms: torch.Tensor
    pose_frame: torch.Tensor
    pose_hand: torch.Tensor
    pose_valid: torch.Tensor
    camera_to_world: torch.Tensor

    def pin(self) -> "PoseBatch":
        return PoseBatch(*[getattr(self, f).pin_memory() for f in self.__dataclass_fields__])

    def small_nbytes(self) -> int:
        return sum(getattr(self, f).numel() * getattr(self, f).element_size() for f in self.__dataclass_fields__ if f != "frames")


@dataclass
class BatchResult:
    joint_angles: torch.Tensor   # [n,22]
    wrist_xforms: torch.Tensor   # [n,4,4] wrist->world, millimetres, right hands mirrored
    landmarks: torch.Tensor      # [n,21,3] world, millimetres
    memory_idx: Optional[torch.Tensor] = None  # [n] (sequence slot*2 + hand) - set by step_poses, where n is data dependent


class BatchedHandTracker:
    def __init__(self, model, hand_model_mm: HandModel, max_slots: int, map_mode: str = "fp32",
                 crop_dtype: torch.dtype = torch.uint8, ingest: str = "pull") -> None:
        """ingest: how step() brings HOST frames to the GPU.
        "pull" (needs pinned frames): a small side-stream kernel copies only each crop's source
        bounding box (~37 KB per crop instead of the 305 KB view) from the pinned frames into a
        device frame buffer of the same layout, overlapped with the previous step's kernels; the
        resampler then runs on device memory, in any map mode.
        "zero_copy" (pinned frames, map_mode "fp32"): the resampler itself reads the pinned frames
        in place through shared memory - no side stream, the PCIe time is inside the step.
        "copy": the whole frame tensor is copied on the side stream."""
        if not torch.cuda.is_available():
            raise _cabi.AbsoluteTrackError("BatchedHandTracker needs a CUDA device (no CPU fallback)")
        self._device = torch.device("cuda", torch.cuda.current_device())
        self._model = model.to(self._device)
        self._model.memory_capacity_hint = int(max_slots)
        self._model.strict = False  # status flag is read by check() instead of once per call
        self._map_mode = map_mode
        self._crop_dtype = crop_dtype
        f32 = dict(device=self._device, dtype=torch.float32)
        self._hm = hand_model_mm
        self._skel = InputSkeletonData(
            joint_rotation_axes=hand_model_mm.joint_rotation_axes.to(**f32).contiguous(),
            joint_rest_positions=(hand_model_mm.joint_rest_positions.to(**f32) * 0.001).contiguous(),
        )
        self._hm_dev = HandModel(*[
            (v.to(self._device) if isinstance(v, torch.Tensor) else v) for v in hand_model_mm
        ])
        # the dtypes at_fk_skin_f32 takes, so that skin_landmarks_cuda converts nothing per step
        self._hm_dev = self._hm_dev._replace(
            landmark_rest_bone_indices=self._hm_dev.landmark_rest_bone_indices.to(torch.int32).contiguous())
        # host->device ingest: a copy stream and two device staging slots, so that the H2D copy of
        # step i+1 overlaps the kernels of step i (frames are 1.22 MB each: ingest is the e2e bound)
        if ingest not in ("pull", "zero_copy", "copy"):
            raise ValueError("ingest must be 'pull', 'zero_copy' or 'copy'")
        self._ingest = ingest
        self.pulled_bytes = torch.zeros(1, dtype=torch.int64, device=self._device)  # bytes the resampler pulled from host frames
        self.pull_ctas = 0  # CTAs of the pull kernel (0 = one per SM)
        self.trace = None   # debug: a list -> step() appends (ingest start, ingest end, compute start, compute end) events
        self.encoder_events = None   # measurement: a list -> step_device() appends (start, end) timing events around the encoder
        # high priority: the ingest kernels are tiny and must not queue behind the encoder's CTAs
        # The ingest stream is released from INSIDE the encoder (at_encoder_forward_release: the copy stream waits until
        # layer1.0.conv1 of the current step has been launched and finished).  Released at a kernel boundary (behind the
        # resampler, as in round 1) the pull kernel's one-warp CTAs are placed on whichever SMs are free at that moment
        # and pile up there; a conv grid of 148 persistent CTAs that loses one SM to them runs a second wave (measured
        # on B200: compute at half speed while the pull was in flight, e2e 234k -> 305k hand-frames/s on the same box,
        # profiles/r2_e2e_ingest_release_ab.log).  AT_INGEST_RELEASE=crops restores the old release point (A/B).
        self._copy_stream = torch.cuda.Stream(device=self._device, priority=int(os.environ.get("AT_INGEST_PRIORITY", "-1")))
        self._release_in_encoder = os.environ.get("AT_INGEST_RELEASE", "encoder") == "encoder"
        if self._release_in_encoder:
            self._model.release_stream = self._copy_stream
        self._stage = [None, None]
        self._stage_free = [None, None]
        self._slot = 0
        self._crops_done = None
        # Front stream: the resampler and the encoder's stem of step i+1 (ordinary grids, 8 % of a step) run there, NEXT TO
        # the tail of step i on the main stream (18 small latency-bound launches, 14 % of a step) instead of behind it;
        # the main stream joins before layer 1 (at_encoder_forward_front).  Opt-in (AT_FRONT_OVERLAP=1): measured on B200 the
        # overlap happens but buys nothing - the tail stretches from 0.34 to 0.54 ms while the resampler and the stem share
        # the SMs with it (profiles/r2_front_stream_overlap_trace.txt), so the default keeps everything on one stream.
        self._front = torch.cuda.Stream(device=self._device) if os.environ.get("AT_FRONT_OVERLAP", "0") == "1" else None
        self._enc_done = None   # main-stream event behind the previous step's encoder (its level-0 planes and crops are free)

    @property
    def device(self):
        return self._device

    def step_device(self, b: FrameBatch, inputs_ready="current") -> BatchResult:
        """All inputs already on the device.  `inputs_ready` says when `b.frames` / `b.crop_params` are complete, for the
        front stream (the resampler runs there): "current" (default) = produced by work queued on the current stream
        up to this call (safe, but the resampler then also waits for the previous step's tail: no overlap); a
        torch.cuda.Event = after that event; None = already complete (resident inputs, e.g. a replayed recording)."""
        S, H, V, W = b.frames.shape
        cur = torch.cuda.current_stream(self._device)
        front = self._front
        if front is not None:
            if isinstance(inputs_ready, str):
                front.wait_stream(cur)
            else:
                if inputs_ready is not None:
                    front.wait_event(inputs_ready)
                if self._enc_done is not None:
                    front.wait_event(self._enc_done)
                else:
                    front.wait_stream(cur)
        with torch.cuda.stream(front if front is not None else cur):
            crops = crop_resample_cuda(
                b.frames, H * V * W, W, V * W, b.crop_params, (96, 96), self._crop_dtype, self._map_mode,
                staged_bytes=None if b.frames.is_cuda else self.pulled_bytes,
            )
            # the next step's ingest kernel is released after this point: the resampler runs with its own
            # (L1-heavy) shared-memory configuration and must not meet resident ingest CTAs
            if not self._release_in_encoder:
                self._crops_done = torch.cuda.Event()
                self._crops_done.record()
        self._model.front_stream = front
        enc_ev = None
        if self.encoder_events is not None and front is None:
            enc_ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            enc_ev[0].record(cur)
        try:
            feats = self._model.encode(crops)
        finally:
            self._model.front_stream = None
        if enc_ev is not None:
            enc_ev[1].record(cur)
            self.encoder_events.append(enc_ev)
        if front is not None:
            self._enc_done = torch.cuda.Event()
            self._enc_done.record(cur)
        out = self._model.tail(
            feats, InputFrameData(None, b.intrinsics, b.extrinsics),
            InputFrameDesc(b.sample_range, b.memory_idx, b.use_memory, b.hand_idx), self._skel, _cabi.AT_HEAD_KNOWN_SKELETON)
        # metres -> millimetres on the wrist translations (tracker.py:379) inside the FK launch.
        # The model already mirrored right hands (umetrack_model.py:88-89); landmarks_from_hand_pose
        # mirrors them back before FK (perspective_crop.py:46-48) -> hand_idx does that in-kernel
        wrist_mm = torch.empty_like(out.wrist_xfs)
        lm = skin_landmarks_cuda(self._hm_dev, out.joint_angles, out.wrist_xfs, hand_idx=b.hand_idx,
                                 translation_scale=M_TO_MM, wrist_out=wrist_mm)
        return BatchResult(out.joint_angles, wrist_mm, lm)

    def step(self, host_batch: FrameBatch, out_host: Optional[dict] = None) -> dict:
        """Host buffers in, host buffers out: H2D on the copy stream into one of two staging slots
        (overlaps the previous call's kernels), kernels + D2H on the current stream.  Nothing here
        synchronises the host; results in `out_host` are valid after a stream synchronize."""
        slot, self._slot = self._slot, self._slot ^ 1
        cur = torch.cuda.current_stream(self._device)
        cs = self._copy_stream
        st = self._stage[slot]
        pinned = host_batch.frames.is_pinned()
        zero_copy = self._ingest == "zero_copy" and self._map_mode == "fp32" and pinned
        pull = self._ingest == "pull" and pinned
        fields = [f for f in host_batch.__dataclass_fields__ if not (zero_copy and f == "frames")]
        if st is None or any(getattr(st, f) is None or getattr(st, f).shape != getattr(host_batch, f).shape for f in fields):
            st = FrameBatch(*[
                (torch.empty_like(getattr(host_batch, f), device=self._device) if f in fields else None)
                for f in host_batch.__dataclass_fields__])
            self._stage[slot] = st
            self._stage_free[slot] = None
            # the caching allocator may hand out blocks whose previous users are still running on
            # the current stream: order the first copy-stream write after them
            cs.wait_stream(cur)
        if zero_copy:
            st.frames = host_batch.frames  # read in place by the resampler; the caller keeps it alive and unchanged
        if self._stage_free[slot] is not None:
            cs.wait_event(self._stage_free[slot])  # the kernels that last read this slot are done
        if self._crops_done is not None:
            cs.wait_event(self._crops_done)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)] if self.trace is not None else None
        with torch.cuda.stream(cs):
            if ev: ev[0].record(cs)
            for f in fields:
                if not (pull and f == "frames"):
                    getattr(st, f).copy_(getattr(host_batch, f), non_blocking=True)
            if pull:
                S, H, V, W = host_batch.frames.shape
                crop_pull_source_boxes_cuda(host_batch.frames, st.frames, H * V * W, W, V * W, st.crop_params,
                                            (96, 96), self.pull_ctas, self.pulled_bytes)
            ready = ev[1] if ev else torch.cuda.Event()
            ready.record(cs)
        cur.wait_event(ready)
        if ev: ev[2].record(cur)
        r = self.step_device(st, inputs_ready=ready)
        return self._finish(slot, r, ev, out_host)

    def _finish(self, slot, r: BatchResult, ev, out_host):
        cur = torch.cuda.current_stream(self._device)
        free = ev[3] if ev else torch.cuda.Event()
        free.record(cur)
        if ev: self.trace.append(ev)
        self._stage_free[slot] = free
        res = {k: v for k, v in vars(r).items() if v is not None}
        if out_host is None:
            out_host = {k: torch.empty(v.shape, dtype=v.dtype, pin_memory=True) for k, v in res.items()}
        for k, v in res.items():
            out_host[k][: v.shape[0]].copy_(v, non_blocking=True)
        return out_host

    # ------------------------------------------------------------------ poses in, poses out
    def attach_rig(self, cameras, camera_angles, opts: Optional[CropCameraOpts] = None) -> None:
        """Enable step_poses: the rig's cameras (constant per recording) go to the device once."""
        self._rig = DeviceRig(cameras, camera_angles, self._hm, self._device)
        self._cc_opts = opts or CropCameraOpts()
        self._history = None
        self._pstage = [None, None]

    def reset_history(self) -> None:
        if getattr(self, "_history", None) is not None:
            self._history.zero_()

    def step_poses(self, pb: PoseBatch, out_host: Optional[dict] = None) -> dict:
        """One time step from what a recording provides (frames, camera poses, the pose each crop is
        centred on): crop cameras + every model input are generated ON THE DEVICE
        (at_gen_crop_cameras: the reference's HandTracker.gen_crop_cameras + _make_inputs,
        tracker.py:222-260,315-368), then source-box pull -> resampler -> network -> FK as in step().
        The host waits once per call for the two output counts (n hands, c crops) of the generator,
        which ran on the side stream ahead of the previous step's kernels.
        Returns pinned host buffers at capacity size plus out["n"]: the first n rows are valid after
        a stream synchronize (rows follow out["memory_idx"] = sequence slot*2 + hand)."""
        if getattr(self, "_rig", None) is None:
            raise _cabi.AbsoluteTrackError("step_poses: call attach_rig(cameras, camera_angles) first")
        if not pb.frames.is_pinned():
            raise _cabi.AbsoluteTrackError("step_poses reads the frames in place: they must be pinned host memory")
        slot, self._slot = self._slot, self._slot ^ 1
        cur = torch.cuda.current_stream(self._device)
        cs = self._copy_stream
        S, H, V, W = pb.frames.shape
        P = int(pb.joint_angles.shape[0])
        small = [f for f in pb.__dataclass_fields__ if f != "frames"]
        st = self._pstage[slot]
        if st is None or st["frames"].shape != pb.frames.shape or st["joint_angles"].shape[0] != P:
            dev = self._device
            st = {f: torch.empty_like(getattr(pb, f), device=dev) for f in small}
            st["frames"] = torch.empty(pb.frames.shape, dtype=torch.uint8, device=dev)
            st["out"] = alloc_crop_inputs(P, self._cc_opts.max_view_num, dev)
            st["ws"] = torch.empty((int(_cabi.lib().at_crop_cameras_workspace_bytes(P, self._cc_opts.max_view_num)),),
                                   dtype=torch.uint8, device=dev)
            st["counts_host"] = torch.zeros((4,), dtype=torch.int32).pin_memory()
            self._pstage[slot] = st
            self._stage_free[slot] = None
            cs.wait_stream(cur)  # allocator reuse: see step()
        if self._history is None or self._history.numel() != 2 * S:
            self._history = torch.zeros((2 * S,), dtype=torch.bool, device=self._device)
            cs.wait_stream(cur)
        if self._stage_free[slot] is not None:
            cs.wait_event(self._stage_free[slot])
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)] if self.trace is not None else None
        with torch.cuda.stream(cs):
            if ev: ev[0].record(cs)
            for f in small:
                st[f].copy_(getattr(pb, f), non_blocking=True)
            gen_crop_inputs_cuda(self._rig, st["joint_angles"], st["wrist_xforms"], st["pose_frame"], st["pose_hand"],
                                 st["camera_to_world"], self._cc_opts, pose_valid=st["pose_valid"],
                                 tracking_history=self._history, out=st["out"], workspace=st["ws"])
            st["counts_host"].copy_(st["out"].counts, non_blocking=True)
            counted = torch.cuda.Event()
            counted.record(cs)
        counted.synchronize()  # the only host wait: (n, c) size the launches below
        ci = st["out"].resolve(st["counts_host"])
        if ci.n == 0:
            self._stage_free[slot] = None
            if out_host is not None:
                out_host["n"] = 0
            return out_host if out_host is not None else {"n": 0}
        if self._crops_done is not None:
            cs.wait_event(self._crops_done)
        with torch.cuda.stream(cs):
            crop_pull_source_boxes_cuda(pb.frames, st["frames"], H * V * W, W, V * W, ci.crop_params, (96, 96),
                                        self.pull_ctas, self.pulled_bytes)
            ready = ev[1] if ev else torch.cuda.Event()
            ready.record(cs)
        cur.wait_event(ready)
        if ev: ev[2].record(cur)
        r = self.step_device(FrameBatch(st["frames"], ci.crop_params, ci.intrinsics, ci.extrinsics, ci.sample_range,
                                        ci.hand_idx, ci.memory_idx, ci.use_memory), inputs_ready=ready)
        r.memory_idx = ci.memory_idx
        if out_host is None:
            out_host = {
                "joint_angles": torch.empty((P, 22), dtype=torch.float32, pin_memory=True),
                "wrist_xforms": torch.empty((P, 4, 4), dtype=torch.float32, pin_memory=True),
                "landmarks": torch.empty((P, 21, 3), dtype=torch.float32, pin_memory=True),
                "memory_idx": torch.empty((P,), dtype=torch.int64, pin_memory=True),
            }
        out_host["n"] = ci.n
        out_host["c"] = ci.c
        return self._finish(slot, r, ev, out_host)

    def check(self) -> None:
        self._model.check_status()

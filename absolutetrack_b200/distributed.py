"""Multi-GPU plumbing of the hot path: sequences (recordings) are independent, so they are sharded
across ranks with no data-path collective; the only exchange is the sum of per-shard metric
accumulators at the end.

shard rule = the reference Sampler's rank striding (lib/data_utils/async_dataset.py:558-559):
rank r of `world` takes items r, r + world, r + 2*world, ...
"""
from typing import Sequence

import torch


def shard_indices(n_items: int, rank: int, world: int) -> range:
    """Indices of the items (sequences / recordings) rank `rank` processes."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world of {world}")
    return range(rank, n_items, world)


def shard(items: Sequence, rank: int, world: int) -> list:
    return [items[i] for i in shard_indices(len(items), rank, world)]


def all_reduce_metric_sums(sums: torch.Tensor) -> torch.Tensor:
    """Sum the float64 accumulators {sum err_mm, sum accel, sum |dtheta|, n_valid, n_triples, ...}
    of all ranks in place (one all_reduce over NCCL / NVLink on GPUs, gloo on CPU tests).  Returns
    `sums`; a no-op when torch.distributed is not initialised (single process)."""
    import torch.distributed as dist

    if sums.dtype != torch.float64:
        raise TypeError("metric accumulators must be float64 (they are sums over up to 1e6 frames)")
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    return sums


def mpjpe_accumulate(sums: torch.Tensor, tracked: torch.Tensor, gt: torch.Tensor, valid: torch.Tensor) -> None:
    """sums[0] += sum over valid hands of the mean-over-21-landmarks error (mm), sums[1] += count
    (run_eval_known_skeleton.py:92-93 accumulated over a shard)."""
    d = (tracked.double() - gt.double()).norm(dim=-1).mean(dim=-1)
    v = valid.to(d.dtype)
    sums[0] += (d * v).sum()
    sums[1] += v.sum()


def accumulate_metric_sums_cuda(landmarks, gt_landmarks, sums, valid=None, joint_angles=None, gt_joint_angles=None):
    """sums (float64 [3], CUDA) += {sum of per-hand mean keypoint error, sum of per-hand mean |dtheta| in
    degrees, number of valid hands} of one batch, in ONE launch of csrc/fk_skin.cu::eval_accumulate_kernel
    (C-ABI at_eval_accumulate_f32; the bookkeeping of run_eval_known_skeleton.py:92-93 on the device).
    Feed the result to all_reduce_metric_sums at the end of a sharded run."""
    import torch

    from . import _cabi

    lib = _cabi.lib()
    n = int(landmarks.shape[0])
    if not (sums.is_cuda and sums.dtype == torch.float64 and sums.numel() >= 3):
        raise TypeError("sums must be a CUDA float64 tensor with >= 3 elements")
    f = lambda t: None if t is None else t.to(device=sums.device, dtype=torch.float32).contiguous()  # noqa: E731
    lm, gt, ja, gja = f(landmarks), f(gt_landmarks), f(joint_angles), f(gt_joint_angles)
    v = None if valid is None else valid.to(device=sums.device, dtype=torch.uint8).contiguous()
    with torch.cuda.device(sums.device):
        rc = lib.at_eval_accumulate_f32(lm.data_ptr(), gt.data_ptr(), _cabi.ptr(ja), _cabi.ptr(gja), _cabi.ptr(v), n,
                                        sums.data_ptr(), _cabi.stream_ptr(sums.device))
    _cabi.check(rc, "at_eval_accumulate_f32")
    return sums

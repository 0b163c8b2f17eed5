"""Evaluation table on the GPU (mirror of the reference's load_eval.py).

The reference walks the result pickles of run_eval_*.py, computes per-frame errors and accelerations
in numpy (`_compute_metrics`, load_eval.py:26-44), concatenates them over recordings and prints the
success rate, mean keypoint error, PCK-AUC and mean accelerations (`aggregate_metrics`, :47-92).
Here one launch per recording (C-ABI at_eval_recording_f32) turns the keypoints into SUMS and integer
PCK counts that add up across recordings and - with one all-reduce - across ranks; the table is
computed from them at the end.  There is no CPU path for the per-frame work.
"""
import fnmatch
import os
import pickle
from typing import Dict, Optional

import numpy as np
import torch

try:
    from . import _cabi
    from .lib.common import metric_utils
except ImportError:  # found as top-level `load_eval` (absolutetrack_b200/ itself on sys.path)
    import sys

    _root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if _root not in sys.path:
        sys.path.append(_root)
    from absolutetrack_b200 import _cabi
    from absolutetrack_b200.lib.common import metric_utils


class EvalSums:
    """Device accumulators: sums f64 [3] = {sum error, sum accel, sum gt accel}; counts i64 [3 + K] =
    {valid hand-frames, hand-frames, valid triples, #(error <= threshold_k)}."""

    def __init__(self, device=None, thresholds: Optional[np.ndarray] = None):
        if not torch.cuda.is_available():
            raise _cabi.AbsoluteTrackError("EvalSums needs a CUDA device (no CPU fallback)")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        thr = metric_utils.PCK_THRESHOLDS if thresholds is None else np.asarray(thresholds, np.float64)
        if thr.ndim != 1 or thr.size > 128:
            raise ValueError("thresholds: 1-D, at most 128 entries")
        self.thresholds_host = thr.copy()
        self.thresholds = torch.from_numpy(self.thresholds_host).to(self.device)
        self.sums = torch.zeros(3, dtype=torch.float64, device=self.device)
        self.counts = torch.zeros(3 + thr.size, dtype=torch.int64, device=self.device)

    def all_reduce(self) -> "EvalSums":
        """Sum the accumulators of all ranks (NCCL over NVLink; no-op in a single process)."""
        all_reduce_eval_sums(self.sums, self.counts)
        return self

    def table(self) -> Dict[str, object]:
        """What aggregate_metrics prints (load_eval.py:62-92).  PCK is in percent and the AUC is
        normalised with y_max = 1, exactly as the reference calls it (:63-70)."""
        return table_from_sums(self.sums.cpu().numpy(), self.counts.cpu().numpy(), self.thresholds_host)


def all_reduce_eval_sums(sums: torch.Tensor, counts: torch.Tensor) -> None:
    """The only exchange of a sharded evaluation: float64 sums and int64 counts, summed in place."""
    import torch.distributed as dist

    if sums.dtype != torch.float64 or counts.dtype != torch.int64:
        raise TypeError("evaluation accumulators are float64 sums and int64 counts")
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
        dist.all_reduce(counts, op=dist.ReduceOp.SUM)


def table_from_sums(sums: np.ndarray, counts: np.ndarray, thresholds: np.ndarray) -> Dict[str, object]:
    n_valid, n_total, n_acc = int(counts[0]), int(counts[1]), int(counts[2])
    pck = metric_utils.PCK_curve_from_counts(counts[3:], n_valid) * 100.0
    return {
        "n_valid": n_valid,
        "n_total": n_total,
        "success_rate": n_valid / n_total * 100 if n_total else 0.0,
        "mean_keypoint_error": float(sums[0]) / n_valid if n_valid else float("nan"),
        "pck_curve": pck,
        "auc_score": float(metric_utils.normalized_AUC(thresholds, pck)) if thresholds.size > 1 else float("nan"),
        "mean_keypoint_acceleration": float(sums[1]) / n_acc if n_acc else float("nan"),
        "gt_mean_keypoint_acceleration": float(sums[2]) / n_acc if n_acc else float("nan"),
    }


def compute_metric_sums_cuda(gt_keypoints, tracked_keypoints, valid_tracking, acc: EvalSums) -> EvalSums:
    """`_compute_metrics` (load_eval.py:26-44) of one recording, accumulated into `acc` on the device.
    gt_keypoints / tracked_keypoints [n_hands, T, 21, 3] (numpy or torch, any float dtype: the
    reference's arrays hold float32 values), valid_tracking [n_hands, T] bool."""
    lib = _cabi.lib()
    dev = acc.device
    f = lambda a: torch.as_tensor(a).to(device=dev, dtype=torch.float32).contiguous()  # noqa: E731
    gt, tr = f(gt_keypoints), f(tracked_keypoints)
    v = torch.as_tensor(valid_tracking).to(device=dev, dtype=torch.uint8).contiguous()
    if gt.shape != tr.shape or gt.dim() != 4 or tuple(gt.shape[2:]) != (21, 3) or tuple(v.shape) != tuple(gt.shape[:2]):
        raise ValueError(f"expected [H,T,21,3] keypoints and [H,T] validity, got {tuple(gt.shape)}, {tuple(tr.shape)}, {tuple(v.shape)}")
    with torch.cuda.device(dev):
        rc = lib.at_eval_recording_f32(tr.data_ptr(), gt.data_ptr(), v.data_ptr(), int(gt.shape[0]), int(gt.shape[1]),
                                       acc.thresholds.data_ptr(), int(acc.thresholds.numel()), acc.sums.data_ptr(),
                                       acc.counts.data_ptr(), _cabi.stream_ptr(dev))
    _cabi.check(rc, "at_eval_recording_f32")
    return acc


def aggregate_metrics(output_dir: str, rank: int = 0, world: int = 1, verbose: bool = True) -> Optional[Dict[str, object]]:
    """load_eval.aggregate_metrics (:47-92): every `*.npy` result pickle under `output_dir`
    ({'tracked_keypoints', 'gt_keypoints', 'valid_tracking'}, written by run_eval_*.py:96-104).
    With world > 1 the files are strided across ranks (distributed.shard) and the accumulators
    all-reduced; every rank returns the same table."""
    from absolutetrack_b200.distributed import shard

    files = []
    for cur_dir, _, filenames in os.walk(output_dir):
        files += [os.path.join(cur_dir, f) for f in fnmatch.filter(filenames, "*.npy")]
    files.sort()
    if not files:
        return None
    acc = EvalSums()
    for path in shard(files, rank, world):
        with open(path, "rb") as fp:
            d = pickle.load(fp)
        compute_metric_sums_cuda(d["gt_keypoints"], d["tracked_keypoints"], d["valid_tracking"], acc)
    t = acc.all_reduce().table()
    if verbose and rank == 0:
        print(f"  Tracked {t['n_valid']} out of {t['n_total']}, success rate: {t['success_rate']}%")
        print(f"  Mean keypoint error: {t['mean_keypoint_error']}")
        print(f"  AUC score: {t['auc_score']}")
        print(f"  Mean keypoint accelerations: {t['mean_keypoint_acceleration']}")
        print(f"  GT mean keypoint accelerations: {t['gt_mean_keypoint_acceleration']}")
    return t


if __name__ == "__main__":
    # same walk as the reference's `python load_eval.py` (:95-110): tmp/eval_results_<mode>/real/<protocol>
    # under the given root (default: the current directory); under torchrun the files of each
    # directory are strided over the ranks and the sums all-reduced
    import sys

    root = sys.argv[1] if len(sys.argv) > 1 else os.getcwd()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        import torch.distributed as dist

        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        dist.init_process_group("nccl")
    for eval_mode in ("known_skeleton", "unknown_skeleton"):
        for protocol in ("separate_hand", "hand_hand"):
            if rank == 0:
                print(f"Evaluation for {eval_mode} on protocol {protocol}")
            aggregate_metrics(os.path.join(root, "tmp", f"eval_results_{eval_mode}", "real", protocol), rank, world)
    if world > 1:
        dist.destroy_process_group()

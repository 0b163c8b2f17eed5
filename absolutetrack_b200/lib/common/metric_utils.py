"""PCK / AUC bookkeeping (mirror of the reference's lib/common/metric_utils.py).

The per-frame work (errors, threshold counts) runs on the GPU (csrc/fk_skin.cu::eval_recording_kernel,
see absolutetrack_b200/load_eval.py); what is left here is arithmetic on the 101 counts.
"""
import numpy as np

MAX_LANDMARK_ERROR_MM = 50                                   # metric_utils.py:16
PCK_THRESHOLDS = np.linspace(0, MAX_LANDMARK_ERROR_MM, 101)  # metric_utils.py:17


def PCK_curve_from_counts(counts_le_threshold: np.ndarray, n_samples: int, eps: float = 1e-6) -> np.ndarray:
    """metric_utils._PCK_curve (:34-44) with the comparisons already counted: counts / n, 0 where
    n < eps (_safe_div :20-31)."""
    counts = np.asarray(counts_le_threshold, np.float64)
    if n_samples < eps:
        return np.zeros_like(counts)
    return counts / float(n_samples)


def normalized_AUC(x: np.ndarray, y: np.ndarray, y_max: float = 1.0) -> np.ndarray:
    """metric_utils.normalized_AUC (:95-121): trapezoid area of curves sharing the x axis, divided by
    (x[-1] - x[0]) * y_max; shape = y.shape[:-1]."""
    x = np.asarray(x, np.float64)
    y = np.asarray(y, np.float64)
    out_shape = y.shape[:-1]
    y = y.reshape(-1, y.shape[-1])
    auc = ((x[1:] - x[:-1]).reshape(1, -1) * ((y[..., 1:] + y[..., :-1]) * 0.5)).sum(axis=-1)
    return (auc / ((x[-1] - x[0]) * y_max)).reshape(out_shape)

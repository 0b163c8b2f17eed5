"""TEST FIXTURE - a stand-in for PyAV (`import av`), which is not installed in this image.

The reference's entry scripts read their frames through lib/tracker/video_pose_data.py:37-55:
`av.open(path)`, `container.streams.video[0].frames`, `container.decode(stream)` and
`np.array(frame.to_image())[..., 0]`.  This module serves exactly that surface from a tiny JSON
descriptor stored under the `.mp4` name,

    {"absolutetrack_fake_mp4": 1, "seed": 7, "n_frames": 20, "height": 480, "width": 2544}

and synthesises the frames deterministically with integer arithmetic only (a smooth random
texture plus a little noise), so that the reference run that produced the goldens
(oracle/make_golden_scripts.py, build container) and the run under test (GPU box) decode
bit-identical "videos".  Never imported by the product.
"""
import builtins
import json

import numpy as np

CELL = 16


def synth_video_frame(seed: int, index: int, height: int, width: int) -> np.ndarray:
    """uint8 [height, width]: 16-px random lattice, integer bilinear interpolation, +-6 noise."""
    rng = np.random.default_rng([int(seed), int(index)])
    gh, gw = height // CELL + 2, width // CELL + 2
    g = rng.integers(0, 256, (gh, gw), dtype=np.int64)
    y, x = np.arange(height), np.arange(width)
    gy, fy = (y // CELL)[:, None], (y % CELL)[:, None]
    gx, fx = (x // CELL)[None, :], (x % CELL)[None, :]
    acc = (g[gy, gx] * (CELL - fy) * (CELL - fx) + g[gy, gx + 1] * (CELL - fy) * fx
           + g[gy + 1, gx] * fy * (CELL - fx) + g[gy + 1, gx + 1] * fy * fx)
    img = (acc + CELL * CELL // 2) // (CELL * CELL) + rng.integers(-6, 7, (height, width), dtype=np.int64)
    return np.clip(img, 0, 255).astype(np.uint8)


class _Frame:
    def __init__(self, mono: np.ndarray):
        self._mono = mono

    def to_image(self):
        return np.broadcast_to(self._mono[..., None], (*self._mono.shape, 3))


class _Stream:
    def __init__(self, desc):
        self.frames = int(desc["n_frames"])
        self.average_rate = 30


class _Streams:
    def __init__(self, desc):
        self.video = [_Stream(desc)]


class _Container:
    def __init__(self, path: str):
        with builtins.open(path, "rb") as f:
            self._desc = json.load(f)
        assert self._desc.get("absolutetrack_fake_mp4") == 1, f"{path} is not a fake-av descriptor"
        self.streams = _Streams(self._desc)

    def decode(self, stream):
        d = self._desc
        for i in range(int(d["n_frames"])):
            yield _Frame(synth_video_frame(d["seed"], i, d["height"], d["width"]))


def open(path: str, *args, **kwargs):  # noqa: A001 (PyAV's name)
    return _Container(path)

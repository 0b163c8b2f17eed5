"""Reader (and fixture writer) for the `.torch.idx` / `.torch.bin` pairs of the torch_data sets
(mirror of the reference's lib/data_utils/idxbinfile.py; SURVEY 8f rank 3, host side).

Format (idxbinfile.py:111-123): the idx file is an int64 array
    [magic 0x584449544E54, version 1, dtype code, itemsize, N, S,
     N+1 dim offsets (into `sizes`), N+1 data offsets (units of itemsize, into the .bin), S sizes]
dtype codes 1..7 = uint8, int8, int16, int32, int64, float32, float64 tensors, 8 = msgpack objects
(:358-367).  Version 0 files carry magic 0 (:137-139).  This is host IO: numpy views over the bytes,
no copies, no GPU work; `write_torch_idx_bin` exists for fixtures and tests (the reference ships no
writer).
"""
import math
import os
from enum import Enum
from typing import Any, Dict, List, Optional, Sequence, Tuple, Union

import numpy as np

# type names the reference's own readers import from this module (async_dataset.py:42, idxbinfile.py:24-33)
Buffer = Union[bytes, memoryview]
MsgpackObject = Dict[str, Any]
RawField = Union[np.ndarray, MsgpackObject]


class BinFormat(Enum):
    TENSOR = 0
    MSGPACK = 1


MAGIC = 0x584449544E54
_CODE_TO_DTYPE = {1: "uint8", 2: "int8", 3: "int16", 4: "int32", 5: "int64", 6: "float32", 7: "float64", 8: "object"}
_DTYPE_TO_CODE = {v: k for k, v in _CODE_TO_DTYPE.items()}
OBJECT_DTYPE = np.dtype("object")
IDX_SUFFIX, BIN_SUFFIX = ".torch.idx", ".torch.bin"


def _get_bin_path_for_idx(idx_path: str) -> str:
    assert idx_path.endswith(IDX_SUFFIX)  # idxbinfile.py:36-42
    return idx_path[: -len(IDX_SUFFIX)] + BIN_SUFFIX


def _as_int64(buffer) -> np.ndarray:
    mv = memoryview(buffer)
    if mv.ndim != 1:
        raise ValueError(f".torch.idx data has invalid shape: {mv.shape}; require ndim=1")
    if mv.format == "B":
        if len(mv) % 8 != 0:
            raise ValueError(f".torch.idx data has invalid length: {len(mv)}%8 != 0")
    elif mv.format not in ("q", "l") or mv.itemsize != 8:
        raise ValueError(f".torch.idx data has invalid format {mv.format}: expected 'B' (bytes)  or 'q' or 'l' (int64)")
    return np.frombuffer(mv, dtype=np.int64)


class TorchIdx:
    """Layout of one .torch.bin file: `len(idx)` frames, `idx.dims[i]` their shapes, `idx.dtype`,
    `idx.shape` (whole-file shape when every frame has the same dims, else None), byte offsets.
    Same attributes and methods as the reference class (idxbinfile.py:48-347)."""

    def __init__(self, path: str, bin_path: Optional[str] = None, buffer=None) -> None:
        self.source = path
        self.bin_path = _get_bin_path_for_idx(path) if bin_path is None else bin_path
        if buffer is None:
            with open(path, "rb") as fp:
                buffer = fp.read()
        head = _as_int64(buffer)
        magic, version = int(head[0]), int(head[1])
        if version not in (0, 1):
            raise ValueError(f"unsupported version {version} in .torch.idx file {path}")
        if magic != (MAGIC if version == 1 else 0):
            raise ValueError(f"bad magic number in .torch.idx file {path}")
        code = int(head[2])
        self.itemsize = int(head[3])
        if code not in _CODE_TO_DTYPE:
            raise KeyError(f"unrecognized type {code}")
        self.dtype = np.dtype(_CODE_TO_DTYPE[code])
        self._msgpack = code == 8
        if not self._msgpack and self.dtype.itemsize != self.itemsize:
            raise ValueError(f"item size {self.itemsize} not compatible with dtype {self.dtype.name}.itemsize={self.dtype.itemsize}")
        n, s = int(head[4]), int(head[5])
        dim_off = head[6: 7 + n]
        data_off = head[7 + n: 8 + 2 * n]
        sizes = head[8 + 2 * n: 8 + 2 * n + s]
        dims: Any = [tuple(int(v) for v in sizes[dim_off[i]: dim_off[i + 1]]) for i in range(n)]
        self.is_uniform = False
        self.shape: Optional[Tuple[int, ...]] = None
        self.byte_offsets_array: Optional[np.ndarray] = data_off * self.itemsize
        if not self._msgpack and n > 0 and all(len(d) == len(dims[0]) for d in dims):
            arr = np.array(dims, dtype=np.int64).reshape(n, len(dims[0]))
            if (arr == arr[0]).all():
                # every frame alike: one shape for the file, offsets recomputed on the fly (:171-186)
                dims = np.broadcast_to(arr[0].copy(), arr.shape)
                self.is_uniform = True
                self.shape = (n, *(int(v) for v in arr[0]))
                assert np.all(np.diff(data_off) == math.prod(self.shape[1:]))
                if int(data_off[0]) == 0:
                    self.byte_offsets_array = None
            else:
                dims = arr
        self.dims = dims

    def __len__(self) -> int:
        return len(self.dims)

    def __repr__(self) -> str:
        if self.is_uniform:
            return f"TorchIdx(shape={self.shape}, dtype={self.dtype}, source={self.source})"
        return f"TorchIdx(N={len(self)}, dtype={self.dtype}, is_uniform=False, source={self.source})"

    # ---- offsets
    def _frame_bytes(self) -> int:
        assert self.shape is not None
        return math.prod(self.shape[1:]) * self.itemsize

    def byte_offset(self, i: int) -> int:
        if self.byte_offsets_array is not None:
            return int(self.byte_offsets_array[i])
        n = self.shape[0]
        i = n if i == -1 else i
        assert i <= n
        return self._frame_bytes() * i

    def byte_offsets(self, start: int, end: int) -> np.ndarray:
        if self.byte_offsets_array is not None:
            return self.byte_offsets_array[start:end]
        n = self.shape[0]
        end = n if end == -1 else end
        assert start <= n and end <= n + 1
        return self._frame_bytes() * np.arange(start, end, dtype=np.ulonglong)

    def data_size_bytes(self) -> int:
        return self.byte_offset(-1) - self.byte_offset(0)

    def frame_size_bytes(self) -> int:
        if not self.is_uniform:
            raise ValueError("frame_size_bytes requires that all frames are the same size")
        return self.byte_offset(1) - self.byte_offset(0)

    def item_shape(self, i: Optional[int] = None) -> Tuple[int, ...]:
        if i is None:
            if self.shape is None:
                raise ValueError(f"Dataset does not have uniform shape: {self.source}")
            return self.shape[1:]
        return tuple(int(v) for v in self.dims[i])

    # ---- data
    def _check(self, istart: int, istop: int, buffer) -> None:
        got = memoryview(buffer).nbytes
        want = self.byte_offset(istop) - self.byte_offset(istart)
        if want != got:
            raise ValueError(f"expected {want} bytes but got {got} for {self.bin_path}[{istart}:{istop}]")

    def read_bin(self):
        """The whole .bin file: one ndarray for uniform files, else a list-like of frames."""
        s0, s1 = self.byte_offset(0), self.byte_offset(-1)
        with open(self.bin_path, "rb") as fp:
            fp.seek(s0)
            buf = fp.read(s1 - s0)
        return self.view_buffer(buf)

    def view_buffer(self, buffer):
        self._check(0, -1, buffer)
        if self.shape is not None:
            return np.ndarray(shape=self.shape, dtype=self.dtype, buffer=buffer)
        return _LazyFrames(self, buffer)

    def view_buffer_at(self, index: int, buffer):
        lo, hi = (int(v) for v in self.byte_offsets(index, index + 2))
        return self.view_frame(index, buffer[lo:hi])

    def view_frame(self, index: int, buffer):
        self._check(index, index + 1, buffer)
        if self._msgpack:
            import msgpack

            return msgpack.unpackb(buffer)
        return np.ndarray(shape=tuple(int(v) for v in self.dims[index]), dtype=self.dtype, buffer=buffer)


class _LazyFrames:
    """Frames of a non-uniform / msgpack file, decoded on access (the reference's map_dataset view)."""

    def __init__(self, idx: TorchIdx, buffer) -> None:
        self._idx, self._buf = idx, memoryview(buffer)
        self._base = idx.byte_offset(0)

    def __len__(self) -> int:
        return len(self._idx)

    def __getitem__(self, i: int):
        if i < 0:
            i += len(self)
        if not 0 <= i < len(self):
            raise IndexError(i)
        lo, hi = self._idx.byte_offset(i) - self._base, self._idx.byte_offset(i + 1) - self._base
        return self._idx.view_frame(i, self._buf[lo:hi])

    def __iter__(self):
        return (self[i] for i in range(len(self)))


def write_torch_idx_bin(idx_path: str, frames: Sequence[Union[np.ndarray, dict]], dtype: Optional[str] = None) -> None:
    """Write `frames` (ndarrays of one dtype, or msgpack-able dicts) as a version-1 pair.  Not in the
    reference (it only reads); used to build fixtures and in tests."""
    import msgpack

    is_obj = len(frames) > 0 and not isinstance(frames[0], np.ndarray)
    name = "object" if is_obj else (dtype or str(frames[0].dtype))
    code = _DTYPE_TO_CODE[name]
    itemsize = 1 if is_obj else np.dtype(name).itemsize
    blobs: List[bytes] = []
    dims: List[Tuple[int, ...]] = []
    for f in frames:
        if is_obj:
            b = msgpack.packb(f)
            dims.append((len(b),))
        else:
            a = np.ascontiguousarray(f, dtype=name)
            b = a.tobytes()
            dims.append(tuple(a.shape))
        blobs.append(b)
    dim_off = np.concatenate([[0], np.cumsum([len(d) for d in dims])]).astype(np.int64)
    data_off = np.concatenate([[0], np.cumsum([len(b) // itemsize for b in blobs])]).astype(np.int64)
    sizes = np.array([v for d in dims for v in d], dtype=np.int64)
    head = np.array([MAGIC, 1, code, itemsize, len(frames), sizes.size], dtype=np.int64)
    os.makedirs(os.path.dirname(os.path.abspath(idx_path)), exist_ok=True)
    with open(idx_path, "wb") as fp:
        fp.write(np.concatenate([head, dim_off, data_off, sizes]).tobytes())
    with open(_get_bin_path_for_idx(idx_path), "wb") as fp:
        for b in blobs:
            fp.write(b)

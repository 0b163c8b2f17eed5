"""Generic helpers over dicts / dataclasses / NamedTuples of tensors (mirror of the subset of the
reference's lib/data_utils/bundles.py the hot path calls: map_fields :113-142, to_device :145-149,
collate :152-201)."""
import dataclasses
from typing import Any, Mapping

import numpy as np
import torch


def is_dictlike(obj: Any) -> bool:
    return dataclasses.is_dataclass(obj) or hasattr(obj, "_asdict") or isinstance(obj, Mapping)


def asdict(obj):
    if dataclasses.is_dataclass(obj):
        return {f.name: getattr(obj, f.name) for f in dataclasses.fields(obj)}
    if hasattr(obj, "_asdict"):
        return obj._asdict()
    if isinstance(obj, Mapping):
        return obj
    raise TypeError("asdict() requires a Mapping, dataclass, or NamedTuple")


def map_fields(func, obj, only_type=object):
    """Apply `func` to every leaf (of type `only_type`) of a nested container, keeping its types."""
    if is_dictlike(obj):
        if isinstance(obj, Mapping):
            return type(obj)((k, map_fields(func, v, only_type)) for k, v in obj.items())
        return type(obj)(**{k: map_fields(func, v, only_type) for k, v in asdict(obj).items()})
    if isinstance(obj, tuple):
        return tuple(map_fields(func, v, only_type) for v in obj)
    if isinstance(obj, list):
        return [map_fields(func, v, only_type) for v in obj]
    return func(obj) if isinstance(obj, only_type) else obj


def to_device(obj, device):
    return map_fields(lambda t: t.to(device), obj, only_type=torch.Tensor)


def group(batch, group_fn):
    first = batch[0]
    if isinstance(first, (np.ndarray, np.generic, torch.Tensor)):
        return group_fn(batch)
    if is_dictlike(first):
        rows = [asdict(x) for x in batch]
        return type(first)(**{k: group([r[k] for r in rows], group_fn) for k in rows[0]})
    if isinstance(first, tuple):
        return tuple(group([v[i] for v in batch], group_fn) for i in range(len(first)))
    if isinstance(first, list):
        return [group([v[i] for v in batch], group_fn) for i in range(len(first))]
    if any(v is None for v in batch):
        if all(v is None for v in batch):
            return None
        raise TypeError("Some items are None and others are not")
    return batch


def collate(batch, device=None):
    """Stack a list of N like-typed items into one item of size-N arrays."""

    def stack(items):
        if isinstance(items[0], (np.ndarray, np.generic)):
            return np.stack(items)
        if isinstance(items[0], torch.Tensor):
            t = torch.stack(items)
            return t.to(device) if device is not None else t
        raise TypeError(f"Can't stack tensors: unknown type {type(items[0])} found")

    return group(batch, stack)

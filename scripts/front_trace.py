"""Where do the resampler + stem of step i+1 run relative to the tail of step i?  CUDA events on both streams of
BatchedHandTracker (front stream / main stream) over a few device-resident steps; prints one steady-state step."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from absolutetrack_b200 import synthetic
from absolutetrack_b200.lib.models.model_loader import make_random_state_dict, model_from_state_dict
from absolutetrack_b200.lib.tracker import batch_tracker as BT
from absolutetrack_b200.lib.tracker.batch_tracker import BatchedHandTracker

dev = torch.device("cuda", 0)
S = 512
model = model_from_state_dict(make_random_state_dict(0), precision="fp16")
tr = BatchedHandTracker(model, synthetic.hand_model(), max_slots=2 * S, map_mode="fp32", ingest="pull")
seqs = np.arange(S)
g = torch.Generator(device=dev).manual_seed(1)
frames = torch.randint(0, 256, (S, synthetic.H, synthetic.V, synthetic.W), dtype=torch.uint8, device=dev, generator=g)
bs = [synthetic.make_batch(seqs, t, frames=torch.zeros(1, dtype=torch.uint8)).to(dev) for t in range(2)]
for b in bs:
    b.frames = frames
marks = []   # (label, event)


def mark(label, stream=None):
    e = torch.cuda.Event(enable_timing=True)
    e.record(stream if stream is not None else torch.cuda.current_stream(dev))
    marks.append((label, e))


# instrument: wrap the three stages
orig_crop, orig_encode, orig_tail = BT.crop_resample_cuda, model.encode, model.tail


def crop(*a, **k):
    mark("crop start")
    r = orig_crop(*a, **k)
    mark("crop end")
    return r


def encode(c):
    r = orig_encode(c)
    if tr._front is not None:
        mark("stem end (front stream idle)", tr._front)
    mark("encoder end")
    return r


def tail(*a, **k):
    r = orig_tail(*a, **k)
    mark("tail end")
    return r


BT.crop_resample_cuda, model.encode, model.tail = crop, encode, tail
for i in range(6):
    if i == 4:
        torch.cuda.synchronize()
        marks.clear()
        mark("t0")
    tr.step_device(bs[i & 1], inputs_ready=None)
    mark("fk end")
torch.cuda.synchronize()
t0 = marks[0][1]
print("AT_FRONT_OVERLAP =", os.environ.get("AT_FRONT_OVERLAP", "1"))
for label, e in marks[1:]:
    print(f"{t0.elapsed_time(e):8.3f} ms  {label}")

"""Weight ingestion (mirror of the reference's lib/models/model_loader.py API).

`load_pretrained_model(path)` consumes the reference's checkpoint unchanged (flat state dict,
252 tensors, model_loader.py:84-87), folds every eval-mode BatchNorm (eps 1e-5) into the
preceding conv / linear and packs one float32 blob in the layout the CUDA library declares
through at_model_layer_table() ([kh][kw][Cin][CoutP] + bias[CoutP] per layer).
"""
import math
import os
from typing import Dict, List, Tuple

import torch

from .. import _cabi
from .model_opts import ModelOpts
from .umetrack_model import UmeTrackModel

BN_EPS = 1e-5
_BB = "_feature_extractor._image_backbone"


def _layer_keys(name: str) -> Tuple[str, str]:
    """C layer name -> (conv/linear key prefix, batchnorm key prefix or '') in the state dict."""
    p = name.split(".")
    if name == "backbone.stem":
        return f"{_BB}.0._layers.0.0", f"{_BB}.0._layers.0.1"
    if name == "backbone.proj":
        return f"{_BB}.1", ""
    if p[0] == "backbone":
        base = f"{_BB}.0._layers.{p[1][len('layer'):]}.{p[2]}"
        if p[3] == "down":
            return base + ".downsample.0", base + ".downsample.1"
        return f"{base}.{p[3]}", f"{base}.bn{p[3][-1]}"
    if p[0] == "fusion":
        i = int(p[1])
        conv = f"_feature_extractor._multi_view_fusion.{3 * i}"
        return conv, (f"_feature_extractor._multi_view_fusion.{3 * i + 1}" if i < 2 else "")
    if p[0] == "temporal":
        return f"_temporal._temporal_module.{2 * int(p[1])}", ""
    if name == "skeleton":
        return "_skeleton_enc._layers.0", "_skeleton_enc._layers.2"
    if p[0] in ("head_k", "head_u"):
        base = f"_regressor_{p[0][-1]}._pose_regression_layers"
        if p[1] == "out":
            return base + ".2", ""
        return f"{base}.{p[1]}.{p[2]}", f"{base}.{p[1]}.bn{p[2][-1]}"
    raise KeyError(name)


def _has_bias(name: str) -> bool:
    p = name.split(".")
    if p[0] == "backbone":
        return name in ("backbone.stem", "backbone.proj")
    if p[0] in ("head_k", "head_u"):
        return p[1] == "out"
    return True


def network_layers() -> List[Dict]:
    """The 46 conv / linear layers of the network in blob order: name, cin, cout, ksize, stride.
    Pure Python (the architecture is fixed: model_opts.py:12 `resnet_layers_2352-f32`,
    model_utils.py:107-208); tests assert that it equals the C library's at_model_layer_table()."""
    out = []

    def add(name, cin, cout, k, stride=1):
        out.append(dict(name=name, cin=cin, cout=cout, ksize=k, stride=stride))

    add("backbone.stem", 1, 32, 3)
    cin = 32
    for li, (nb, cout, stride) in enumerate(((2, 32, 1), (3, 64, 2), (5, 128, 2), (2, 256, 2)), start=1):
        for b in range(nb):
            st = stride if b == 0 else 1
            add(f"backbone.layer{li}.{b}.conv1", cin, cout, 3, st)
            add(f"backbone.layer{li}.{b}.conv2", cout, cout, 3)
            if st != 1 or cin != cout:
                add(f"backbone.layer{li}.{b}.down", cin, cout, 1, st)
            cin = cout
    add("backbone.proj", 256, 72, 1)
    for i, (ci, co) in enumerate(((144, 108), (108, 72), (72, 72))):
        add(f"fusion.{i}", ci, co, 1)
    for i in range(3):
        add(f"temporal.{i}", 90, 90, 1)
    add("skeleton", 132, 144, 1)
    for head, c, n_out in (("head_k", 76, 62), ("head_u", 72, 63)):
        for blk in range(2):
            add(f"{head}.{blk}.conv1", c, c, 3)
            add(f"{head}.{blk}.conv2", c, c, 3)
        add(f"{head}.out", c, n_out, 1)
    return out


def state_dict_spec() -> List[Tuple[str, Tuple[int, ...]]]:
    """(key, shape) of every tensor of the reference checkpoint (252 entries).  No native code involved."""
    spec = []
    for L in network_layers():
        conv, bn = _layer_keys(L["name"])
        if L["name"] == "skeleton":
            spec.append((conv + ".weight", (L["cout"], L["cin"])))
        else:
            spec.append((conv + ".weight", (L["cout"], L["cin"], L["ksize"], L["ksize"])))
        if _has_bias(L["name"]):
            spec.append((conv + ".bias", (L["cout"],)))
        if bn:
            nbn = 4 if L["name"] == "skeleton" else L["cout"]
            for s in ("weight", "bias", "running_mean", "running_var"):
                spec.append((f"{bn}.{s}", (nbn,)))
            spec.append((bn + ".num_batches_tracked", ()))
    return spec


def make_random_state_dict(seed: int = 0) -> Dict[str, torch.Tensor]:
    """Seeded random checkpoint with the reference's key set (the real weights are not
    distributed with the reference: .MISSING_LARGE_BLOBS).  Conv weights use the reference's
    init N(0, sqrt(2/(k*k*cout))) (backbone_resnet.py:119-122); BN statistics are non-trivial.
    The residual-branch BN gains and the output layers are scaled so that activations stay O(1)
    and the decoded pose is plausible (joint angles ~0.3 rad, wrist ~0.35 m in front of cam0):
    errors measured on it are then meaningful in mm / degrees."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for key, shape in state_dict_spec():
        leaf = key.rsplit(".", 1)[1]
        if leaf == "num_batches_tracked":
            sd[key] = torch.tensor(100, dtype=torch.long)
        elif leaf in ("running_var",) or (leaf == "weight" and len(shape) == 1):
            sd[key] = torch.rand(shape, generator=g) + 0.5
            if leaf == "weight" and (".bn2." in key or ".downsample.1." in key):
                sd[key] *= 0.5  # keep the residual sums from growing through 12 blocks
        elif leaf in ("running_mean", "bias"):
            sd[key] = torch.randn(shape, generator=g) * 0.1
        else:
            fan = shape[0] * (shape[2] * shape[3] if len(shape) == 4 else 1)
            sd[key] = torch.randn(shape, generator=g) * math.sqrt(2.0 / fan)
    sd[_BB + ".1.weight"] *= 0.25
    # rigid 7-point template (regressor.py:19-47) under a fixed plausible wrist pose
    q = 0.1 / math.sqrt(2.0)
    tmpl = torch.tensor([[0, 0, 0], [0.1, 0, 0], [0, 0.1, 0], [0, 0, 0.1], [-q, -q, 0], [-q, 0, -q], [0, -q, -q]])
    ax = torch.tensor([1.0, 2.0, 3.0]) / math.sqrt(14.0) * 0.5
    K = torch.tensor([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    R0 = torch.matrix_exp(K)
    pts = (tmpl @ R0.T + torch.tensor([0.05, -0.03, 0.35])).reshape(-1)
    for h, n_out in (("k", 62), ("u", 63)):
        base = f"_regressor_{h}._pose_regression_layers.2"
        sd[base + ".weight"] *= 0.3
        b = torch.randn(n_out, generator=g) * 0.3
        b[20:41] = pts + torch.randn(21, generator=g) * 0.005
        if h == "u":
            b[41] = 0.05
        sd[base + ".bias"] = b
    return sd


def pack_state_dict(sd: Dict[str, torch.Tensor]) -> torch.Tensor:
    """Fold BN and pack the float32 blob (CPU tensor of at_model_blob_floats() floats)."""
    lib = _cabi.lib()
    expected = {k for k, _ in state_dict_spec()}
    missing = expected - set(sd.keys())
    unexpected = set(sd.keys()) - expected
    if missing or unexpected:
        raise RuntimeError(
            f"Error(s) in loading state_dict: missing {sorted(missing)[:4]}... ({len(missing)}), "
            f"unexpected {sorted(unexpected)[:4]}... ({len(unexpected)})"
        )
    blob = torch.zeros(lib.at_model_blob_floats(), dtype=torch.float32)
    for L in _cabi.layer_table():
        conv, bn = _layer_keys(L["name"])
        w = sd[conv + ".weight"].detach().to("cpu", torch.float64)
        cout, cin, k = L["cout"], L["cin"], L["ksize"]
        if tuple(w.shape) not in ((cout, cin, k, k), (cout, cin)):
            raise RuntimeError(f"size mismatch for {conv}.weight: {tuple(w.shape)}")
        w = w.reshape(cout, cin, k, k)
        b = sd[conv + ".bias"].detach().to("cpu", torch.float64) if _has_bias(L["name"]) else torch.zeros(cout, dtype=torch.float64)
        if bn:
            gamma = sd[bn + ".weight"].double()
            beta = sd[bn + ".bias"].double()
            mean = sd[bn + ".running_mean"].double()
            var = sd[bn + ".running_var"].double()
            scale = gamma / torch.sqrt(var + BN_EPS)
            if L["name"] == "skeleton":  # BatchNorm2d(4) over View((-1,4,6,6)): channel = out // 36
                rep = cout // scale.numel()
                scale, beta, mean = (t.repeat_interleave(rep) for t in (scale, beta, mean))
            w = w * scale.reshape(-1, 1, 1, 1)
            b = (b - mean) * scale + beta
        coutp = (cout + 3) // 4 * 4
        wp = torch.zeros(k, k, cin, coutp, dtype=torch.float64)
        wp[..., :cout] = w.permute(2, 3, 1, 0)
        blob[L["weight_offset"] : L["weight_offset"] + wp.numel()] = wp.reshape(-1).float()
        blob[L["bias_offset"] : L["bias_offset"] + cout] = b.float()
    return blob


def model_from_state_dict(sd: Dict[str, torch.Tensor], **kwargs) -> UmeTrackModel:
    return UmeTrackModel(pack_state_dict(sd), model_opts=ModelOpts(), **kwargs)


DEFAULT_PRECISION = "fp16"


def load_pretrained_model(model_path: str) -> UmeTrackModel:
    """Reference model_loader.py:53-88: build the model and load the checkpoint at `model_path`.

    The encoder format is chosen without editing a script: environment variable AT_PRECISION =
    "fp16" (default: the tcgen05 encoder, inside north_star's 0.1 mm end-to-end tolerance, and
    with an overflow flag, UmeTrackModel.check_status), "bf16" (same kernels, 8-bit mantissa) or
    "fp32" (CUDA-core reference-grade encoder)."""
    with open(model_path, "rb") as fp:
        sd = torch.load(fp, map_location="cpu")
    return model_from_state_dict(sd, precision=os.environ.get("AT_PRECISION", DEFAULT_PRECISION))

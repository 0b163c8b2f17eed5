"""CPU ORACLE of the network (torch CPU float32, functional) - TEST INFRASTRUCTURE, NOT PRODUCT.

Restates lib/models/* of the reference as pure functions of the reference's flat state dict
(252 keys, model_loader.py:84-87).  Pinned against the imported reference in
oracle/make_golden_model.py and tests/test_oracle_cpu.py; committed outputs of the REFERENCE
itself are in tests/golden/model.npz.  Only tests/, __graft_entry__.smoke() and bench.py's CPU
baseline legs may import this.
"""
import numpy as np
import torch
import torch.nn.functional as F

BB = "_feature_extractor._image_backbone"
EPS = 1e-5


def _bn(sd, p, x):
    return F.batch_norm(x, sd[p + ".running_mean"], sd[p + ".running_var"], sd[p + ".weight"], sd[p + ".bias"], False, 0.0, EPS)


def _basic_block(sd, p, x, stride, down):
    """backbone_resnet.py:56-72."""
    out = F.conv2d(x, sd[p + ".conv1.weight"], None, stride, 1)
    out = F.relu(_bn(sd, p + ".bn1", out))
    out = F.conv2d(out, sd[p + ".conv2.weight"], None, 1, 1)
    out = _bn(sd, p + ".bn2", out)
    res = x
    if down:
        res = _bn(sd, p + ".downsample.1", F.conv2d(x, sd[p + ".downsample.0.weight"], None, stride, 0))
    return F.relu(out + res)


def backbone(sd, images):
    """model_utils.py:107-138 + backbone_resnet.py:75-158: [c,96,96] -> [c,72,6,6]."""
    x = images.unsqueeze(1)
    p = BB + ".0._layers"
    x = F.conv2d(x, sd[p + ".0.0.weight"], sd[p + ".0.0.bias"], 1, 1)
    x = F.max_pool2d(F.relu(_bn(sd, p + ".0.1", x)), 2, 2)
    for li, nb in enumerate([2, 3, 5, 2], start=1):
        for b in range(nb):
            first = li > 1 and b == 0
            x = _basic_block(sd, f"{p}.{li}.{b}", x, 2 if first else 1, first)
    return F.conv2d(x, sd[BB + ".1.weight"], sd[BB + ".1.bias"])


def apply_ftl(xfs, fmaps):
    """model_utils.py:57-104 with ftl_ratio = 1."""
    n = fmaps.shape[0]
    pts = fmaps.reshape(n, 3, -1)
    out = torch.matmul(xfs[:, :3, :3], pts) + xfs[:, :3, 3].unsqueeze(-1)
    return out.reshape(fmaps.shape)


def singlev_xfs(intrinsics):
    """model_utils.py:166-192."""
    S = torch.eye(4).unsqueeze(0).repeat(intrinsics.shape[0], 1, 1)
    S[:, 2, 2] = intrinsics[:, 0, 0] / 200
    return S


def fuse_two_views(sd, feats2, S2, E2):
    """feature_extractor.py:61-141 (useUnscaledAsCanonical=False). feats2 [n,2,72,6,6]."""
    xf0 = E2[:, 0:1]
    to_world = torch.inverse(E2) @ S2
    s0 = torch.inverse(S2[:, 0:1])
    to_canon = s0 @ xf0 @ to_world
    canon = apply_ftl(to_canon.reshape(-1, 4, 4), feats2.flatten(0, 1)).reshape(feats2.shape)
    x = canon.flatten(1, 2)
    p = "_feature_extractor._multi_view_fusion"
    x = F.relu(_bn(sd, p + ".1", F.conv2d(x, sd[p + ".0.weight"], sd[p + ".0.bias"])))
    x = F.relu(_bn(sd, p + ".4", F.conv2d(x, sd[p + ".3.weight"], sd[p + ".3.bias"])))
    x = F.conv2d(x, sd[p + ".6.weight"], sd[p + ".6.bias"])
    return apply_ftl(S2[:, 0], x)


class TemporalState:
    """temporal.py:41-42."""

    def __init__(self):
        self.mem = torch.empty(0)
        self.prev_ext = torch.empty(0)


def temporal_step(sd, state, feats, E0, memory_idx, use_memory):
    """temporal.py:51-139."""
    need = int(memory_idx.max()) + 1
    if len(state.mem) < need:
        mem = torch.zeros(need, 18, 6, 6)
        ext = torch.zeros(need, 4, 4)
        if len(state.mem):
            mem[: len(state.mem)] = state.mem
            ext[: len(state.prev_ext)] = state.prev_ext
        state.mem, state.prev_ext = mem, ext
    use_idx = memory_idx[use_memory]
    if len(use_idx) != len(use_memory):
        z = memory_idx[~use_memory]
        state.mem[z] = 0
        state.prev_ext[z] = 0
    if len(use_idx):
        xf = E0[use_memory].bmm(torch.inverse(state.prev_ext[use_idx]))
        state.mem[use_idx] = apply_ftl(xf, state.mem[use_idx])
    state.prev_ext[memory_idx] = E0
    x = torch.cat([state.mem[memory_idx], feats], 1)
    p = "_temporal._temporal_module"
    x = F.relu(F.conv2d(x, sd[p + ".0.weight"], sd[p + ".0.bias"]))
    x = F.relu(F.conv2d(x, sd[p + ".2.weight"], sd[p + ".2.bias"]))
    x = F.conv2d(x, sd[p + ".4.weight"], sd[p + ".4.bias"])
    state.mem[memory_idx] = x[:, :18]
    return x[:, 18:]


def skeleton_features(sd, axes, rest):
    """skeleton_encoder.py:43-53."""
    x = torch.cat((axes.reshape(-1, 22, 3), rest.reshape(-1, 22, 3)), -1).reshape(-1, 132)
    x = F.linear(x, sd["_skeleton_enc._layers.0.weight"], sd["_skeleton_enc._layers.0.bias"]).view(-1, 4, 6, 6)
    return F.relu(_bn(sd, "_skeleton_enc._layers.2", x))


def rigid_template():
    """regressor.py:19-47."""
    pts = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1], [-1, -1, 0], [-1, 0, -1], [0, -1, -1]], np.float64)
    out = np.empty_like(pts)
    for i, p in enumerate(pts):
        nrm = np.linalg.norm(p)
        out[i] = p if nrm == 0 else p / nrm * 0.1
    return torch.from_numpy(out).float()


def procrustes(a, b):
    """model_utils.py:17-54."""
    n = a.shape[0]
    am, bm = a.mean(1), b.mean(1)
    H = torch.matmul((a - am[:, None]).transpose(1, 2), b - bm[:, None])
    u, _, v = H.svd()
    vut = v @ u.transpose(1, 2)
    w = torch.eye(3).unsqueeze(0).repeat(n, 1, 1)
    w[:, 2, 2] = torch.det(vut)
    xf = torch.eye(4).unsqueeze(0).repeat(n, 1, 1)
    xf[:, :3, :3] = v @ w @ u.transpose(1, 2)
    xf[:, :3, 3] = bm - torch.matmul(xf[:, :3, :3], am.unsqueeze(-1)).squeeze(-1)
    return xf


def pose_head(sd, which, x):
    """regressor.py:163-186 + model_utils.py:195-208. which in 'k','u'."""
    p = f"_regressor_{which}._pose_regression_layers"
    x = _basic_block(sd, p + ".0", x, 1, False)
    x = _basic_block(sd, p + ".1", x, 1, False)
    f = F.conv2d(x, sd[p + ".2.weight"], sd[p + ".2.bias"]).mean((2, 3))
    n = f.shape[0]
    out = {"joint_angles": torch.cat([f[:, :20], torch.zeros(n, 2)], 1)}
    out["wrist_xfs"] = procrustes(rigid_template().unsqueeze(0).expand(n, 7, 3).clone(), f[:, 20:41].reshape(n, 7, 3))
    o = 41
    if which == "u":
        out["skel_scales"] = torch.exp(f[:, 41])
        o = 42
    out["sigmas"] = torch.clamp(F.softplus(f[:, o : o + 21]), min=1e-5)
    return out


def forward(sd, state, images, intrinsics, extrinsics, sample_range, memory_idx, use_memory, hand_idx, skel=None):
    """umetrack_model.py:123-242.  skel = (axes, rest) -> known-skeleton head, None -> scale head.
    Returns dict(joint_angles, wrist_xfs (world), sigmas[, skel_scales], features)."""
    with torch.no_grad():
        feats = backbone(sd, images.float())
        S = singlev_xfs(intrinsics)
        fused = []
        for r0, r1 in sample_range.tolist():
            if r1 - r0 == 1:
                fused.append(apply_ftl(S[r0:r1], feats[r0:r1]))
            else:
                fused.append(fuse_two_views(sd, feats[r0:r1].unsqueeze(0), S[r0:r1].unsqueeze(0), extrinsics[r0:r1].unsqueeze(0)))
        fused = torch.cat(fused, 0)
        E0 = extrinsics[sample_range[:, 0]]
        t = temporal_step(sd, state, fused, E0, memory_idx, use_memory)
        if skel is not None:
            sk = skeleton_features(sd, skel[0].float(), skel[1].float())
            if sk.shape[0] == 1 and t.shape[0] > 1:
                sk = sk.expand(t.shape[0], *sk.shape[1:])
            out = pose_head(sd, "k", torch.cat([t, sk], 1))
        else:
            out = pose_head(sd, "u", t)
        w = torch.inverse(E0) @ out["wrist_xfs"]
        w[hand_idx == 1, :, 0] *= -1
        out["wrist_xfs"] = w
        out["features"] = feats
    return out

"""Parity gate on the EXACT configuration bench.py times: BatchedHandTracker with the
fp16 tcgen05 encoder, the fp32 lattice crop map and pinned-host frames pulled on the side stream, S = 256
lock-step sequences x T = 8 steps with carried temporal memory = 4,096 hand-frames / 8,192 crops, on noise
and smooth frames, against the CPU oracle on the same inputs (oracle/parity.py)."""
import pytest

import parity

pytestmark = pytest.mark.gpu


def test_timed_path_meets_north_star_tolerances(cuda_lib):
    r = parity.timed_path_parity(S=256, T=8, precision="fp16", map_mode="fp32", ingest="pull")
    print({k: v for k, v in r.items() if k not in ("path", "against")})
    assert r["n_hand_frames"] >= 1024
    assert r["crop_pixels_differing"] == 0 and r["crop_max_abs_diff"] == 0     # bit-identical to tracker._warp_image
    assert r["crops_checked_against_numpy_warp"] >= 64
    assert r["d_mpjpe_mm"] <= 0.1 and r["d_theta_deg"] <= 0.05                 # north_star, means over hand-frames
    assert r["d_mpjpe_mm_max"] <= 1.0 and r["d_theta_deg_max"] <= 0.5          # no outlier hand-frame
    assert r["pass"]


def test_fp32_encoder_on_the_same_path_is_accumulation_noise(cuda_lib):
    r = parity.timed_path_parity(S=32, T=4, precision="fp32", map_mode="fp32", ingest="pull", seq_offset=40)
    print({k: v for k, v in r.items() if k not in ("path", "against")})
    assert r["crop_pixels_differing"] == 0
    assert r["d_mpjpe_mm_max"] <= 0.05 and r["d_theta_deg_max"] <= 0.02

"""The loss oracle (oracle.c orc_l1_ssim) against the reference's own utils/loss_utils.py outputs and autograd
gradients (tests/golden/data/loss_*.npz, made by tests/golden/make_loss_golden.py)."""
import glob
import os

import numpy as np
import pytest

DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "data")
CASES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(DATA, "loss_*.npz")) if "batch" not in p)


def test_fixtures_present():
    assert len(CASES) == 5


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_loss(oracle, name):
    d = np.load(os.path.join(DATA, name + ".npz"))
    lam = float(d["lambda_dssim"])
    out3, g = oracle.l1_ssim(d["img"], d["gt"], lam)
    # the reference works in float32 (sigma = E[x^2] - mu^2 cancels ~3 digits); the oracle in float64
    np.testing.assert_allclose(out3[0], float(d["l1"]), rtol=2e-6)
    np.testing.assert_allclose(out3[1], float(d["ssim"]), rtol=2e-5)
    np.testing.assert_allclose(out3[2], float(d["loss"]), rtol=2e-5)
    ref = d["grad_loss"]
    assert np.abs(g - ref).max() <= 1e-4 * np.abs(ref).max()
    # lambda = 1 isolates the SSIM term: d(1 - SSIM) = -dSSIM
    _, g1 = oracle.l1_ssim(d["img"], d["gt"], 1.0)
    ref = d["grad_ssim"]
    assert np.abs(-g1 - ref).max() <= 1e-4 * np.abs(ref).max()


def test_oracle_identical_images(oracle):
    d = np.load(os.path.join(DATA, "loss_smooth_64x64.npz"))
    out3, g = oracle.l1_ssim(d["gt"], d["gt"], 0.2)
    assert out3[0] == 0.0 and abs(out3[1] - 1.0) < 1e-12 and abs(out3[2]) < 1e-12
    assert np.abs(g).max() < 1e-9  # SSIM is at its maximum, |x - y| at its kink (sign(0) = 0)


def test_oracle_batch_size_average_false(oracle):
    d = np.load(os.path.join(DATA, "loss_batch_2x3x20x24.npz"))
    for b in range(2):
        out3, _ = oracle.l1_ssim(d["img"][b], d["gt"][b], 1.0, grad=False)
        np.testing.assert_allclose(out3[1], d["ssim"][b], rtol=2e-5)

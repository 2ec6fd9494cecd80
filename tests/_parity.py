"""Shared comparison helpers for the GPU parity tests."""
import numpy as np

from raytracingrust_b200 import _abi as A


def assert_hits_identical(dev, ora, what, allow_t_ulp=0):
    """Level 1 (north_star): closest-hit ids exact; t / point / normal bit-identical (the device
    arithmetic is contraction-free and order-identical, so no tolerance is needed)."""
    bad = np.flatnonzero((dev["object"] != ora["object"]) | (dev["primitive"] != ora["primitive"]))
    assert len(bad) == 0, f"{what}: {len(bad)} id mismatches, first {bad[:8]}: dev {dev[bad[:4]]} ora {ora[bad[:4]]}"
    hit = ora["object"] != A.RTR_NO_HIT
    for f in ("t", "point", "normal"):
        a, b = dev[f][hit], ora[f][hit]
        if allow_t_ulp == 0:
            same = (a == b) | (np.isnan(a) & np.isnan(b))
            assert same.all(), f"{what}: {f} not bit-identical in {np.count_nonzero(~same)} values, max |d| {np.nanmax(np.abs(a - b))}"
        else:  # libm-dependent path (ln in Volume::hit): a few ulp of the vector's magnitude
            a2, b2 = a.reshape(len(a), -1).astype(np.float64), b.reshape(len(b), -1).astype(np.float64)
            err = np.linalg.norm(a2 - b2, axis=1)
            tol = allow_t_ulp * 1.2e-7 * (np.linalg.norm(b2, axis=1) + 1.0)
            assert (err <= tol).all(), f"{what}: {f} off by up to {(err / tol).max():.2f} x tolerance"
    assert np.array_equal(dev["front_face"][hit], ora["front_face"][hit]), what
    assert np.array_equal(dev["material"][hit], ora["material"][hit]), what
    return int(hit.sum())


def assert_images_close(dev_rgba, dev_acc, ora_rgba, ora_acc, what, rtol=1e-5, frac_pixels_off=0.0):
    """Level 3 against the oracle with the SAME Philox stream: accumulation within 1e-5 relative (the device
    multiplies throughput front-to-back, the reference back-to-front: few-ulp), RGBA8 within +-1."""
    assert np.array_equal(dev_acc[..., 3], ora_acc[..., 3]), f"{what}: sample counts differ"
    close = np.isclose(dev_acc[..., :3], ora_acc[..., :3], rtol=rtol, atol=1e-6).all(-1)
    off = 1.0 - close.mean()
    assert off <= frac_pixels_off, f"{what}: {off:.2e} of pixels outside rtol={rtol} (allowed {frac_pixels_off:.1e})"
    d = np.abs(dev_rgba.astype(int) - ora_rgba.astype(int))[close]
    assert d.size == 0 or d.max() <= 1, f"{what}: RGBA8 differs by {d.max()}"

"""CPU tests (-m "not gpu"): the oracle against the committed golden vectors.

The goldens are outputs of the reference's own third-party entry points (cv2.phaseCorrelate,
cv2.logPolar; oracle/gen_golden.py) or, for FastSpacedBM (no CPU implementation in the
reference), an independent brute-force numpy evaluation.  Tolerances: FFT shifts 0.01 px,
scale 1e-3, rotation 1e-3 rad (BASELINE.json north_star); observed agreement is far tighter and
is asserted at 2e-3 px / 1e-4 so that a regression of the oracle shows up.
"""
import os

import numpy as np
import pytest

from helpers import ill_conditioned_sections
from mrs_optic_flow_b200 import synth

FFT_TOL_PX = 0.01
ORACLE_SELF_TOL_PX = 2e-3


@pytest.fixture(scope="module")
def fft_golden(golden_dir):
    return np.load(os.path.join(golden_dir, "fft_cv2.npz"))


@pytest.mark.parametrize("key", ["F240_N60", "F240_N120", "F240_N240", "F60_N60", "F120_N120"])
def test_fft_pairs_vs_cv2_golden(oracle_mod, fft_golden, key):
    N = int(key.split("_N")[1])
    curr, prev, exp = fft_golden[key + "_curr"], fft_golden[key + "_prev"], fft_golden[key + "_cv2"]
    got = oracle_mod.fft_pairs(curr, prev, N, max_px_speed=1e9)
    assert got.shape == exp.shape
    err = np.abs(got - exp).max(axis=2)
    bad = ill_conditioned_sections(curr, prev, N)
    assert np.nanmax(err) < FFT_TOL_PX, np.nanmax(err)
    assert np.nanmax(err[~bad]) < ORACLE_SELF_TOL_PX, np.nanmax(err[~bad])


def test_fft_known_answers(oracle_mod, fft_golden):
    curr, prev = fft_golden["ka_curr"], fft_golden["ka_prev"]
    got = oracle_mod.fft_pairs(curr, prev, 120, max_px_speed=80.0).reshape(3, 2)
    exp = fft_golden["ka_cv2"].reshape(3, 2)
    # identical frames -> (0, 0); integer circular shift (3, -2) -> (3, -2); 50 px -> valid at 80
    assert np.allclose(got[0], 0.0, atol=1e-5)
    assert np.allclose(got[1], [3.0, -2.0], atol=1e-4)
    assert np.allclose(got, exp, atol=1e-4)
    # max_px_speed = 20 -> the 50 px shift becomes NaN (src/FftMethod.cpp:1841-1853)
    got20 = oracle_mod.fft_pairs(curr, prev, 120, max_px_speed=20.0).reshape(3, 2)
    assert np.isnan(got20[2]).all() and np.isnan(fft_golden["ka_cv2_speed20"].reshape(3, 2)[2]).all()
    assert np.allclose(got20[:2], exp[:2], atol=1e-4)


def test_fft_geometry_fixups(oracle_mod):
    # odd frame size -> F-1; non-divisible section size -> one window (src/FftMethod.cpp:1706-1716)
    curr, prev, _ = synth.shifted_pairs(2, 120, 120, seed=5)
    pad_c = np.zeros((2, 121, 121), np.uint8)
    pad_p = np.zeros((2, 121, 121), np.uint8)
    pad_c[:, :120, :120], pad_p[:, :120, :120] = curr, prev
    a = oracle_mod.fft_pairs(curr, prev, 120)
    b = oracle_mod.fft_pairs(pad_c, pad_p, 120)
    assert np.allclose(a, b, atol=1e-9)
    c = oracle_mod.fft_pairs(curr, prev, 50)  # 120 % 50 != 0 -> N := 120
    assert c.shape == (2, 1, 2) and np.allclose(a, c, atol=1e-9)


def test_fft_section_order_and_multithread(oracle_mod):
    curr, prev, true = synth.shifted_pairs(3, 240, 60, seed=11)
    one = oracle_mod.fft_pairs(curr, prev, 60, nthreads=1)
    many = oracle_mod.fft_pairs(curr, prev, 60, nthreads=3)
    assert np.array_equal(one, many)
    # index i + j*sq: section (i=1, j=2) is the tile at x 60..120, y 120..180
    s = oracle_mod.fft_pairs(curr[:, 120:180, 60:120], prev[:, 120:180, 60:120], 60)
    assert np.allclose(one[:, 1 + 2 * 4], s[:, 0], atol=1e-9)
    # the centroid estimator is biased (SURVEY "things to know" 6) but tracks the true shift
    assert np.nanmax(np.abs(one - true)) < 0.5


def test_hann_window_matches_cv2_formula(oracle_mod):
    w = oracle_mod.hann(60)
    c = 0.5 * (1 - np.cos(2 * np.pi * np.arange(60) / 59))
    assert np.allclose(w, np.sqrt(np.outer(c, c)), atol=1e-6)


def test_bm_vs_bruteforce_golden(oracle_mod, golden_dir):
    g = np.load(os.path.join(golden_dir, "bm_bruteforce.npz"))
    for name in ("small", "default", "grid480"):
        F, B, step, R = [int(v) for v in g[name + "_geom"]]
        curr, prev = g[name + "_curr"], g[name + "_prev"]
        for p in range(curr.shape[0]):
            fx, fy, voted, sad = oracle_mod.bm(curr[p], prev[p], B, step, R, want_sad=True)
            assert np.array_equal(fx, g[name + "_fx"][p])
            assert np.array_equal(fy, g[name + "_fy"][p])
            assert np.array_equal(voted, g[name + "_voted"][p])
            if name == "small":
                assert np.array_equal(sad, g[name + "_sad"][p])
        flow, voted = oracle_mod.bm_pairs(curr, prev, B, step, R, nthreads=2)
        assert np.array_equal(flow[:, 0], g[name + "_fx"]) and np.array_equal(flow[:, 1], g[name + "_fy"])
        assert np.array_equal(voted, g[name + "_voted"])


def test_bm_recovers_integer_shift_and_noise_gate(oracle_mod, golden_dir):
    g = np.load(os.path.join(golden_dir, "bm_bruteforce.npz"))
    # default geometry: the voted vector equals the injected integer shift
    assert np.array_equal(g["default_voted"].astype(np.int32), g["default_true"])
    # flat pair -> SAD differences below 0.2 R^2 -> (0, 0) (FastSpacedBMMethod.cl:77-82)
    fx, fy, voted = oracle_mod.bm(g["small_curr"][3], g["small_prev"][3], 16, 8, 8)
    assert not fx.any() and not fy.any() and not voted.any()


def test_logpolar_and_scalerot_vs_cv2_golden(oracle_mod, golden_dir):
    g = np.load(os.path.join(golden_dir, "scalerot_cv2.npz"))
    res, M = int(g["res"]), float(g["M"])
    for p in range(g["curr"].shape[0]):
        lp = oracle_mod.logpolar(g["prev"][p], M, 2)
        assert np.array_equal(lp, g["lp_cubic"][p])  # bit-exact uint8
        lp = oracle_mod.logpolar(g["curr"][p], M, 4, dst=lp)
        assert np.array_equal(lp, g["lp_lanczos"][p])
    got = oracle_mod.scalerot_pairs(g["curr"], g["prev"], M, nthreads=2)
    assert np.abs(got[:, 0] - g["cv2"][:, 0]).max() < 1e-4  # north-star tolerance 1e-3
    assert np.abs(got[:, 1] - g["cv2"][:, 1]).max() < 1e-4  # north-star tolerance 1e-3 rad
    # sign convention: magnified by s -> ~1/s ; rotation sign flipped
    true = g["true"]
    assert np.abs(got[:, 0] - 1.0 / true[:, 0]).max() < 0.02
    assert np.abs(got[:, 1] + true[:, 1]).max() < 0.02


def test_live_cv2_crosscheck_if_available(oracle_mod):
    cv2 = pytest.importorskip("cv2")
    cv2.setNumThreads(1)
    curr, prev, _ = synth.shifted_pairs(6, 120, 120, seed=21)
    for p in range(6):
        a, b = curr[p].astype(np.float32), prev[p].astype(np.float32)
        (cx, cy), _ = cv2.phaseCorrelate(a, b)
        (ox, oy), _ = oracle_mod.phase_correlate(a, b)
        assert abs(cx - ox) < ORACLE_SELF_TOL_PX and abs(cy - oy) < ORACLE_SELF_TOL_PX


def test_cv2_reference_arm_matches_oracle(oracle_mod):
    """oracle/cv2_ref.py (what bench.py --impl reference times: cv2.phaseCorrelate per section with the
    reference's conversions and NaN rules, worker processes) against the C restatement."""
    from oracle import cv2_ref

    if not cv2_ref.available():
        pytest.skip("cv2 not importable")
    from helpers import assert_fft_parity

    for F, N in ((240, 240), (240, 120), (120, 60)):
        curr, prev, _ = synth.shifted_pairs(6, F, N, seed=77 + N)
        prev[1] = 0                                     # all-tie surface -> (-N/2, -N/2) -> NaN at 40 px/frame for N >= 60
        a = cv2_ref.fft_pairs(curr, prev, N, max_px_speed=40.0, nproc=2)
        b = oracle_mod.fft_pairs(curr, prev, N, max_px_speed=40.0)
        assert_fft_parity(oracle_mod, a, b, curr, prev, N, max_px_speed=40.0, max_outlier_frac=0.05)


def test_expand_bank_keeps_sections_paired():
    from helpers import expand_bank

    c, p, _ = synth.shifted_pairs(3, 120, 60, seed=9)
    out = expand_bank(np.stack([c, p], axis=1), 7, 60, seed=1)
    assert out.shape == (7, 2, 120, 120)
    # every section of item 4 (bank item 1) is a section of bank item 1, same tile in both frames
    tiles = lambda a: {a[j * 60:(j + 1) * 60, i * 60:(i + 1) * 60].tobytes() for j in range(2) for i in range(2)}  # noqa: E731
    assert len(tiles(out[4, 0]) & tiles(c[1])) >= 2 and len(tiles(out[4, 1]) & tiles(p[1])) >= 2
    assert not np.array_equal(out[1], out[4])


@pytest.mark.parametrize("kind", ["fft60", "fft120", "fft240"])
def test_oracle_vs_full_size_cv2_golden(oracle_mod, golden_dir, kind):
    """SURVEY 8d sizes: 256 frame pairs per section size (tie guard 1e-4), outputs of the reference's own
    cv2.phaseCorrelate call sequence."""
    from helpers import assert_fft_parity, full_golden

    N = int(kind[3:])
    curr, prev, exp = full_golden(golden_dir, kind)
    got = oracle_mod.fft_pairs(curr, prev, N, max_px_speed=80.0, nthreads=os.cpu_count() or 1)
    assert_fft_parity(oracle_mod, got, exp, curr, prev, N, max_px_speed=80.0)
    err = np.abs(got - exp)
    assert np.nanpercentile(err, 99) < ORACLE_SELF_TOL_PX


def test_oracle_vs_full_size_bm_and_scalerot_goldens(oracle_mod, golden_dir):
    from helpers import full_golden

    curr, prev, exp = full_golden(golden_dir, "bm")
    flow, voted = oracle_mod.bm_pairs(curr, prev, 120, 24, 21, nthreads=os.cpu_count() or 1)
    assert np.array_equal(flow[:, :, 0, 0], exp) and np.array_equal(voted, exp.astype(np.float32))
    curr, prev, exp = full_golden(golden_dir, "scalerot")
    got = oracle_mod.scalerot_pairs(curr, prev, 49.9, nthreads=os.cpu_count() or 1)
    off = (np.abs(got[:, 0] - exp[:, 0]) >= 1e-3) | (np.abs(got[:, 1] - exp[:, 1]) >= 1e-3)
    assert off.sum() <= 1, (off.sum(), np.abs(got - exp).max(axis=0))

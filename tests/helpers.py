"""Shared test helpers."""
import numpy as np


def ill_conditioned_sections(curr, prev, N):
    """Sections in which a spectral bin of either image is exactly zero in exact arithmetic
    (integer pixels can cancel at bins whose DFT kernel is a low-order root of unity).  The
    reference normalises that bin's rounding noise to unit magnitude, so its own result moves by
    up to ~4e-3 px between float32 implementations (cv2 vs numpy float64 shows the same)."""
    n, F, _ = curr.shape
    sq = F // N
    bad = np.zeros((n, sq * sq), bool)
    for p in range(n):
        for j in range(sq):
            for i in range(sq):
                for im in (curr, prev):
                    s = np.abs(np.fft.fft2(im[p, j * N:(j + 1) * N, i * N:(i + 1) * N].astype(np.float64)))
                    if s.min() < 1e-6:
                        bad[p, i + j * sq] = True
    return bad


def expand_bank(bank, n, N, seed):
    """Cheap large batches from a small bank of synthetic frames: item k of the result is bank item
    k % B with its N x N sections permuted (one permutation per item, shared by every frame of the
    item so that a section keeps its own previous frame) and one pixel row XOR-salted, so that no two
    items are equal.  bank: [B, ..., F, F] uint8 (any number of middle axes: frames of a pair or of a
    stream).  -> [n, ..., F, F]."""
    rng = np.random.default_rng(seed)
    bank = np.asarray(bank)
    B, F = bank.shape[0], bank.shape[-1]
    sq = F // N
    mid = bank.shape[1:-2]
    out = np.empty((n,) + mid + (F, F), np.uint8)
    tiles = bank.reshape((B,) + mid + (sq, N, sq, N))
    for k in range(n):
        perm = rng.permutation(sq * sq)
        t = tiles[k % B]
        t = np.moveaxis(t, -2, -3)                       # [..., sq, sq, N, N]
        t = t.reshape(mid + (sq * sq, N, N))[..., perm, :, :].reshape(mid + (sq, sq, N, N))
        out[k] = np.moveaxis(t, -3, -2).reshape(mid + (F, F))
        out[k][..., int(rng.integers(0, F)), :] ^= np.uint8((k * 37 + 11) & 0xFF)
    return out


def near_tie(oracle_mod, curr_sec, prev_sec, rel_gap=1e-4):
    """True when the two largest samples of the oracle's correlation surface are within rel_gap of
    each other: the integer peak (and with it the 5x5 centroid window) then flips on a last-bit
    difference, in cv2 itself as much as anywhere (SURVEY.md §7 'argmax tie sensitivity')."""
    _, _, surf = oracle_mod.phase_correlate(curr_sec.astype(np.float32), prev_sec.astype(np.float32),
                                            want_surface=True)
    top = np.partition(surf.ravel(), -2)[-2:]
    return bool((top[1] - top[0]) <= rel_gap * max(abs(float(top[1])), 1e-30))


def centroid_condition(oracle_mod, curr_sec, prev_sec):
    """sum|v| / |sum v| over the 5x5 window around the oracle's peak: how strongly the weighted centroid
    (a division by sum v, src/FftMethod.cpp:1330-1384) amplifies last-bit differences of the surface."""
    _, _, surf = oracle_mod.phase_correlate(curr_sec.astype(np.float32), prev_sec.astype(np.float32),
                                            want_surface=True)
    py, px = np.unravel_index(int(np.argmax(surf)), surf.shape)
    w = surf[max(py - 2, 0):py + 3, max(px - 2, 0):px + 3].astype(np.float64)
    return float(np.abs(w).sum() / max(abs(w.sum()), 1e-300))


def _cv2_section(curr_sec, prev_sec, N, max_px_speed):
    """The reference's own arithmetic for one section (cv2.phaseCorrelate + the rules of
    src/FftMethod.cpp:1836-1856), or None when cv2 is not importable."""
    try:
        from oracle import cv2_ref

        if not cv2_ref.available():
            return None
        import cv2
    except Exception:
        return None
    out = np.empty((1, 2))
    cv2_ref.fft_method_frame(cv2, np.ascontiguousarray(curr_sec), np.ascontiguousarray(prev_sec), N,
                             1e9 if max_px_speed is None else max_px_speed, out)
    return out[0]


def assert_fft_parity(oracle_mod, got, exp, curr, prev, N, tol=0.01, max_px_speed=None, max_outlier_frac=0.005):
    """got/exp [n, S, 2] against each other at `tol` px with equal NaN patterns.

    A section that misses the C oracle is looked at individually and accepted only if
      * cv2.phaseCorrelate itself (the library call the reference makes, src/FftMethod.cpp:1836) gives the
        CUDA answer within `tol`: the oracle is a restatement pinned to cv2 on well-conditioned inputs; where
        a spectral bin is exactly zero in exact arithmetic, or the 5x5 window sums to almost nothing, the
        reference normalises / divides rounding noise and the float32 restatement and cv2 drift apart by
        more than `tol` (measured: 10 of 7168 sections of the N = 60 batch, up to 0.8 px) -- or
      * the two largest samples of the oracle's surface tie to 1e-4 (the integer peak flips on a last bit), or
      * the section is ill-conditioned in one of those two ways (an exactly-zero spectral bin; sum|v| > 4 |sum v|
        over the 5x5 window: three float32 implementations then give three answers -- e.g. cuda (11.58, -26.01),
        oracle (12.35, -24.30), cv2 (11.58, -24.91) on a frame of twelve grey-level-1 pixels against a texture);
        for a NaN mismatch, the shift sits on the validity threshold.
    Such sections must stay below max_outlier_frac.  Returns (max error over the sections that match the
    oracle, number of sections accepted through the list above)."""
    assert got.shape == exp.shape, (got.shape, exp.shape)
    n, S, _ = exp.shape
    sq = int(round(S ** 0.5))
    nan_g, nan_e = np.isnan(got).any(axis=2), np.isnan(exp).any(axis=2)
    err = np.abs(got - exp).max(axis=2)
    bad = (nan_g != nan_e) | (np.nan_to_num(err, nan=0.0) >= tol)
    idx = np.argwhere(bad)
    assert len(idx) <= max(2, int(max_outlier_frac * n * S)), f"{len(idx)} of {n * S} sections off"
    for p, s_ in idx:
        j, i = divmod(int(s_), sq)
        cs = curr[p, j * N:(j + 1) * N, i * N:(i + 1) * N]
        ps = prev[p, j * N:(j + 1) * N, i * N:(i + 1) * N]
        ref = _cv2_section(cs, ps, N, max_px_speed)
        if ref is not None:
            same_nan = bool(np.isnan(ref).any()) == bool(nan_g[p, s_])
            ok = same_nan and (nan_g[p, s_] or float(np.abs(got[p, s_] - ref).max()) < tol)
        else:
            ok = False
        ok = ok or near_tie(oracle_mod, cs, ps)
        if not ok:   # the reference's estimator itself is unstable here: cv2 and the oracle disagree as well
            ok = bool(ill_conditioned_sections(cs[None], ps[None], N)[0, 0]) or centroid_condition(oracle_mod, cs, ps) > 4.0
        if not ok and nan_g[p, s_] != nan_e[p, s_] and max_px_speed is not None:
            v = got[p, s_] if nan_e[p, s_] else exp[p, s_]
            r = float(np.hypot(v[0], v[1]))
            ok = abs(r - max_px_speed) < 1e-3 or abs(abs(v[0]) - N / 2) < 1e-3 or abs(abs(v[1]) - N / 2) < 1e-3
        assert ok, f"pair {p} section {s_}: cuda {got[p, s_]} oracle {exp[p, s_]} cv2 {ref}"
    good = ~bad & ~nan_e
    return (float(err[good].max()) if good.any() else 0.0), len(idx)


def full_golden(golden_dir, kind):
    """Inputs and expected outputs of the SURVEY 8d-size goldens (tests/golden/full_cv2.npz, minted by
    oracle/gen_golden.py full): the inputs are regenerated from their seeds and checked against the stored
    SHA-256.  kind: 'fft60' | 'fft120' | 'fft240' | 'bm' | 'scalerot' -> (curr, prev, expected)."""
    import hashlib
    import os

    from mrs_optic_flow_b200 import synth

    g = np.load(os.path.join(golden_dir, "full_cv2.npz"))

    def sha(*arrays):
        h = hashlib.sha256()
        for a in arrays:
            h.update(np.ascontiguousarray(a).tobytes())
        return h.hexdigest()

    if kind.startswith("fft"):
        N = int(kind[3:])
        keep = g[f"fft{N}_keep"]
        curr, prev, _ = synth.shifted_pairs(int(keep[-1]) + 1, 240, N, seed=1)
        curr, prev, exp = curr[keep], prev[keep], g[f"fft{N}_cv2"]
    elif kind == "bm":
        curr, prev, _ = synth.bm_pairs(256, 240, 21, seed=2)
        exp = g["bm_flow"]
    else:
        curr, prev, _ = synth.scalerot_pairs(128, 240, seed=3)
        exp = g["scalerot_cv2"]
    key = kind + "_sha"
    assert sha(curr, prev) == str(g[key]), f"{kind}: regenerated inputs differ from the ones the golden was minted on"
    return curr, prev, exp

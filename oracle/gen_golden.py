"""Mint the golden vectors under tests/golden/ (run once, in the build container).

The reference ships no tests or fixtures ("parity unpinned").  Its CPU path is
a loop over the third-party call cv::phaseCorrelate (src/FftMethod.cpp:1836) and, for scale /
rotation, cv::logPolar + cv::phaseCorrelate (src/scaleRotationEstimator.cpp:45,112,117).  The
same library entry points are importable here through Python (cv2 4.13.0), so the goldens are the
outputs of those entry points on seeded synthetic inputs; cv2 may be absent on the GPU box, hence
the committed .npz files.  FastSpacedBM has no CPU implementation in the reference at all
(OpenCL only, and dead code), so its golden is an independent brute-force numpy evaluation of the
kernel's formula (src/FastSpacedBMMethod.cl:28-35, 50-57, 66-73, 77-82).

Usage:  python oracle/gen_golden.py [fft|bm|scalerot|frontend|blockmethod|refine|full ...]   (writes tests/golden/*.npz)
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import cv2  # noqa: E402

from mrs_optic_flow_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def cv2_fft_method(curr: np.ndarray, prev: np.ndarray, N: int, max_px_speed: float) -> np.ndarray:
    """FftMethod::processImage, CPU branch, through the real cv::phaseCorrelate."""
    F = curr.shape[0]
    sq = F // N
    cf, pf = curr.astype(np.float32), prev.astype(np.float32)  # convertTo, :1805-1806
    out = np.zeros((sq * sq, 2))
    for j in range(sq):
        for i in range(sq):
            a = np.ascontiguousarray(cf[j * N:(j + 1) * N, i * N:(i + 1) * N])
            b = np.ascontiguousarray(pf[j * N:(j + 1) * N, i * N:(i + 1) * N])
            (x, y), _ = cv2.phaseCorrelate(a, b)
            sx, sy = -x, -y  # :1836
            ok = not (sx * sx + sy * sy > max_px_speed ** 2 or abs(sx) > N / 2 or abs(sy) > N / 2
                      or np.isnan(sx) or np.isnan(sy))  # :1841-1849
            out[i + j * sq] = (sx, sy) if ok else (np.nan, np.nan)
    return out


def peak_gap(curr: np.ndarray, prev: np.ndarray) -> float:
    """Relative gap between the two largest correlation samples (tie guard)."""
    F1 = np.fft.fft2(curr.astype(np.float64))
    F2 = np.fft.fft2(prev.astype(np.float64))
    P = F1 * np.conj(F2)
    c = np.fft.ifft2(P / (np.abs(P) + 1e-30)).real.ravel()
    top = np.partition(c, -2)[-2:]
    return float((top[1] - top[0]) / abs(top[1]))


def gen_fft():
    data = {}
    for N, F, n_pairs in ((60, 240, 3), (120, 240, 6), (240, 240, 10), (60, 60, 24), (120, 120, 8)):
        curr, prev, true = synth.shifted_pairs(n_pairs, F, N, seed=100 + N + F)
        exp = np.stack([cv2_fft_method(curr[p], prev[p], N, 1e9) for p in range(n_pairs)])
        sq = F // N
        gaps = np.array([[peak_gap(curr[p, j * N:(j + 1) * N, i * N:(i + 1) * N],
                                   prev[p, j * N:(j + 1) * N, i * N:(i + 1) * N])
                          for j in range(sq) for i in range(sq)] for p in range(n_pairs)])
        # reorder gaps to section index i + j*sq (loop above is already j-major)
        key = f"F{F}_N{N}"
        data[key + "_curr"], data[key + "_prev"] = curr, prev
        data[key + "_true"], data[key + "_cv2"], data[key + "_gap"] = true, exp, gaps
        print(key, "pairs", n_pairs, "min gap %.2e" % gaps.min(),
              "max |cv2-true| %.3f" % np.nanmax(np.abs(exp - true)))
    # known-answer cases: identical frames, integer circular shift, NaN limits
    rng = np.random.default_rng(7)
    t = synth.to_u8(synth.texture(rng, 120, 120))
    ka_prev = np.stack([t, t, t])
    ka_curr = np.stack([t, np.roll(t, (-2, 3), axis=(0, 1)), np.roll(t, (0, 50), axis=(0, 1))])
    data["ka_curr"], data["ka_prev"] = ka_curr, ka_prev
    data["ka_cv2"] = np.stack([cv2_fft_method(ka_curr[p], ka_prev[p], 120, 80.0) for p in range(3)])
    data["ka_cv2_speed20"] = np.stack(
        [cv2_fft_method(ka_curr[p], ka_prev[p], 120, 20.0) for p in range(3)])
    print("known answers:", data["ka_cv2"].reshape(3, 2), data["ka_cv2_speed20"].reshape(3, 2))
    np.savez_compressed(os.path.join(OUT, "fft_cv2.npz"), **data)


def brute_bm(curr, prev, B, step, R):
    H, W = curr.shape
    D = 2 * R + 1
    pitch = B + step
    gx, gy = (W - 2 * R) // pitch, (H - 2 * R) // pitch
    fx = np.zeros((gy, gx), np.int8)
    fy = np.zeros((gy, gx), np.int8)
    sad = np.zeros((gy, gx, D, D), np.int64)
    c = curr.astype(np.int64)
    p = prev.astype(np.int64)
    for by in range(gy):
        for bx in range(gx):
            x0, y0 = bx * pitch, by * pitch
            blk = c[y0 + R:y0 + R + B, x0 + R:x0 + R + B]
            for sy in range(D):
                for sx in range(D):
                    sad[by, bx, sy, sx] = np.abs(blk - p[y0 + sy:y0 + sy + B, x0 + sx:x0 + sx + B]).sum()
            k = int(np.argmin(sad[by, bx]))  # first minimum, row-major (sy outer, sx inner)
            my, mx = divmod(k, D)
            vx, vy = mx - R, my - R
            if sad[by, bx, R, R] - sad[by, bx].min() <= R * R * 0.2:
                vx, vy = 0, 0
            fx[by, bx], fy[by, bx] = vx, vy

    def mode(a):
        h = np.bincount(a.ravel().astype(np.int64) + R, minlength=D)
        return int(np.argmax(h)) - R  # argmax -> lowest bin among ties

    return fx, fy, np.array([mode(fx), mode(fy)], np.float32), sad.astype(np.int32)


def gen_bm():
    data = {}
    # small-block variant exercising the grid (B=16, step=8, R=8) and the default geometry
    for name, F, B, step, R, n in (("small", 96, 16, 8, 8, 4), ("default", 240, 120, 24, 21, 2),
                                   ("grid480", 480, 120, 24, 21, 1)):
        curr, prev, true = synth.bm_pairs(n, F, R, seed=200 + F)
        if name == "small":  # a flat pair to trip the noise gate, and a tie case
            curr[3] = 90
            prev[3] = 91
        fx, fy, vt, sads = [], [], [], []
        for p in range(n):
            a, b, v, s = brute_bm(curr[p], prev[p], B, step, R)
            fx.append(a); fy.append(b); vt.append(v); sads.append(s)
        data[name + "_curr"], data[name + "_prev"], data[name + "_true"] = curr, prev, true
        data[name + "_fx"], data[name + "_fy"] = np.stack(fx), np.stack(fy)
        data[name + "_voted"] = np.stack(vt)
        if name == "small":
            data[name + "_sad"] = np.stack(sads)
        data[name + "_geom"] = np.array([F, B, step, R], np.int32)
        print("bm", name, "voted", np.stack(vt).tolist(), "true", true.tolist())
    np.savez_compressed(os.path.join(OUT, "bm_bruteforce.npz"), **data)


def gen_scalerot():
    res, M = 240, 49.9
    curr, prev, true = synth.scalerot_pairs(6, res, seed=300)
    lp_cubic, lp_lanczos, out = [], [], []
    for p in range(curr.shape[0]):
        # first call (scaleRotationEstimator.cpp:27,45,48): tempIm zeros, INTER_CUBIC
        temp = np.zeros((res, res), np.uint8)
        temp = cv2.logPolar(prev[p], (res // 2, res // 2), M, cv2.INTER_CUBIC, dst=temp)
        lp_cubic.append(temp.copy())
        prev_f = temp.astype(np.float32)
        # later call (:112-117): same tempIm, INTER_LANCZOS4
        temp = cv2.logPolar(curr[p], (res // 2, res // 2), M, cv2.INTER_LANCZOS4, dst=temp)
        lp_lanczos.append(temp.copy())
        cur_f = temp.astype(np.float32)
        (x, y), _ = cv2.phaseCorrelate(cur_f, prev_f)
        if abs(x) > res // 2:
            out.append((1.0, 0.0))
        else:
            out.append((np.exp(x / M), (y / (res / 360.0)) * (np.pi / 180.0)))
    out = np.array(out)
    print("scalerot cv2:", out.tolist())
    print("scalerot true (scale, angle):", true.tolist())
    np.savez_compressed(os.path.join(OUT, "scalerot_cv2.npz"), curr=curr, prev=prev, true=true,
                        lp_cubic=np.stack(lp_cubic), lp_lanczos=np.stack(lp_lanczos), cv2=out,
                        res=res, M=M)


def full_inputs(kind: str):
    """Deterministic inputs of the SURVEY 8d-size goldens (regenerated by the tests; the .npz keeps their
    SHA-256 so that a drifting generator is noticed rather than silently compared)."""
    if kind.startswith("fft"):
        N = int(kind[3:])
        curr, prev, _ = synth.shifted_pairs(FULL_CANDIDATES, 240, N, seed=1)   # cfg1: F = 240, seed 1
        return curr, prev
    if kind == "bm":
        curr, prev, _ = synth.bm_pairs(256, 240, 21, seed=2)                    # cfg2: seed 2
        return curr, prev
    if kind == "scalerot":
        curr, prev, _ = synth.scalerot_pairs(128, 240, seed=3)                  # cfg3: seed 3
        return curr, prev
    raise ValueError(kind)


FULL_CANDIDATES = 420   # candidates per FFT size; the first 256 whose every section passes the tie guard are kept


def sha(*arrays) -> str:
    import hashlib

    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def gen_full():
    """Goldens at the sizes SURVEY 8d names: cfg1 256 pairs at N = 60 / 120 / 240 (tie guard: top-1/top-2 gap
    of every section >= 1e-4), cfg2 256 pairs of FastSpacedBM (brute-force numpy), cfg3 128 scale/rotation
    pairs (cv2.logPolar + cv2.phaseCorrelate).  Only outputs, kept indices and input hashes are stored."""
    from numpy.lib.stride_tricks import sliding_window_view

    data = {}
    for N in (60, 120, 240):
        curr, prev = full_inputs(f"fft{N}")
        sq = 240 // N
        keep = []
        for p in range(curr.shape[0]):
            g = min(peak_gap(curr[p, j * N:(j + 1) * N, i * N:(i + 1) * N], prev[p, j * N:(j + 1) * N, i * N:(i + 1) * N])
                    for j in range(sq) for i in range(sq))
            if g >= 1e-4:
                keep.append(p)
            if len(keep) == 256:
                break
        assert len(keep) == 256, (N, len(keep))
        keep = np.array(keep, np.int32)
        data[f"fft{N}_keep"] = keep
        data[f"fft{N}_cv2"] = np.stack([cv2_fft_method(curr[p], prev[p], N, 80.0) for p in keep])
        data[f"fft{N}_sha"] = np.array(sha(curr[keep], prev[keep]))
        print("full fft", N, "kept 256 of", int(keep[-1]) + 1)
    curr, prev = full_inputs("bm")
    B, R, D = 120, 21, 43
    flow = np.zeros((256, 2), np.int8)
    for p in range(256):
        blk = curr[p, R:R + B, R:R + B].astype(np.int32)
        win = sliding_window_view(prev[p, :B + 2 * R, :B + 2 * R].astype(np.int32), (B, B))   # [sy, sx, B, B]
        sad = np.abs(win - blk).sum(axis=(2, 3))
        k = int(np.argmin(sad))
        my, mx = divmod(k, D)
        vx, vy = mx - R, my - R
        if sad[R, R] - sad.min() <= R * R * 0.2:
            vx, vy = 0, 0
        flow[p] = (vx, vy)
    data["bm_flow"] = flow          # one block per pair at F = 240: the vote equals the block's flow
    data["bm_sha"] = np.array(sha(curr, prev))
    print("full bm: moved", int((flow != 0).any(axis=1).sum()), "of 256")
    curr, prev = full_inputs("scalerot")
    res, M = 240, 49.9
    out = []
    for p in range(128):
        temp = np.zeros((res, res), np.uint8)
        temp = cv2.logPolar(prev[p], (res // 2, res // 2), M, cv2.INTER_CUBIC, dst=temp)
        prev_f = temp.astype(np.float32)
        temp = cv2.logPolar(curr[p], (res // 2, res // 2), M, cv2.INTER_LANCZOS4, dst=temp)
        (x, y), _ = cv2.phaseCorrelate(temp.astype(np.float32), prev_f)
        out.append((1.0, 0.0) if abs(x) > res // 2 else (np.exp(x / M), (y / (res / 360.0)) * (np.pi / 180.0)))
    data["scalerot_cv2"] = np.array(out)
    data["scalerot_sha"] = np.array(sha(curr, prev))
    np.savez_compressed(os.path.join(OUT, "full_cv2.npz"), **data)


def gen_frontend():
    """Front end (src/optic_flow.cpp:1603-1622) and long-range mode (src/FftMethod.cpp:1905-1990)
    through the real cv2 entry points: cv2.resize, crop, cv2.cvtColor(COLOR_RGB2GRAY),
    cv2.phaseCorrelate."""
    rng = np.random.default_rng(400)
    data = {}
    W, H, F = 200, 128, 24
    tex = np.stack([synth.to_u8(synth.texture(rng, H, W)) for _ in range(3)], axis=-1)  # smooth BGR
    noise = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)                              # all rounding cases
    raw = np.stack([tex, noise])
    data["raw"] = raw
    for k, cx in ((1, 101), (2, 47), (4, 25)):
        g = []
        for im in raw:
            sc = cv2.resize(im, (W // k, H // k)) if k != 1 else im.copy()   # :1603-1607
            xi, yi = cx - F // 2, sc.shape[0] // 2 - F // 2                    # :1611-1614
            g.append(cv2.cvtColor(np.ascontiguousarray(sc[yi:yi + F, xi:xi + F]), cv2.COLOR_RGB2GRAY))
        data[f"gray_k{k}_cx{cx}"] = np.stack(g)
    mono = raw[..., 1].copy()
    data["mono_k2_cx50"] = np.stack([cv2.resize(im, (W // 2, H // 2))[32 - 12:32 + 12, 50 - 12:50 + 12] for im in mono])
    # long range: F = 240, N = 60 -> sqNum 4 -> sqNum_lr 1, frames down-scaled to 60 x 60
    curr, prev, true = synth.shifted_pairs(3, 240, 240, seed=401)
    out = []
    for p in range(curr.shape[0]):
        c4 = cv2.resize(curr[p], None, fx=0.25, fy=0.25)   # :1931
        p4 = cv2.resize(prev[p], None, fx=0.25, fy=0.25)   # :1932
        out.append(cv2_fft_method(c4, p4, 60, 80.0))
    data["lr_curr"], data["lr_prev"], data["lr_true"] = curr, prev, true
    data["lr_cv2"] = np.stack(out)
    print("long range cv2:", data["lr_cv2"].reshape(-1, 2).tolist(), "true/4:", (true.reshape(-1, 2) / 4).tolist())
    np.savez_compressed(os.path.join(OUT, "frontend_cv2.npz"), **data)


def gen_blockmethod():
    """BlockMethod::processImage integer stage (src/BlockMethod.cpp:25-80) with the real cv2 calls the
    reference makes: cv2.absdiff + cv2.sumElems per candidate, cv2.minMaxLoc per block, histogram
    argmax (std::max_element -> first maximum)."""
    F, B, R = 96, 16, 8
    D = 2 * R + 1
    curr, prev, true = synth.bm_pairs(3, F, R, seed=500)
    curr[2] = 77
    prev[2] = 78   # flat pair: every candidate ties -> first minimum (-R, -R)
    side = (F - 2 * R) // B
    fxs, fys, voted = [], [], []
    for p in range(curr.shape[0]):
        xh, yh = np.zeros(D, int), np.zeros(D, int)
        fx, fy = np.zeros((side, side), np.int8), np.zeros((side, side), np.int8)
        for m in range(side):
            for n in range(side):
                sx, sy = n * B + R, m * B + R
                ad = np.zeros((D, D), np.int32)
                blk = curr[p, sy:sy + B, sx:sx + B]
                for i in range(-R, R + 1):
                    for j in range(-R, R + 1):
                        d = cv2.absdiff(blk, prev[p, sy + i:sy + i + B, sx + j:sx + j + B])
                        ad[R + i, R + j] = int(cv2.sumElems(d)[0])
                _, _, min_loc, _ = cv2.minMaxLoc(ad)
                xh[min_loc[0]] += 1
                yh[min_loc[1]] += 1
                fx[m, n], fy[m, n] = min_loc[0] - R, min_loc[1] - R
        fxs.append(fx); fys.append(fy)
        voted.append((int(np.argmax(xh)) - R, int(np.argmax(yh)) - R))
    print("blockmethod voted", voted, "true", true.tolist())
    np.savez_compressed(os.path.join(OUT, "blockmethod_cv2.npz"), curr=curr, prev=prev, true=true,
                        fx=np.stack(fxs), fy=np.stack(fys), voted=np.array(voted, np.float32),
                        geom=np.array([F, B, R], np.int32))


def ref_refine_cv2(im_curr, im_prev, fullpix, passes=2):
    """BlockMethod::Refine (src/BlockMethod.cpp:96-147) statement by statement with the cv2 calls the
    reference makes (cv2.resize into imPrev2x FROM imCurr2x included)."""
    cur2 = im_curr.copy()
    prev2 = im_prev.copy()
    ox, oy = int(fullpix[0]), int(fullpix[1])
    scale = 1
    h, w = im_curr.shape
    for _ in range(passes):
        scale *= 2
        ox *= 2
        oy *= 2
        prev2 = cv2.resize(cur2, (im_prev.shape[1] * 2, im_prev.shape[0] * 2))
        cur2 = cv2.resize(cur2, (w * 2, h * 2))
        sx = -ox + 1 if ox < 0 else 1
        sy = -oy + 1 if oy < 0 else 1
        cw, ch = cur2.shape[1] - (abs(ox) + 2), cur2.shape[0] - (abs(oy) + 2)
        ad = np.zeros((3, 3), np.int32)
        for m in (-1, 0, 1):
            for n in (-1, 0, 1):
                d = cv2.absdiff(cur2[1:1 + ch, 1:1 + cw], prev2[sy + m:sy + m + ch, sx + n:sx + n + cw])
                ad[1 + m, 1 + n] = int(cv2.sumElems(d)[0])
        _, _, min_loc, _ = cv2.minMaxLoc(ad)
        ox += min_loc[0] - 1
        oy += min_loc[1] - 1
    return np.float32(ox) / np.float32(scale), np.float32(oy) / np.float32(scale)


def gen_refine():
    """cv2.resize 2x (8-bit, INTER_LINEAR) on random and textured frames, and BlockMethod::Refine end
    to end on frames with integer flows of every sign combination."""
    rng = np.random.default_rng(900)
    ups_in = [rng.integers(0, 256, (37, 53), dtype=np.uint8), rng.integers(0, 256, (2, 2), dtype=np.uint8),
              synth.to_u8(synth.texture(rng, 96, 96))]
    ups_out = [cv2.resize(a, (a.shape[1] * 2, a.shape[0] * 2)) for a in ups_in]
    F = 96
    n = 10
    frames = np.stack([synth.to_u8(synth.texture(rng, F, F)) for _ in range(n)])
    frames[8] = 0          # flat: every candidate ties -> first minimum (-1, -1) per pass
    frames[9] = rng.integers(0, 256, (F, F), dtype=np.uint8)
    full = np.array([[0, 0], [3, 2], [-3, 2], [3, -2], [-3, -2], [-8, -8], [7, 0], [0, -5], [1, 1], [-2, 4]],
                    np.int32)
    ref = np.array([ref_refine_cv2(frames[k], frames[k], full[k]) for k in range(n)], np.float64)
    print("refine", ref.tolist())
    d = {f"up_in{k}": a for k, a in enumerate(ups_in)}
    d.update({f"up_out{k}": a for k, a in enumerate(ups_out)})
    np.savez_compressed(os.path.join(OUT, "refine_cv2.npz"), frames=frames, full=full, refined=ref, **d)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    cv2.setNumThreads(1)
    print("cv2", cv2.__version__)
    only = sys.argv[1:]
    for name, fn in (("fft", gen_fft), ("bm", gen_bm), ("scalerot", gen_scalerot), ("frontend", gen_frontend),
                     ("blockmethod", gen_blockmethod), ("refine", gen_refine), ("full", gen_full)):
        if not only or name in only:
            fn()

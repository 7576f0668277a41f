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

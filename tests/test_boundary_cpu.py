"""CPU tests (-m "not gpu") of the boundary and the host logic: the C-ABI library loads and exports
every symbol include/mof.h declares, it refuses to run without a CUDA device, and the kernel's
work-item code -- executed thread-by-thread on the CPU by tests/emu -- reproduces the oracle."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from mrs_optic_flow_b200 import _lib, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(_lib.LIB_PATH):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "mrs_optic_flow_b200", "csrc"), "-j4", "-s"])
    return _lib.load()


def test_library_exports_every_declared_symbol(lib):
    hdr = open(os.path.join(ROOT, "include", "mof.h")).read()
    declared = set(re.findall(r"\b(mof_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name
    assert b"sm_100a" in lib.mof_version()


def test_config_struct_matches_header(lib):
    cfg = _lib.MofConfig()
    lib.mof_default_config(C.byref(cfg))
    # config/default.yaml:13,28-32,63
    assert (cfg.frame_size, cfg.sample_point_size, cfg.max_px_speed) == (480, 120, 80.0)
    assert (cfg.bm_block, cfg.bm_step, cfg.bm_radius) == (120, 24, 21)
    assert (cfg.sr_resolution, cfg.sr_m, cfg.hann, cfg.n_streams) == (240, 49.9, 0, 0)


def test_no_cpu_fallback_without_device(lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    cfg = _lib.MofConfig()
    lib.mof_default_config(C.byref(cfg))
    ctx = C.c_void_p()
    rc = lib.mof_create(C.byref(ctx), C.byref(cfg))
    assert rc == _lib.MOF_ERR_CUDA and not ctx.value
    assert b"no CPU fallback" in lib.mof_last_error(None)
    from mrs_optic_flow_b200.api import Backend, MofError

    with pytest.raises(MofError):
        Backend(240, 120, 80.0)


def test_null_arguments_are_rejected(lib):
    assert lib.mof_create(None, None) == _lib.MOF_ERR_INVALID
    assert lib.mof_sync(None) == _lib.MOF_ERR_INVALID
    assert lib.mof_fft_pairs(None, None, None, 0, 0, 0, None) == _lib.MOF_ERR_INVALID
    assert lib.mof_kernel_launches(None) == 0
    lib.mof_destroy(None)  # no-op


# ---- the kernel's item functions on the CPU ----------------------------------------------------
@pytest.fixture(scope="module")
def emu():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "tests", "emu"), "-s"])
    L = C.CDLL(os.path.join(ROOT, "tests", "emu", "libemu_pfa.so"))
    L.emu_check_butterflies.restype = C.c_double
    L.emu_pfa.restype = C.c_int

    def run(N, curr, prev, win=None, max_sq=1e18, order=0, stream=0):
        out = np.zeros(2)
        surf = np.zeros((N, N), np.float32)
        curr, prev = np.ascontiguousarray(curr), np.ascontiguousarray(prev)
        rc = L.emu_pfa(N, curr.ctypes.data_as(C.c_void_p), prev.ctypes.data_as(C.c_void_p),
                      curr.strides[0], None if win is None else win.ctypes.data_as(C.c_void_p),
                      C.c_double(max_sq), order, out.ctypes.data_as(C.c_void_p),
                      surf.ctypes.data_as(C.c_void_p), stream)
        assert rc == 0, rc
        return out, surf

    run.lib = L
    return run


def test_register_butterflies_match_naive_dft(emu):
    assert emu.lib.emu_check_butterflies() < 5e-6


@pytest.mark.parametrize("N", [60, 120, 240])
def test_pipeline_items_match_oracle(emu, oracle_mod, N):
    n = 6 if N < 240 else 3
    curr, prev, _ = synth.shifted_pairs(n, N, N, seed=3)
    for p in range(n):
        (ox, oy), _, osurf = oracle_mod.phase_correlate(
            curr[p].astype(np.float32), prev[p].astype(np.float32), want_surface=True)
        ref = None
        for order in (0, 1, 2):  # item execution order must not matter (no intra-phase hazards)
            out, surf = emu(N, curr[p], prev[p], order=order, stream=int(order == 2))
            assert abs(out[0] + ox) < 0.01 and abs(out[1] + oy) < 0.01
            assert np.abs(surf - osurf).max() < 4.0  # a single ill-conditioned bin moves it by <= 2
            if ref is None:
                ref = (out, surf)
            else:
                assert np.array_equal(ref[0], out) and np.array_equal(ref[1], surf)


def test_pipeline_items_window_and_validity(emu, oracle_mod):
    curr, prev, _ = synth.shifted_pairs(2, 120, 120, seed=9)
    w = oracle_mod.hann(120)
    for p in range(2):
        (ox, oy), _ = oracle_mod.phase_correlate(curr[p].astype(np.float32),
                                                 prev[p].astype(np.float32), window=w)
        out, _ = emu(120, curr[p], prev[p], win=w)
        assert abs(out[0] + ox) < 2e-3 and abs(out[1] + oy) < 2e-3
    out, _ = emu(120, curr[0], prev[0], max_sq=1e-6)  # max_px_speed ~ 0 -> NaN
    assert np.isnan(out).all()
    out, _ = emu(120, curr[0], curr[0])  # identical frames -> zero shift
    assert np.allclose(out, 0.0, atol=1e-5)


def test_multi_create_fails_loudly_without_gpu():
    """mof_multi_create (several GPUs behind one handle) has no CPU fallback either."""
    import ctypes as C

    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from mrs_optic_flow_b200 import _lib

    L = _lib.load()
    cfg = _lib.MofConfig()
    L.mof_default_config(C.byref(cfg))
    m = C.c_void_p()
    dev = (C.c_int * 2)(0, 1)
    rc = L.mof_multi_create(C.byref(m), C.byref(cfg), dev, 2)
    assert rc == _lib.MOF_ERR_CUDA and not m.value
    assert b"no CPU fallback" in L.mof_multi_last_error(None)


def test_config_struct_layout_matches_header():
    """ctypes mirror of mof_config: ocl_compat sits right after bm_no_gate, total size unchanged."""
    import ctypes as C

    from mrs_optic_flow_b200 import _lib

    names = [f[0] for f in _lib.MofConfig._fields_]
    assert names[-3:] == ["bm_no_gate", "ocl_compat", "reserved"]
    assert C.sizeof(_lib.MofConfig) == 4 * 3 + 4 + 8 + 4 * 7 + 4 + 8 + 4 * 8 or C.sizeof(_lib.MofConfig) % 8 == 0


# ---- generic section sizes (pg_generic.cuh), executed lane by lane on the CPU ---------------------------
@pytest.fixture(scope="module")
def emu_pg():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "tests", "emu"), "-s", "libemu_pg.so"])
    L = C.CDLL(os.path.join(ROOT, "tests", "emu", "libemu_pg.so"))
    L.emu_pg.restype = C.c_int
    L.emu_pg_supported.restype = C.c_int
    return L


def test_generic_plan_covers_even_235_smooth_sizes(emu_pg):
    ok = [n for n in range(2, 800) if emu_pg.emu_pg_supported(n)]
    for n in (30, 80, 96, 160, 480, 768, 36, 100, 144, 256):
        assert n in ok
    for n in (70, 75, 45, 14, 2, 770, 800 - 1):
        assert n not in ok


@pytest.mark.parametrize("N", [30, 80, 96, 160, 480, 24, 36, 40, 64, 100, 128, 250, 300, 450, 512])
def test_generic_pipeline_matches_oracle(emu_pg, oracle_mod, N):
    """Stockham passes (radix 4/2/3/5; the second group of sizes walks their combinations, pure powers of two and
    lines shorter than a warp included), packed two-image transform, cross-power on bin pairs, first maximum
    and centroid of pg_generic.cuh against the oracle, 0.01 px (typ. 1e-6)."""
    from mrs_optic_flow_b200 import synth

    curr, prev, _ = synth.shifted_pairs(4, N, N, seed=600 + N, max_shift=min(N / 4.0, 30.0))
    exp = oracle_mod.fft_pairs(curr, prev, N, max_px_speed=80.0)
    worst = 0.0
    for k in range(4):
        out = (C.c_double * 2)()
        rc = emu_pg.emu_pg(N, curr[k].ctypes.data_as(C.c_void_p), prev[k].ctypes.data_as(C.c_void_p), N,
                           C.c_double(80.0 ** 2), out)
        assert rc == 0
        got = np.array([out[0], out[1]])
        assert np.array_equal(np.isnan(got), np.isnan(exp[k, 0]))
        if not np.isnan(got).any():
            worst = max(worst, float(np.abs(got - exp[k, 0]).max()))
    assert worst < (0.01 if N >= 80 else 0.05), worst

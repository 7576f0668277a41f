#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path (BASELINE.json: "frame-pairs/sec, 240px FFT
sections, 1/2/4/8 B200; % of HBM roofline").

    python bench.py --gpus N --steps K --warmup W            # this backend
    python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU path (cv2.phaseCorrelate, all cores)
    python bench.py --workload streams [--streams 64] ...      # cfg5: camera streams sharded over the ranks

A step = one pass of the batched FftMethod path over one batch of synthetic frame pairs
(F = 240, N = 240: one 240 px section per pair).  `value` is whole-job throughput with the inputs
already resident in HBM (device-pointer entry point of the C ABI); `e2e` is the same metric
through the host-pointer entry point with pinned host buffers (H2D and D2H inside the timed
region).  One process per GPU; with N > 1 the independent pairs are sharded across ranks (no
data-path collective) and the time is the max over ranks.  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

F_HEAD, N_HEAD = 240, 240
FP32_PEAK_TFLOPS_NOMINAL = 148 * 128 * 2 * 1.965e9 / 1e12  # 74.4, non-tensor FP32
# tools/ubench/f32x2.cu on this pool's B200: FADD/FFMA/FADD2/FFMA2 all sustain 128.7 lane-ops/clk/SM
FP32_PEAK_TFLOPS_MEASURED = 148 * 128.7 * 2 * 1.9e9 / 1e12
# ncu --set full of pc_pfa_kernel<16,true,true> (profiles/): shared-memory wavefronts (128 B each) and DRAM
# bytes per frame pair of the headline workload; fixed by the pipeline design, not by the run
SMEM_WAVEFRONTS_PER_PAIR_240 = 37682
NCU_DRAM_BYTES_PER_PAIR_240 = (273.27e6 + 6.91e6) / 2368   # profiles/r2j_ncu_brief_pc240.txt
NCU_PROFILE_FILE = "profiles/r2j_ncu_brief_pc240.txt"


def algo_bytes_per_pair(F: int, N: int) -> int:
    """2*F^2 uint8 in + 16 B (double2) per section out."""
    return 2 * F * F + 16 * (F // N) ** 2


def algo_flops_per_pair(F: int, N: int) -> float:
    """SURVEY §8d convention: per section 15 N^2 log2 N (three 2-D real transforms) + 10 N^2."""
    S = (F // N) ** 2
    return S * (15.0 * N * N * np.log2(N) + 10.0 * N * N)


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.lines: list[str] = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "25",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 8:
                continue
            try:
                sm.append(float(p[1]))
                mx.append(float(p[2]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if p[4 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": float(max(mx)) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def make_inputs_host(n: int, F: int, seed: int):
    from mrs_optic_flow_b200 import synth

    return synth.fast_pairs(n, F, seed)


def device_batch(torch, host_curr, host_prev, n: int, dev):
    """Tile a host bank to n pairs on the device and make every pair distinct."""
    reps = (n + host_curr.shape[0] - 1) // host_curr.shape[0]
    dc = torch.from_numpy(host_curr).to(dev).repeat(reps, 1, 1)[:n].contiguous()
    dp = torch.from_numpy(host_prev).to(dev).repeat(reps, 1, 1)[:n].contiguous()
    salt = (torch.arange(n, device=dev) % 251).to(torch.uint8)
    dc[:, 3, :] ^= salt[:, None]
    dp[:, 7, :] ^= salt[:, None]
    return dc, dp


def timed_device_loop(torch, fn, steps: int, warmup: int, barrier):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    barrier()
    return e0.elapsed_time(e1)  # ms


def parity_sample(torch, dc, dp, res, N, max_px_speed, n_sample):
    """Oracle check of the TIMED output: a random sample of the batch the last timed step wrote is
    recomputed by the CPU oracle (the checker, never the thing measured).  The NaN pattern must agree;
    max |cuda - oracle| over the sample is reported (tolerance of the path: 0.01 px)."""
    from oracle import oracle

    oracle.build()
    n = dc.shape[0]
    idx = np.sort(np.random.default_rng(2024).choice(n, size=min(n_sample, n), replace=False))
    ti = torch.from_numpy(idx).to(dc.device)
    hc, hp = dc[ti].cpu().numpy(), dp[ti].cpu().numpy()
    exp = oracle.fft_pairs(hc, hp, N, max_px_speed=max_px_speed, nthreads=os.cpu_count() or 1)
    got = res[idx]
    nan_equal = bool(np.array_equal(np.isnan(got), np.isnan(exp)))
    err = np.abs(got - exp)
    finite = np.isfinite(err)
    return {"pairs_checked": int(len(idx)), "sections_checked": int(exp.shape[0] * exp.shape[1]),
            "max_err_px": float(err[finite].max()) if finite.any() else None,
            "sections_over_0p01px": int((np.nan_to_num(err, nan=0.0).max(axis=2) > 0.01).sum()),
            "nan_pattern_equal": nan_equal, "checker": "oracle/mof_oracle.c (CPU restatement pinned to cv2 4.13)"}


def side_workload(torch, api, dev_index, dev, name, F, N, n, steps=5, **kw):
    """Short device-resident measurement of another configuration (reported, not the headline)."""
    hc, hp = make_inputs_host(min(n, 64), F, seed=11)
    dc, dp = device_batch(torch, hc, hp, n, dev)
    be = api.Backend(F, N, 80.0, device=dev_index, **kw)
    out = torch.empty((n, be.S, 2), dtype=torch.float64, device=dev)
    ms = timed_device_loop(torch, lambda: be.fft_pairs(dc, dp, out=out), steps, 3, lambda: None)
    be.close()
    pps = n * steps / (ms * 1e-3)
    hbm, _ = measured_peaks()
    return {"workload": name, "pairs_per_s": pps, "ms_per_step": ms / steps, "batch": n,
            "hbm_frac": algo_bytes_per_pair(F, N) * pps / 1e9 / hbm,
            "fp32_frac_nominal": algo_flops_per_pair(F, N) * pps / 1e12 / FP32_PEAK_TFLOPS_NOMINAL}


def run_ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the backend has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # keep stdout to the one JSON line (NCCL_DEBUG=VERSION prints a banner there)
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()

    from mrs_optic_flow_b200 import api

    F, N = F_HEAD, N_HEAD
    # 56 sections per resident CTA (one per SM at N = 240): 8 288 pairs on a 148-SM B200, no partly filled last
    # round of the persistent grid (8 192 pairs, the earlier default, left the 56th round 35 % full: -1.2 %)
    batch = args.batch if args.batch > 0 else 56 * torch.cuda.get_device_properties(local).multi_processor_count
    hc, hp = make_inputs_host(256, F, seed=100 + rank)
    dc, dp = device_batch(torch, hc, hp, batch, dev)
    be = api.Backend(F, N, 80.0, device=local, max_batch=args.e2e_chunk)
    out = torch.empty((batch, be.S, 2), dtype=torch.float64, device=dev)

    def step():
        be.fft_pairs(dc, dp, out=out)

    # ---- device-resident headline ---------------------------------------------------------------
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    be.enable_kernel_timing(True)
    be.fft_kernel_time(reset=True)
    launches0 = be.kernel_launches
    sampler = ClockSampler(local)
    sampler.start()
    ms = timed_device_loop(torch, step, args.steps, 0, barrier)
    clocks = sampler.stop()
    launches = be.kernel_launches - launches0
    kavg_ms, kn = be.fft_kernel_time(reset=True)
    be.enable_kernel_timing(False)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    pairs_total = batch * args.steps * world
    value = pairs_total / (ms_max * 1e-3)

    # sanity: the timed kernel really produced shifts (finite for most pairs)
    res = out.cpu().numpy()
    finite_frac = float(np.isfinite(res).all(axis=(1, 2)).mean())
    parity = None
    if rank == 0 and not args.no_parity:
        parity = parity_sample(torch, dc, dp, res, N, 80.0, args.parity_pairs)

    # ---- end to end through the host-pointer C-ABI call -----------------------------------------
    n_e2e = args.e2e_pairs
    pc = torch.from_numpy(np.ascontiguousarray(np.resize(hc, (n_e2e, F, F)))).pin_memory()
    pp = torch.from_numpy(np.ascontiguousarray(np.resize(hp, (n_e2e, F, F)))).pin_memory()
    po = torch.empty((n_e2e, be.S, 2), dtype=torch.float64).pin_memory()
    for _ in range(3):
        be.fft_pairs(pc, pp, out=po)
    barrier()
    t0 = time.perf_counter()
    e2e_steps = max(2, min(3 * args.steps, 30))   # ~4.5 ms each: long enough to average PCIe jitter
    for _ in range(e2e_steps):
        be.fft_pairs(pc, pp, out=po)  # blocking: returns with the results on the host
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = n_e2e * e2e_steps * world / float(t.item())
    e2e = {"value": e2e_value, "unit": "frame-pairs/s",
           "h2d_bytes_per_step": int(2 * n_e2e * F * F), "d2h_bytes_per_step": int(n_e2e * be.S * 16),
           "pairs_per_step_per_gpu": n_e2e, "steps": e2e_steps,
           "timing": "host clock around blocking mof_fft_pairs calls (each returns after its D2H), max over ranks"}

    hbm_peak, peak_src = measured_peaks()
    bytes_per_launch = algo_bytes_per_pair(F, N) * batch
    achieved = bytes_per_launch / (kavg_ms * 1e-3) / 1e9 if kavg_ms > 0 else 0.0
    roofline = {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                "frac": achieved / hbm_peak,
                "traffic": args.ncu_traffic_bytes if args.ncu_traffic_bytes else NCU_DRAM_BYTES_PER_PAIR_240 * batch,
                "traffic_source": f"constant from {NCU_PROFILE_FILE}: ncu dram__bytes_read+write of one launch, scaled to this batch",
                "kernel": "mof::pc_pfa_kernel<16,true,true>", "kernel_avg_ms": kavg_ms, "launches_timed": kn,
                "algorithmic_bytes_per_launch": bytes_per_launch, "peak_source": peak_src,
                "note": "the fused kernel is FP32/shared-memory bound, not HBM bound (see roofline_fp32)"}
    flops_per_launch = algo_flops_per_pair(F, N) * batch
    ach_tf = flops_per_launch / (kavg_ms * 1e-3) / 1e12 if kavg_ms > 0 else 0.0
    roofline_fp32 = {"bound": "fp32", "achieved": ach_tf, "peak": FP32_PEAK_TFLOPS_NOMINAL,
                     "unit": "TFLOP/s", "frac": ach_tf / FP32_PEAK_TFLOPS_NOMINAL,
                     "peak_source": "nominal 148 SM x 128 lanes x 2 x 1.965 GHz (non-tensor FP32)",
                     "peak_measured": FP32_PEAK_TFLOPS_MEASURED,
                     "flops_convention": "15 N^2 log2 N + 10 N^2 per section (SURVEY §8d)"}
    # the roof that binds in practice: every FFT stage goes through shared memory (128 B/clk/SM)
    sm_clock = (clocks.get("sm_mhz") or 1965.0) * 1e6
    smem_peak = 148 * 128 * sm_clock / 1e12
    smem_ach = SMEM_WAVEFRONTS_PER_PAIR_240 * 128 * batch / (kavg_ms * 1e-3) / 1e12 if kavg_ms > 0 else 0.0
    roofline_smem = {"bound": "shared-memory bandwidth", "achieved": smem_ach, "peak": smem_peak, "unit": "TB/s",
                     "frac": smem_ach / smem_peak,
                     "peak_source": "148 SM x 128 B/clk x SM clock under load",
                     "wavefronts_per_pair": SMEM_WAVEFRONTS_PER_PAIR_240,
                     "wavefronts_source": f"constant from {NCU_PROFILE_FILE} (l1tex__data_pipe_lsu_wavefronts_mem_shared per pair)"}

    line = {
        "metric": "frame_pairs_per_s", "value": value, "unit": "frame-pairs/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": "FftMethod batched, F=240 N=240 (one 240px section per frame pair), "
                               f"{batch} independent pairs per GPU per step",
                   "frame_size": F, "sample_point_size": N, "pairs_per_step_per_gpu": batch,
                   "sharding": f"independent pairs x{world}, no collective",
                   "l2": f"inputs {2 * batch * F * F / 1e6:.0f} MB per step > 126 MB L2 (no flush needed)",
                   "valid_fraction": finite_frac},
        "roofline": roofline, "roofline_fp32": roofline_fp32, "roofline_smem": roofline_smem, "e2e": e2e,
        "gpu_launches": int(launches), "clocks": clocks,
    }
    if parity is not None:
        line["parity_max_err_px"] = parity["max_err_px"]
        line["parity"] = parity

    if rank == 0 and world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(F, N, args.cpu_seconds)
    if rank == 0 and world == 1 and not args.no_extras:
        extras = []
        try:
            extras.append(side_workload(torch, api, local, dev, "FftMethod F=480 N=120 (default.yaml, 4x4 sections)", 480, 120, 2048))
            extras.append(side_workload(torch, api, local, dev, "FftMethod F=480 N=60 (8x8 sections, cfg4: 4096 pairs)", 480, 60, 4096))
            extras.append(side_workload(torch, api, local, dev, "FftMethod F=240 N=60 (4x4 sections)", 240, 60, 8192))
            extras.append(side_workload(torch, api, local, dev, "FftMethod F=480 N=480 (constructor fallback N := frameSize; generic-size kernel)", 480, 480, 1024))
            extras.append(side_workload(torch, api, local, dev, "FftMethod F=480 N=160 (3x3 sections; generic-size kernel)", 480, 160, 1024))
            extras.append(stream_workload(torch, api, local, dev))
            extras.append(stream240_workload(torch, api, local, dev))
            extras.append(stream240_many_workload(torch, api, local, dev))
            extras.append(stream240_e2e_workload(torch, api, local, dev))
            extras.append(raw_stream_workload(torch, api, local, dev))
            extras.append(longrange_workload(torch, api, local, dev))
            extras.append(bm_workload(torch, api, local, dev))
            extras.append(scalerot_workload(torch, api, local, dev))
            extras.append(refine_workload(torch, api, local, dev))
        except Exception as ex:  # extras must never take the headline down
            extras.append({"error": repr(ex)})
        line["other_workloads"] = extras
    be.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        emit(line)


def stream_workload(torch, api, dev_index, dev):
    """cfg5-style: 64 camera streams, 480^2 crop, N=120; stateful push (cached previous spectrum)."""
    n_streams, F, N = 64, 480, 120
    hc, hp = make_inputs_host(64, F, seed=5)
    dc, _ = device_batch(torch, hc, hp, n_streams, dev)
    dn, _ = device_batch(torch, hp, hc, n_streams, dev)
    be = api.Backend(F, N, 80.0, device=dev_index, n_streams=n_streams)
    ids = np.arange(n_streams, dtype=np.int32)
    out = torch.empty((n_streams, be.S, 2), dtype=torch.float64, device=dev)
    be.fft_stream_push(ids, dc, out=out)
    flip = [dn, dc]
    k = [0]

    def step():
        be.fft_stream_push(ids, flip[k[0] & 1], out=out)
        k[0] += 1

    ms = timed_device_loop(torch, step, 20, 3, lambda: None)
    be.close()
    return {"workload": "FftMethod stream push, 64 streams x 480^2, N=120 (one frame per stream per step)",
            "frames_per_s": n_streams * 20 / (ms * 1e-3), "ms_per_step": ms / 20}


def run_streams(args):
    """cfg5 (BASELINE.json configs[4]): 64 camera streams of 752 x 480 BGR8 frames sharded over the
    ranks (stream s -> rank s % world, its spectrum cache stays on that GPU); a step = one new frame
    on every stream: device front end (crop 480^2 at cx = 376, RGB2GRAY) + stateful push, N = 120.
    Raw frames are device resident; `value` = frames/s over all ranks, time = max over ranks."""
    import torch
    import torch.distributed as dist

    from mrs_optic_flow_b200 import api, shard

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the backend has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()

    n_streams, F, N, W, H, T = args.streams, 480, 120, 752, 480, args.frames
    mine = shard.local_streams(n_streams, rank, world)
    n = len(mine)
    g = torch.Generator(device=dev).manual_seed(5 + rank)
    # T frames per stream per step, device resident ([n, T, H, W, 3]: 64 x 64 frames = 4.4 GB on one GPU)
    raws = torch.randint(0, 256, (max(n, 1), T, H, W, 3), dtype=torch.uint8, device=dev, generator=g)
    be = api.Backend(F, N, 80.0, device=local, n_streams=max(n, 1))
    ids = np.arange(n, dtype=np.int32)   # local slot of global stream s is s // world
    out = torch.empty((max(n, 1), T, be.S, 2), dtype=torch.float64, device=dev)

    def step():
        if n:
            if T == 1:
                be.fft_stream_push_raw(ids, raws[:n, 0], 1, 376, out=out[:n, 0])
            else:
                be.fft_stream_push_many_raw(ids, raws[:n], 1, 376, out=out[:n])

    launches0 = be.kernel_launches
    sampler = ClockSampler(local)
    sampler.start()
    ms = timed_device_loop(torch, step, args.steps, max(args.warmup, 3), barrier)
    clocks = sampler.stop()
    launches = be.kernel_launches - launches0
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    # latency of the single-frame push (one frame on every local stream per call), host clock around a
    # call + synchronize: what a camera callback would see
    lat_us = None
    if n:
        for _ in range(5):
            be.fft_stream_push_raw(ids, raws[:n, 0], 1, 376, out=out[:n, 0])
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for k in range(50):
            be.fft_stream_push_raw(ids, raws[:n, k % T], 1, 376, out=out[:n, 0])
            torch.cuda.synchronize()
        lat_us = (time.perf_counter() - t0) / 50 * 1e6
    # e2e: the same call with HOST frames (pinned), H2D of the crop windows and D2H of the shifts inside the
    # timed region; a bounded sequence (Te frames per stream per call) keeps the pinned buffer near 1 GB
    e2e = None
    Te = min(T, 16)
    if n and not args.no_e2e:
        host = torch.empty((n, Te, H, W, 3), dtype=torch.uint8).pin_memory()
        host.copy_(raws[:n, :Te])
        hout = torch.empty((n, Te, be.S, 2), dtype=torch.float64).pin_memory()
        hb = api.Backend(F, N, 80.0, device=local, n_streams=n, max_batch=n * Te)
        for _ in range(2):
            hb.fft_stream_push_many_raw(ids, host.numpy(), 1, 376, out=hout.numpy())
        barrier()
        t0 = time.perf_counter()
        reps = max(3, args.steps)
        for _ in range(reps):
            hb.fft_stream_push_many_raw(ids, host.numpy(), 1, 376, out=hout.numpy())
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        # frames 1.. of a call depend on that call's frames only: compare with the device-resident call
        chk = be.fft_stream_push_many_raw(ids, raws[:n, :Te].contiguous(), 1, 376)
        same = bool(torch.equal(hout[:, 1:].nan_to_num(nan=-1e9), chk[:, 1:].cpu().nan_to_num(nan=-1e9)))
        full = bool(os.environ.get("MOF_FE_FULL_H2D"))
        e2e = {"value": n_streams * Te * reps / float(dt.item()), "unit": "frames/s",
               "h2d_bytes_per_step": int(n * Te * (H * W * 3 if full else F * F * 3)),
               "d2h_bytes_per_step": int(n * Te * be.S * 16),
               "frames_per_stream_per_call": Te,
               "note": "mof_fft_stream_push_many_raw with pinned host frames: only the 480 x 1440-byte crop window "
                       "of each 752x480 BGR8 frame crosses PCIe, chunk i+1 is copied while chunk i runs"}
        e2e["equals_device_resident_result"] = same
        hb.close()
    hbm, peak_src = measured_peaks()
    bytes_per_frame = F * F * 3 + F * F + F * F + 16 * be.S   # crop read + gray written, gray read, shifts out
    frames_total = n_streams * T * args.steps
    line = {"metric": "frames_per_s", "value": frames_total / (ms_max * 1e-3), "unit": "frames/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_max / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"cfg5: {n_streams} camera streams x {T} frames of 752x480 BGR8 per step, crop 480^2, "
                                   f"N=120, streams sharded s % {world}; one mof_fft_stream_push_many_raw call per step",
                       "streams_per_gpu": int(n), "frames_per_stream_per_step": T,
                       "sharding": "stream s -> rank s % world, no collective",
                       "l2": f"raw frames {n * T * H * W * 3 / 1e6:.0f} MB per step per GPU"},
            "roofline": {"bound": "hbm", "achieved": bytes_per_frame * frames_total / (ms_max * 1e-3) / 1e9,
                         "peak": hbm * world, "unit": "GB/s",
                         "frac": bytes_per_frame * frames_total / (ms_max * 1e-3) / 1e9 / (hbm * world),
                         "traffic": None, "peak_source": peak_src,
                         "note": "front end is HBM bound, the correlation FP32/shared-memory bound (N=120)"},
            "single_frame_push_latency_us": lat_us,
            "gpu_launches": int(launches), "clocks": clocks}
    if e2e is not None:
        line["e2e"] = e2e
    be.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        emit(line)


def run_multi(args):
    """Single process, several GPUs behind ONE handle (mof_multi_*: one context + one host worker thread per
    device inside libmof_b200.so, contiguous pair slices, D2H into disjoint slices of the caller's array).
    End to end only -- the multi-device API takes host buffers: `value` == `e2e.value`.  Also reports the
    box's concurrent pinned-H2D ceiling (tools/ubench/h2d_sweep) when that binary is present."""
    import torch

    from mrs_optic_flow_b200 import api

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the backend has no CPU fallback")
    G = min(args.gpus, torch.cuda.device_count())
    F, N = F_HEAD, N_HEAD
    n = args.e2e_pairs * G
    hc, hp = make_inputs_host(256, F, seed=100)
    pc = torch.from_numpy(np.ascontiguousarray(np.resize(hc, (n, F, F)))).pin_memory()
    pp = torch.from_numpy(np.ascontiguousarray(np.resize(hp, (n, F, F)))).pin_memory()
    po = torch.empty((n, 1, 2), dtype=torch.float64).pin_memory()
    mb = api.MultiBackend(list(range(G)), F, N, 80.0, max_batch=args.e2e_chunk)
    for _ in range(max(3, args.warmup)):
        mb.fft_pairs(pc, pp, out=po)
    launches0 = mb.kernel_launches
    steps = max(2, min(3 * args.steps, 30))
    t0 = time.perf_counter()
    for _ in range(steps):
        mb.fft_pairs(pc, pp, out=po)
    dt = time.perf_counter() - t0
    launches = mb.kernel_launches - launches0
    value = n * steps / dt
    one = api.Backend(F, N, 80.0, device=0, max_batch=args.e2e_chunk)
    ref = one.fft_pairs(pc[:512], pp[:512])
    same = bool(np.array_equal(ref, po[:512].numpy(), equal_nan=True))
    one.close()
    mb.close()
    ceiling = None
    exe = os.path.join(ROOT, "tools", "ubench", "h2d_sweep")
    if os.path.exists(exe):
        try:
            out = subprocess.run([exe, "256"], capture_output=True, text=True, timeout=120).stdout
            ceiling = json.loads(out.strip().splitlines()[-1])
        except Exception as ex:
            ceiling = {"error": repr(ex)}
    line = {"metric": "frame_pairs_per_s", "value": value, "unit": "frame-pairs/s", "n_gpus": G, "steps": steps,
            "warmup": max(3, args.warmup), "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"FftMethod batched, F=240 N=240, {args.e2e_pairs} pairs per GPU per call through "
                                   f"mof_multi_fft_pairs (one process, {G} contexts + worker threads in the library)",
                       "bitwise_equal_to_single_context": same},
            "e2e": {"value": value, "unit": "frame-pairs/s", "h2d_bytes_per_step": int(2 * n * F * F),
                    "d2h_bytes_per_step": int(n * 16)},
            "h2d_ceiling": ceiling, "gpu_launches": int(launches)}
    emit(line)


def stream240_workload(torch, api, dev_index, dev):
    """Stateful mode at the headline geometry: 4096 camera streams of 240^2 frames, one 240 px section
    each; every pushed frame is correlated against its stream's cached previous spectrum (one forward
    transform per frame instead of two)."""
    n_streams, F, N = 4096, 240, 240
    hc, hp = make_inputs_host(64, F, seed=8)
    da, db = device_batch(torch, hc, hp, n_streams, dev)
    be = api.Backend(F, N, 80.0, device=dev_index, n_streams=n_streams)
    ids = np.arange(n_streams, dtype=np.int32)
    out = torch.empty((n_streams, be.S, 2), dtype=torch.float64, device=dev)
    be.fft_stream_push(ids, da, out=out)
    flip = [db, da]
    k = [0]

    def step():
        be.fft_stream_push(ids, flip[k[0] & 1], out=out)
        k[0] += 1

    ms = timed_device_loop(torch, step, 6, 3, lambda: None)
    be.close()
    return {"workload": "FftMethod stream mode F=240 N=240, 4096 streams, one frame per stream per step "
                        "(previous frame cached as its spectrum)",
            "frames_per_s": n_streams * 6 / (ms * 1e-3), "ms_per_step": ms / 6}


def stream240_many_workload(torch, api, dev_index, dev):
    """Stateful mode at the headline geometry with T frames per stream per call (mof_fft_stream_push_many):
    one forward transform per frame, the previous spectrum never leaves the chip / L2 between frames."""
    n_streams, T, F, N = 592, 16, 240, 240
    hc, hp = make_inputs_host(64, F, seed=9)
    da, _ = device_batch(torch, hc, hp, n_streams * T, dev)
    frames = da.view(n_streams, T, F, F)
    be = api.Backend(F, N, 80.0, device=dev_index, n_streams=n_streams)
    ids = np.arange(n_streams, dtype=np.int32)
    out = torch.empty((n_streams, T, be.S, 2), dtype=torch.float64, device=dev)
    ms = timed_device_loop(torch, lambda: be.fft_stream_push_many(ids, frames, out=out), 6, 3, lambda: None)
    be.close()
    fps = n_streams * T * 6 / (ms * 1e-3)
    return {"workload": f"FftMethod stream mode F=240 N=240, {n_streams} streams x {T} frames per call (push_many)",
            "frames_per_s": fps, "ms_per_step": ms / 6,
            "fp32_frac_by_the_pair_convention": algo_flops_per_pair(F, N) * fps / 1e12 / FP32_PEAK_TFLOPS_NOMINAL,
            "note": "every frame forms a pair with its predecessor (the reference's stateful processImage); the "
                    "cached spectrum saves one of the three transforms the 15 N^2 log2 N convention counts, so this "
                    "fraction is NOT comparable with roofline_fp32 of the independent-pairs headline: executed "
                    "flops are 2/3 of the convention's"}


def stream240_e2e_workload(torch, api, dev_index, dev):
    """The e2e line with half the PCIe bytes: consecutive frames of a stream cross the bus once
    (mof_fft_stream_push_many with pinned HOST frames; chunk i+1 copied while chunk i runs)."""
    n_streams, T, F, N = 296, 16, 240, 240
    hc, _ = make_inputs_host(64, F, seed=10)
    frames = torch.from_numpy(np.ascontiguousarray(np.resize(hc, (n_streams, T, F, F)))).pin_memory()
    hout = torch.empty((n_streams, T, F // N * (F // N), 2), dtype=torch.float64).pin_memory()
    be = api.Backend(F, N, 80.0, device=dev_index, n_streams=n_streams, max_batch=n_streams * T)
    ids = np.arange(n_streams, dtype=np.int32)
    for _ in range(2):
        be.fft_stream_push_many(ids, frames, out=hout)
    t0 = time.perf_counter()
    reps = 8
    for _ in range(reps):
        be.fft_stream_push_many(ids, frames, out=hout)
    dt = time.perf_counter() - t0
    be.close()
    return {"workload": f"e2e stream mode F=240 N=240, {n_streams} streams x {T} frames per call, pinned host frames "
                        "(mof_fft_stream_push_many: one frame per pair over PCIe)",
            "frames_per_s": n_streams * T * reps / dt, "ms_per_step": dt / reps * 1e3,
            "h2d_bytes_per_step": n_streams * T * F * F, "d2h_bytes_per_step": n_streams * T * 16,
            "h2d_gbs": n_streams * T * F * F * reps / dt / 1e9}


def raw_stream_workload(torch, api, dev_index, dev):
    """cfg5: 64 camera streams of 752 x 480 BGR8 frames -> device front end (crop 480^2 at cx = 376,
    RGB2GRAY) -> stateful push, N = 120.  Raw frames are device resident."""
    n_streams, F, N, W, H = 64, 480, 120, 752, 480
    g = torch.Generator(device=dev).manual_seed(5)
    raws = [torch.randint(0, 256, (n_streams, H, W, 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(2)]
    be = api.Backend(F, N, 80.0, device=dev_index, n_streams=n_streams)
    ids = np.arange(n_streams, dtype=np.int32)
    out = torch.empty((n_streams, be.S, 2), dtype=torch.float64, device=dev)
    k = [0]

    def step():
        be.fft_stream_push_raw(ids, raws[k[0] & 1], 1, 376, out=out)
        k[0] += 1

    ms = timed_device_loop(torch, step, 20, 3, lambda: None)
    gray = torch.empty((n_streams, F, F), dtype=torch.uint8, device=dev)
    ms_fe = timed_device_loop(torch, lambda: be.frontend(raws[0], 1, 376, out=gray), 20, 3, lambda: None)
    be.close()
    fe_bytes = n_streams * (F * F * 3 + F * F)   # crop read once + gray written once
    hbm, _ = measured_peaks()
    return {"workload": "cfg5 raw: 64 streams x 752x480 BGR8 -> front end + stream push, N=120",
            "frames_per_s": n_streams * 20 / (ms * 1e-3), "ms_per_step": ms / 20,
            "frontend_ms_per_step": ms_fe / 20, "frontend_gbs": fe_bytes / (ms_fe / 20 * 1e-3) / 1e9,
            "frontend_hbm_frac": fe_bytes / (ms_fe / 20 * 1e-3) / 1e9 / hbm}


def longrange_workload(torch, api, dev_index, dev):
    n, F, N = 2048, 480, 60
    hc, hp = make_inputs_host(32, F, seed=6)
    dc, dp = device_batch(torch, hc, hp, n, dev)
    be = api.Backend(F, N, 80.0, device=dev_index)
    out = torch.empty((n, be.longrange_geometry()[1], 2), dtype=torch.float64, device=dev)
    ms = timed_device_loop(torch, lambda: be.fft_pairs_longrange(dc, dp, out=out), 5, 3, lambda: None)
    be.close()
    return {"workload": "FftMethod long range, F=480 N=60 (1/4 down-scale, 2x2 sections of 60 px)",
            "pairs_per_s": n * 5 / (ms * 1e-3), "ms_per_step": ms / 5, "batch": n}


def bm_workload(torch, api, dev_index, dev):
    from mrs_optic_flow_b200 import synth

    n, F = 2048, 240
    c, p, _ = synth.bm_pairs(16, F, 21, seed=2)
    dc, dp = device_batch(torch, c, p, n, dev)
    be = api.Backend(F, F, 1e9, device=dev_index, bm_block=120, bm_step=24, bm_radius=21)
    ms = timed_device_loop(torch, lambda: be.bm_pairs(dc, dp), 5, 3, lambda: None)
    be.close()
    blocks = n * 5 / (ms * 1e-3)
    return {"workload": "FastSpacedBM F=240 B=120 step=24 R=21 (1 block/pair)", "pairs_per_s": blocks,
            "ms_per_step": ms / 5, "sad_pixel_ops_per_s": blocks * 120 * 120 * 43 * 43,
            # integer roof: VABSDIFF4 issues at half rate, 4 pixel-ops per lane
            "vabsdiff4_roof_frac": blocks * 120 * 120 * 43 * 43 / (148 * 64 * 4 * 1.965e9)}


def refine_workload(torch, api, dev_index, dev):
    """BlockMethod::Refine (src/BlockMethod.cpp:96-147): 2x up-scale + two 3x3 SAD passes per frame."""
    n, F = 1024, 240
    hc, _ = make_inputs_host(32, F, seed=8)
    dc, _ = device_batch(torch, hc, hc, n, dev)
    full = torch.randint(-21, 22, (n, 2), dtype=torch.int32, device=dev)
    be = api.Backend(F, F, 1e9, device=dev_index, bm_block=120, bm_step=0, bm_radius=21, bm_no_gate=True)
    ms = timed_device_loop(torch, lambda: be.bm_refine(dc, full), 5, 3, lambda: None)
    be.close()
    fps = n * 5 / (ms * 1e-3)
    # per frame: read F^2, write and re-read the 2x image (4 F^2 each way, L2 permitting), 2 x 9 SADs of ~4 F^2
    return {"workload": "BlockMethod::Refine F=240, 2 passes", "frames_per_s": fps, "ms_per_step": ms / 5,
            "sad_pixel_ops_per_s": fps * 2 * 9 * 4 * F * F}


def scalerot_workload(torch, api, dev_index, dev):
    n, res = 1024, 240
    hc, hp = make_inputs_host(32, res, seed=3)
    dc, dp = device_batch(torch, hc, hp, n, dev)
    be = api.Backend(res, res, 1e9, device=dev_index, sr_resolution=res, sr_m=49.9, max_batch=n)
    ms = timed_device_loop(torch, lambda: be.scalerot_pairs(dc, dp), 5, 3, lambda: None)
    be.close()
    return {"workload": "scaleRotationEstimator res=240 M=49.9 (log-polar x2 + phase correlation)",
            "pairs_per_s": n * 5 / (ms * 1e-3), "ms_per_step": ms / 5}


def cpu_baseline(F: int, N: int, seconds: float) -> dict:
    """The reference's CPU path timed on this box's host cores (bounded sample): cv2.phaseCorrelate per
    section exactly as src/FftMethod.cpp:1805-1806, :1829-1856 calls it (oracle/cv2_ref.py), one worker
    process per core.  The figures of the plain-C oracle port are reported beside it; if cv2 cannot be
    imported the port is the baseline and `kind` says so."""
    from oracle import cv2_ref, oracle

    oracle.build()
    cores = os.cpu_count() or 1
    hc, hp = make_inputs_host(64, F, seed=1)
    res = {"unit": "frame-pairs/s", "cores": cores}
    # the oracle port: all threads, then one thread
    t0 = time.perf_counter()
    oracle.fft_pairs(hc[:cores], hp[:cores], N, max_px_speed=80.0, nthreads=cores)
    per_pair_wall = (time.perf_counter() - t0) / max(1, cores)
    n_port = int(max(cores, min(20000, 0.35 * seconds / max(per_pair_wall, 1e-6))))
    c = np.resize(hc, (n_port, F, F))
    p = np.resize(hp, (n_port, F, F))
    t0 = time.perf_counter()
    oracle.fft_pairs(c, p, N, max_px_speed=80.0, nthreads=cores)
    dt_port = time.perf_counter() - t0
    res["port_value"] = n_port / dt_port
    m = max(4, n_port // (8 * cores))
    t0 = time.perf_counter()
    oracle.fft_pairs(c[:m], p[:m], N, max_px_speed=80.0, nthreads=1)
    res["port_value_1core"] = m / (time.perf_counter() - t0)
    if cv2_ref.available():
        pool = cv2_ref.Pool(c[:cores * 8], p[:cores * 8], N, 80.0, cores)
        pool.run()   # fork + first-touch outside the timed part
        t0 = time.perf_counter()
        pool.run()
        rate = cores * 8 / (time.perf_counter() - t0)
        pool.close()
        n = int(max(cores * 8, min(40000, 0.6 * seconds * rate)))
        c2, p2 = np.resize(hc, (n, F, F)), np.resize(hp, (n, F, F))
        pool = cv2_ref.Pool(c2, p2, N, 80.0, cores)
        pool.run()
        t0 = time.perf_counter()
        pool.run()
        dt = time.perf_counter() - t0
        pool.close()
        one = cv2_ref.Pool(c2[:max(8, n // (8 * cores))], p2[:max(8, n // (8 * cores))], N, 80.0, 1)
        import cv2

        cv2.setNumThreads(1)
        one.run()
        t0 = time.perf_counter()
        one.run()
        res["value_1core"] = one.n / (time.perf_counter() - t0)
        one.close()
        res.update({"value": n / dt, "kind": "reference",
                    "via": f"cv2.phaseCorrelate {cv2_ref.version()} per section (the call at src/FftMethod.cpp:1836; "
                           "python binding of the library the reference links), one process per core",
                    "sample": f"{n} pairs of the same workload (F={F}, N={N}), {dt:.1f} s, {cores} worker processes, "
                              "cv2.setNumThreads(1) in each"})
    else:
        res.update({"value": res["port_value"], "kind": "port", "value_1core": res["port_value_1core"],
                    "sample": f"{n_port} pairs of the same workload (F={F}, N={N}), {dt_port:.1f} s, all {cores} host "
                              "threads, one pair per thread (cv2 not importable: oracle port)"})
    return res


def run_reference(args):
    """Reference arm: the reference's own CPU implementation of the path (useOCL=false branch), i.e.
    cv2.phaseCorrelate per section with the conversions and validity rules around it
    (src/FftMethod.cpp:1805-1806, :1829-1856; oracle/cv2_ref.py), one worker process per host core, on a
    bounded sample of the same workload per step.  The reference's C++ translation units cannot be built
    here (ROS + OpenCV C++ headers absent); if cv2 is not importable either, the plain-C oracle port runs."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import cv2_ref, oracle

    F, N = F_HEAD, N_HEAD
    cores = os.cpu_count() or 1
    n = args.ref_pairs if args.ref_pairs > 0 else 512 * cores
    hc, hp = make_inputs_host(64, F, seed=100)
    c = np.resize(hc, (n, F, F))
    p = np.resize(hp, (n, F, F))
    use_cv2 = cv2_ref.available() and not args.ref_port
    if use_cv2:
        pool = cv2_ref.Pool(c, p, N, 80.0, cores)
        step = pool.run
        kind = "reference"
        via = (f"cv2.phaseCorrelate {cv2_ref.version()} per section (src/FftMethod.cpp:1836), {cores} worker "
               "processes, cv2.setNumThreads(1) in each")
    else:
        oracle.build()
        pool = None
        step = lambda: oracle.fft_pairs(c, p, N, max_px_speed=80.0, nthreads=cores)  # noqa: E731
        kind = "port"
        via = f"oracle/mof_oracle.c, {cores} threads (one pair per thread)"
    for _ in range(max(1, args.warmup)):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    if pool:
        pool.close()
    value = n * args.steps / dt
    world = int(os.environ.get("WORLD_SIZE", "1"))
    line = {
        "impl": "reference", "metric": "frame_pairs_per_s", "value": value, "unit": "frame-pairs/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "FftMethod batched, F=240 N=240 (one 240px section per frame pair), "
                               f"bounded sample of {n} pairs per step on the host CPU",
                   "frame_size": F, "sample_point_size": N, "pairs_per_step": n},
        "cpu_baseline": {"value": value, "unit": "frame-pairs/s", "cores": cores, "kind": kind, "via": via,
                         "sample": f"{n} pairs per step, {args.steps} steps"},
        "e2e": {"value": value, "unit": "frame-pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


_JSON_FD = None


def guard_stdout():
    """stdout carries exactly one JSON line: whatever libraries print there (NCCL's version banner at
    communicator creation, ...) is sent to stderr; emit() writes to the real stdout."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_JSON_FD, data)


def main():
    guard_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=0, help="frame pairs per GPU per step (0: 56 per SM = 8288 on a B200)")
    ap.add_argument("--e2e-pairs", type=int, default=2048)
    ap.add_argument("--e2e-chunk", type=int, default=256)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--ref-pairs", type=int, default=0)
    ap.add_argument("--workload", default="pairs", choices=["pairs", "streams", "multi"],
                    help="pairs: the headline batched FftMethod; streams: cfg5 camera streams sharded over the ranks")
    ap.add_argument("--streams", type=int, default=64)
    ap.add_argument("--frames", type=int, default=64, help="streams workload: frames per stream per step")
    ap.add_argument("--ref-port", action="store_true", help="reference arm: time the plain-C oracle port instead of cv2")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="streams workload: skip the host-frame leg")
    ap.add_argument("--parity-pairs", type=int, default=256, help="pairs of the timed output re-checked by the oracle")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--ncu-traffic-bytes", type=float, default=None,
                    help="dram bytes per launch from the committed ncu capture (profiles/), if known")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "streams":
        run_streams(args)
    elif args.workload == "multi":
        run_multi(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

"""N > 1 host logic on CPU: world_size-2 gloo processes shard a bag of pairs and a set of camera
streams exactly like bench.py / production do on GPUs; the per-rank compute is the oracle here
(no GPU in this container), so what is tested is the partitioning, the stream->rank ownership and
the optional result gather -- the sharded result must equal the single-process result bit for bit."""
import os
import sys

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from mrs_optic_flow_b200 import shard, synth  # noqa: E402


class OracleStreams:
    """Stand-in for Backend.fft_stream_push with the reference's imPrev semantics
    (src/FftMethod.cpp:1791-1793, :1872), computing through the oracle."""

    def __init__(self, N):
        from oracle import oracle

        self.o, self.N, self.prev = oracle, N, {}

    def fft_stream_push(self, slots, frames):
        out = []
        for s, f in zip(slots, frames):
            p = self.prev.get(int(s), f)
            out.append(self.o.fft_pairs(f[None], p[None], self.N, max_px_speed=80.0)[0])
            self.prev[int(s)] = f.copy()
        return np.stack(out)


def _worker(rank, world, port, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle

    try:
        curr, prev, _ = synth.shifted_pairs(7, 120, 60, seed=21)  # 7 pairs: ragged over 2 ranks
        full = shard.sharded_fft_pairs(lambda c, p: oracle.fft_pairs(c, p, 60, max_px_speed=80.0),
                                       curr, prev, rank, world, gather=True)
        np.save(os.path.join(tmp, f"pairs_{rank}.npy"), full)
        # 5 streams x 3 frames, stream s owned by rank s % world
        frames, _ = synth.camera_streams(5, 3, 120, 120, seed=22)
        ss = shard.ShardedStreams(OracleStreams(60), 5, rank, world)
        res = np.stack([ss.push(frames[:, t]) for t in range(3)])
        np.save(os.path.join(tmp, f"streams_{rank}.npy"), res)
    finally:
        dist.destroy_process_group()


def test_partition_helpers():
    for n in (0, 1, 7, 64, 4096):
        for w in (1, 2, 4, 8):
            sl = [shard.pair_slice(n, r, w) for r in range(w)]
            assert sl[0].start == 0 and sl[-1].stop == n
            assert all(a.stop == b.start for a, b in zip(sl, sl[1:]))
            assert max(s.stop - s.start for s in sl) - min(s.stop - s.start for s in sl) <= 1
    ids = np.concatenate([shard.local_streams(64, r, 8) for r in range(8)])
    assert sorted(ids.tolist()) == list(range(64))
    assert all(shard.stream_owner(s, 8) == s % 8 and shard.local_slot(s, 8) == s // 8 for s in range(64))
    with pytest.raises(ValueError):
        shard.pair_slice(4, 2, 2)


def test_world2_gloo_matches_single_process(tmp_path, oracle_mod):
    import socket

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    curr, prev, _ = synth.shifted_pairs(7, 120, 60, seed=21)
    exp = oracle_mod.fft_pairs(curr, prev, 60, max_px_speed=80.0)
    for r in range(2):
        got = np.load(tmp_path / f"pairs_{r}.npy")
        assert np.array_equal(got, exp, equal_nan=True)
    frames, _ = synth.camera_streams(5, 3, 120, 120, seed=22)
    one = OracleStreams(60)
    exp_s = np.stack([one.fft_stream_push(np.arange(5), frames[:, t]) for t in range(3)])
    assert np.allclose(exp_s[0], 0.0, atol=1e-5)  # first call of every stream: prev := curr
    for r in range(2):
        got = np.load(tmp_path / f"streams_{r}.npy")
        assert np.array_equal(got, exp_s, equal_nan=True)

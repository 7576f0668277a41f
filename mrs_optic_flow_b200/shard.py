"""Independent-frame-stream sharding across the GPUs of one box (north_star (c)).

The path has no exchange step: every (stream, frame pair, section) is independent and the only
state is each stream's previous frame (src/FftMethod.cpp:1872).  So the sharding is pure
partitioning -- one process per GPU, each with its own backend context -- plus an OPTIONAL gather
of the (tiny) result vectors.  No collective sits on the data path.

  * bags of frame pairs: rank r takes the contiguous slice `pair_slice(n, r, world)`;
  * camera streams:      stream s belongs to rank `s % world` (its previous frame / spectrum stays
                         resident on that GPU for the stream's lifetime).
"""
from __future__ import annotations

import numpy as np


def pair_slice(n_pairs: int, rank: int, world: int) -> slice:
    """Contiguous, balanced slice of a bag of `n_pairs` for `rank` (sizes differ by at most 1)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, rem = divmod(n_pairs, world)
    lo = rank * base + min(rank, rem)
    return slice(lo, lo + base + (1 if rank < rem else 0))


def stream_owner(stream_id: int, world: int) -> int:
    return int(stream_id) % world


def local_streams(n_streams: int, rank: int, world: int) -> np.ndarray:
    """Global ids of the streams `rank` owns, ascending.  Local slot of stream s = s // world."""
    return np.arange(rank, n_streams, world, dtype=np.int32)


def local_slot(stream_id: int, world: int) -> int:
    return int(stream_id) // world


def gather_pairs(local: np.ndarray, n_pairs: int, rank: int, world: int, group=None) -> np.ndarray:
    """Optional result gather for a sharded bag: every rank gets the (n_pairs, ...) array.
    `local` is this rank's slice of results.  Uses torch.distributed (NCCL on GPUs moves the data
    through device tensors, gloo on CPU); with world == 1 it is the identity."""
    if world == 1:
        return local
    import torch
    import torch.distributed as dist

    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    sizes = [pair_slice(n_pairs, r, world) for r in range(world)]
    tail = local.shape[1:]
    mx = max(s.stop - s.start for s in sizes)
    buf = torch.zeros((mx,) + tail, dtype=torch.float64, device=dev)
    buf[: local.shape[0]] = torch.from_numpy(np.ascontiguousarray(local, dtype=np.float64)).to(dev)
    outs = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(outs, buf, group=group)
    full = np.empty((n_pairs,) + tail, dtype=np.float64)
    for r, s in enumerate(sizes):
        full[s] = outs[r][: s.stop - s.start].cpu().numpy()
    return full


def gather_streams(local: np.ndarray, n_streams: int, rank: int, world: int, group=None) -> np.ndarray:
    """Optional gather for stream-sharded results: local[k] belongs to stream rank + k*world."""
    if world == 1:
        return local
    import torch
    import torch.distributed as dist

    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    tail = local.shape[1:]
    mx = (n_streams + world - 1) // world
    buf = torch.zeros((mx,) + tail, dtype=torch.float64, device=dev)
    buf[: local.shape[0]] = torch.from_numpy(np.ascontiguousarray(local, dtype=np.float64)).to(dev)
    outs = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(outs, buf, group=group)
    full = np.empty((n_streams,) + tail, dtype=np.float64)
    for r in range(world):
        ids = local_streams(n_streams, r, world)
        full[ids] = outs[r][: len(ids)].cpu().numpy()
    return full


def sharded_fft_pairs(compute, curr, prev, rank: int, world: int, gather: bool = True, group=None):
    """Run `compute(curr_slice, prev_slice) -> (n_local, S, 2)` on this rank's slice of the bag and
    (optionally) gather.  `compute` is Backend.fft_pairs in production."""
    sl = pair_slice(len(curr), rank, world)
    local = np.asarray(compute(curr[sl], prev[sl]))
    return gather_pairs(local, len(curr), rank, world, group) if gather else local


class ShardedStreams:
    """Stateful multi-camera front: stream s lives on rank s % world; `push(frames)` takes one
    frame per *global* stream (index = stream id), processes the local ones through
    `backend.fft_stream_push` and optionally gathers the per-stream shifts."""

    def __init__(self, backend, n_streams: int, rank: int, world: int, group=None):
        self.be, self.n, self.rank, self.world, self.group = backend, n_streams, rank, world, group
        self.ids = local_streams(n_streams, rank, world)
        self.slots = np.arange(len(self.ids), dtype=np.int32)

    def push(self, frames, gather: bool = True):
        local = np.asarray(self.be.fft_stream_push(self.slots, frames[self.ids]))
        return gather_streams(local, self.n, self.rank, self.world, self.group) if gather else local

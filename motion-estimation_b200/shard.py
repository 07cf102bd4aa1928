"""Frame pairs sharded over the GPUs of one box, one process per GPU (torch.distributed).

The search has no inter-GPU data dependency: rank r takes a contiguous slice of the pair list,
uploads only the frames that slice touches (its frame range plus the one-frame halo a pair needs),
and searches it.  The only exchange is one ``all_gather`` of the per-pair records (32 bytes per
pair) so that every rank -- in particular rank 0, which runs the velocity arithmetic of
estimate_vel_vid.py:540-573 -- sees the whole series.  NCCL over NVLink on GPUs, gloo in CPU tests.

The anchored loop of the video script is sequential (estimate_vel_vid.py:531-534,575) and is NOT
sharded by frame range: recordings are (``shard_items``), which is how the 64-stream configuration
of BASELINE.json runs.  Consecutive pairs (estimate_vel.py:377) shard by range.
"""
from __future__ import annotations

from typing import Callable, List, Sequence, Tuple

import numpy as np

PAIR_FIELDS = 8   # 32-byte records as int32 words (floats by bit pattern): pdx, pdy, flag, nvalid_x, nvalid_y, n_textured, mean_dx, mean_dy


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous [lo, hi) slice of n_items for `rank`; sizes differ by at most one."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_items(items: Sequence, rank: int, world: int) -> List:
    lo, hi = shard_range(len(items), rank, world)
    return list(items[lo:hi])


def frames_needed(pairs: np.ndarray) -> Tuple[int, int]:
    """[lo, hi) frame range a pair slice touches (contiguous upload incl. halo)."""
    if len(pairs) == 0:
        return 0, 0
    return int(pairs.min()), int(pairs.max()) + 1


def records_to_int32(rec: np.ndarray) -> np.ndarray:
    """Structured pair records -> (P, 6) int32 (floats by bit pattern, so the gather is exact)."""
    return np.ascontiguousarray(rec).view(np.int32).reshape(-1, PAIR_FIELDS)


def int32_to_records(a: np.ndarray, dtype) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.int32).reshape(-1).view(dtype)


def search_pairs_sharded(frames: np.ndarray, pairs, search_local: Callable[[np.ndarray, np.ndarray], np.ndarray],
                         record_dtype, group=None, device=None) -> np.ndarray:
    """Search `pairs` (global frame indices into `frames`) across all ranks; every rank returns the
    full, ordered record array.

    ``search_local(frames_slice, local_pairs)`` runs this rank's share: frames_slice is the contiguous
    sub-stack the slice needs, local_pairs index into it.  In the product it is a MotionEstimator
    upload+search; the gloo test passes a CPU stand-in.
    """
    import torch
    import torch.distributed as dist

    pairs = np.ascontiguousarray(np.asarray(pairs, dtype=np.int32).reshape(-1, 2))
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    lo, hi = shard_range(len(pairs), rank, world)
    mine = pairs[lo:hi]
    if len(mine):
        f_lo, f_hi = frames_needed(mine)
        rec = search_local(frames[f_lo:f_hi], mine - f_lo)
    else:
        rec = np.zeros(0, record_dtype)
    if world == 1:
        return rec
    # equal-sized gather: pad every rank's slice to the largest slice
    per = -(-len(pairs) // world)
    buf = np.zeros((per, PAIR_FIELDS), np.int32)
    buf[:len(mine)] = records_to_int32(rec)
    t = torch.from_numpy(buf)
    if device is not None:
        t = t.to(device)
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t, group=group)
    parts = []
    for r in range(world):
        rlo, rhi = shard_range(len(pairs), r, world)
        parts.append(out[r][: rhi - rlo].cpu().numpy())
    return int32_to_records(np.concatenate(parts, axis=0), record_dtype)


# ------------------------------------------------------------------------------------------- GPU ranks
def is_distributed() -> bool:
    try:
        import torch.distributed as dist
    except ImportError:
        return False
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


class RecordGather:
    """all_gather of pair records straight from device memory, off the search's critical path.

    The records of a finished search are copied device-to-device into the send buffer on a side stream
    (``me_copy_pair_records_device``, ordered after the search by an event), gathered with NCCL on that stream and
    copied to pinned host memory there; ``wait()`` returns them.  Nothing blocks the stream the next search runs on,
    so step k's records are gathered and consumed while step k+1 searches.  Two buffers alternate."""

    def __init__(self, per_rank_pairs, record_dtype, group=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.group, self.dtype = torch, dist, group, record_dtype
        self.world = dist.get_world_size(group)
        self.per = int(per_rank_pairs)
        self.side = torch.cuda.Stream()
        self.send = [torch.zeros((self.per, PAIR_FIELDS), dtype=torch.int32, device="cuda") for _ in range(2)]
        self.recv = [torch.zeros((self.world, self.per, PAIR_FIELDS), dtype=torch.int32, device="cuda") for _ in range(2)]
        self.host = [torch.zeros((self.world, self.per, PAIR_FIELDS), dtype=torch.int32).pin_memory() for _ in range(2)]
        self.done = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        self.k = 0

    def stage(self, est, n_pairs, offset=0):
        """Copy the records of est's last search to send[offset : offset + n_pairs] of the buffer the next start() gathers
        (a step searched in several launches stages each of them; start(est, 0) then gathers what was staged)."""
        if n_pairs:
            b = self.k & 1
            est.copy_pair_records_device(self.send[b][offset:].data_ptr(), n_pairs, stream=self.side.cuda_stream)

    def start(self, est, n_pairs):
        """Queue the gather of the records of est's last search; returns a ticket for wait()."""
        b = self.k & 1
        self.k += 1
        if n_pairs:
            est.copy_pair_records_device(self.send[b].data_ptr(), n_pairs, stream=self.side.cuda_stream)
        with self.torch.cuda.stream(self.side):
            self.dist.all_gather_into_tensor(self.recv[b], self.send[b], group=self.group)
            self.host[b].copy_(self.recv[b], non_blocking=True)
            self.done[b].record(self.side)
        return b

    def wait(self, ticket, counts=None):
        """(world * per) records (or the first counts[r] of every rank, concatenated) as a structured array."""
        self.done[ticket].synchronize()
        a = self.host[ticket].numpy()
        if counts is not None:
            a = np.concatenate([a[r, :c] for r, c in enumerate(counts)], axis=0)
        return int32_to_records(a.reshape(-1, PAIR_FIELDS).copy(), self.dtype)


def search_consecutive_sharded(frames: np.ndarray, est_factory, record_dtype, chunk: int = 256, group=None) -> np.ndarray:
    """Records of the consecutive pairs (k, k+1) of a host frame stack, searched on all ranks: rank r takes the
    contiguous pair range shard_range(n - 1, r, world) and uploads only its frames plus the one-frame halo
    (estimate_vel.py:377 -- strictly consecutive pairs are independent).  Every rank returns the whole series.

    ``est_factory(max_frames, max_pairs)`` -> MotionEstimator on this rank's GPU."""
    import torch.distributed as dist
    n_pairs = len(frames) - 1
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    lo, hi = shard_range(n_pairs, rank, world)
    per = -(-n_pairs // world)
    chunk = max(1, min(chunk, per))
    est = est_factory(chunk + 1, chunk)
    gather = RecordGather(chunk, record_dtype, group)
    rounds = -(-per // chunk)                       # every rank takes part in every collective, also with nothing left to search
    tickets, parts = [], [[] for _ in range(world)]
    for i in range(rounds):
        k0 = lo + i * chunk
        k1 = min(k0 + chunk, hi)
        m = max(k1 - k0, 0)
        if m:
            est.upload(frames[k0:k1 + 1], first_slot=0)
            est.launch([(j, j + 1) for j in range(m)])
        tickets.append((gather.start(est, m), i))
        if len(tickets) == 2:                       # consume round i-1 while round i searches
            t, j = tickets.pop(0)
            _collect(gather, t, j, chunk, n_pairs, world, parts)
    for t, j in tickets:
        _collect(gather, t, j, chunk, n_pairs, world, parts)
    return np.concatenate([np.concatenate(p) if p else np.zeros(0, record_dtype) for p in parts])


def _collect(gather, ticket, round_index, chunk, n_pairs, world, parts):
    counts = []
    for r in range(world):
        rlo, rhi = shard_range(n_pairs, r, world)
        k0 = rlo + round_index * chunk
        counts.append(max(min(k0 + chunk, rhi) - k0, 0))
    rec = gather.wait(ticket, counts)
    o = 0
    for r, c in enumerate(counts):
        if c:
            parts[r].append(rec[o:o + c])
        o += c


def gather_frames_sharded(local: "np.ndarray | None", est, first_slot, n_local, n_total, shape, group=None) -> np.ndarray:
    """Reassemble a recording whose contiguous frame ranges were processed on different ranks: rank r holds n_local
    frames in ring slots first_slot.. of `est`; returns the (n_total, H, W) host array on every rank.  The frames go
    ring -> send buffer (device to device) -> NCCL all_gather -> one copy to the host."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    H, W = shape
    per = -(-n_total // world)
    # a stream of our own: torch's default stream has handle 0, which the C ABI reads as "the context's copy stream" --
    # the copy and the collective must be on ONE stream, or the gather can read the send buffer before the copy landed
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        send = torch.zeros((per, H, W), dtype=torch.uint8, device="cuda")
        if n_local:
            est.copy_frames_device(send.data_ptr(), W, first_slot, n_local, stream=side.cuda_stream)
        recv = torch.empty((world, per, H, W), dtype=torch.uint8, device="cuda")
        dist.all_gather_into_tensor(recv, send, group=group)
        host = recv.cpu().numpy()
    side.synchronize()
    out = np.empty((n_total, H, W), np.uint8)
    for r in range(world):
        lo, hi = shard_range(n_total, r, world)
        out[lo:hi] = host[r, :hi - lo]
    return out

"""Frame pairs sharded over the GPUs of one box, one process per GPU (torch.distributed).

The search has no inter-GPU data dependency: rank r takes a contiguous slice of the pair list,
uploads only the frames that slice touches (its frame range plus the one-frame halo a pair needs),
and searches it.  The only exchange is one ``all_gather`` of the per-pair records (24 bytes per
pair) so that every rank -- in particular rank 0, which runs the velocity arithmetic of
estimate_vel_vid.py:540-573 -- sees the whole series.  NCCL over NVLink on GPUs, gloo in CPU tests.

The anchored loop of the video script is sequential (estimate_vel_vid.py:531-534,575) and is NOT
sharded by frame range: recordings are (``shard_items``), which is how the 64-stream configuration
of BASELINE.json runs.  Consecutive pairs (estimate_vel.py:377) shard by range.
"""
from __future__ import annotations

from typing import Callable, List, Sequence, Tuple

import numpy as np

PAIR_FIELDS = 6   # pdx, pdy as float32 bit patterns; flag, nvalid_x, nvalid_y, reserved as int32


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

"""world_size-2 gloo test of the multi-GPU host path on CPU: pair slices, halo frames and the
all_gather of pair records.  The per-rank search is a CPU stand-in (the oracle) -- the collective
and the slicing are what is under test here; the CUDA search is covered by the -m gpu tests."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _local_search(frames_slice, local_pairs):
    from motion_estimation_b200.engine import PAIR_DTYPE
    from oracle import sad
    pd, pi = sad.search_pairs(frames_slice, local_pairs, 16, 8)
    rec = np.zeros(len(local_pairs), PAIR_DTYPE)
    rec["pdx"], rec["pdy"] = pd[:, 0], pd[:, 1]
    rec["flag"], rec["nvalid_x"], rec["nvalid_y"], rec["n_textured"] = pi[:, 0], pi[:, 1], pi[:, 2], pi[:, 3]
    rec["mean_dx"], rec["mean_dy"] = pd[:, 2], pd[:, 3]
    return rec


def _worker(rank, world, port, n_pairs, ret):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from motion_estimation_b200 import shard, synth
    from motion_estimation_b200.engine import PAIR_DTYPE
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    frames, _ = synth.translating_frames(n_pairs + 1, 48, 96, 8)
    pairs = [(k, k + 1) for k in range(n_pairs)]
    rec = shard.search_pairs_sharded(frames, pairs, _local_search, PAIR_DTYPE)
    ret[rank] = rec.tobytes()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n_pairs", [7, 2, 1])
def test_two_ranks_return_the_single_rank_series(n_pairs):
    from motion_estimation_b200 import synth
    from motion_estimation_b200.engine import PAIR_DTYPE
    frames, shifts = synth.translating_frames(n_pairs + 1, 48, 96, 8)
    want = _local_search(frames, np.array([(k, k + 1) for k in range(n_pairs)], np.int32))
    mgr = mp.Manager()
    ret = mgr.dict()
    port = 29500 + (os.getpid() % 2000) + n_pairs
    mp.spawn(_worker, args=(2, port, n_pairs, ret), nprocs=2, join=True)
    for r in range(2):
        got = np.frombuffer(ret[r], PAIR_DTYPE)
        assert got.tobytes() == want.tobytes(), f"rank {r}"
    assert [round(float(v)) for v in want["pdx"]] == [int(s[0]) for s in shifts]


def test_round_bookkeeping_of_the_sharded_consecutive_search():
    """Host logic of shard.search_consecutive_sharded without GPUs: every rank takes part in the same number of gather
    rounds, a round carries min(chunk, what is left of the rank's slice) records per rank, and putting the rounds back
    together gives the pair series in order -- for slices of unequal length and a last round that some ranks sit out."""
    sys.path.insert(0, ROOT)
    from motion_estimation_b200 import shard
    from motion_estimation_b200.engine import PAIR_DTYPE

    class FakeGather:                       # what RecordGather.wait returns: the ranks' records of one round, concatenated
        def __init__(self, n_pairs, world, chunk):
            self.n_pairs, self.world, self.chunk = n_pairs, world, chunk

        def wait(self, ticket, counts):
            parts = []
            for r, c in enumerate(counts):
                lo, _ = shard.shard_range(self.n_pairs, r, self.world)
                rec = np.zeros(c, PAIR_DTYPE)
                rec["nvalid_x"] = np.arange(lo + ticket * self.chunk, lo + ticket * self.chunk + c)     # the global pair index
                parts.append(rec)
            return np.concatenate(parts)

    for n_pairs, world, chunk in ((23, 2, 4), (7, 4, 3), (5, 8, 2), (64, 8, 8), (1, 2, 4)):
        per = -(-n_pairs // world)
        chunk = max(1, min(chunk, per))
        rounds = -(-per // chunk)
        parts = [[] for _ in range(world)]
        g = FakeGather(n_pairs, world, chunk)
        for i in range(rounds):
            shard._collect(g, i, i, chunk, n_pairs, world, parts)
        got = np.concatenate([np.concatenate(p) if p else np.zeros(0, PAIR_DTYPE) for p in parts])
        assert got["nvalid_x"].tolist() == list(range(n_pairs)), (n_pairs, world, chunk)

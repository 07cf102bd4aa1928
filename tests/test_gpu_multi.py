"""Multi-GPU tests of the PRODUCT paths (skipped on a box with fewer than two GPUs): one recording pushed through
two ranks -- consecutive-pair search with the pair records gathered from device memory, non-local-means denoising of
contiguous frame ranges with halo frames, recordings shared out for the anchored loop -- must give exactly the bytes the
single-GPU path gives.  One process per GPU, NCCL, rendezvous on 127.0.0.1."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _inputs():
    from motion_estimation_b200 import synth
    frames, _ = synth.translating_frames(23, 96, 256, 12, seed_texture=4, seed_motion=9)
    stamps_us = [5531 * (k + 1) for k in range(len(frames))]
    stamps_ms = [5.531 * k for k in range(len(frames) + 3)]
    return frames, stamps_us, stamps_ms


def _run(api, me, frames, stamps_us, stamps_ms):
    cfg = me.SearchConfig(16, 16)
    raw = api.end_process_raw(frames, stamps_us, config=cfg, max_pairs=4, return_details=True)       # several rounds per rank
    clean = api.denoise_hf5(frames[:14], n_list=15)
    recs = api.end_process_recordings([dict(clean_arrs=frames[:12], timestamps_ms=stamps_ms),
                                       dict(clean_arrs=frames[5:], timestamps_ms=stamps_ms),
                                       dict(clean_arrs=frames[::2], timestamps_ms=stamps_ms)], config=cfg)
    return dict(raw_info=np.array(raw.out_information), raw_disp=np.array(raw.pair_displacements), clean=clean,
                recs=[np.array(r.out_information) for r in recs])


def _worker(rank, world, port, ret):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import motion_estimation_b200 as me
    from motion_estimation_b200 import api
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    frames, stamps_us, stamps_ms = _inputs()
    out = _run(api, me, frames, stamps_us, stamps_ms)
    ret[rank] = {k: (v.tobytes() if isinstance(v, np.ndarray) else [a.tobytes() for a in v]) for k, v in out.items()}
    api.close_engines()
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_give_the_single_gpu_bytes(me):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from motion_estimation_b200 import api
    frames, stamps_us, stamps_ms = _inputs()
    single = _run(api, me, frames, stamps_us, stamps_ms)
    api.close_engines()
    want = {k: (v.tobytes() if isinstance(v, np.ndarray) else [a.tobytes() for a in v]) for k, v in single.items()}
    assert len(single["raw_info"]) == len(frames) - 1 and single["clean"].shape == (14, 96, 256)
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_worker, args=(2, 29517, ret), nprocs=2, join=True)
        for rank in range(2):
            for key in want:
                assert ret[rank][key] == want[key], f"rank {rank}: {key} differs from the single-GPU result"

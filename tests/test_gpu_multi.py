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


def _make_recording(d):
    """A small synthetic recording as an image sequence + timecode file (what the CLI reads)."""
    import cv2
    from motion_estimation_b200 import synth
    n, H, W = 20, 64, 320
    canvas = synth.texture(H + 8, W + 8 * n + 16, seed=3)
    os.makedirs(d, exist_ok=True)
    for k in range(n):
        cv2.imwrite(os.path.join(d, f"out.{k:04d}.png"), canvas[4:4 + H, 8 * n - 7 * k: 8 * n - 7 * k + W])
    with open(os.path.join(d, "rec_tstamps.txt"), "w") as fh:
        fh.write("# timecode format v2\n" + "\n".join(f"{5.531 * k:.3f}" for k in range(n + 2)) + "\n")


def _cli(rec, out):
    import importlib.util
    spec = importlib.util.spec_from_file_location("estimate_vel_vid", os.path.join(ROOT, "estimate_vel_vid.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.main(["0", "18", "-p", rec, "--out", out, "--range", "16"])


def _worker(rank, world, port, ret, rec=None):
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
    if rec is not None:                                   # the command line under two ranks: non-local means sharded, rank 0 writes
        os.environ.update(WORLD_SIZE=str(world), RANK=str(rank), LOCAL_RANK=str(rank))
        res = _cli(rec, rec + f"_out{world}")
        assert (res is None) == (rank != 0)
    api.close_engines()
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_give_the_single_gpu_bytes(me, tmp_path):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from motion_estimation_b200 import api
    rec = str(tmp_path / "rec.d")
    _make_recording(rec)
    _cli(rec, rec + "_out1")
    frames, stamps_us, stamps_ms = _inputs()
    single = _run(api, me, frames, stamps_us, stamps_ms)
    api.close_engines()
    want = {k: (v.tobytes() if isinstance(v, np.ndarray) else [a.tobytes() for a in v]) for k, v in single.items()}
    assert len(single["raw_info"]) == len(frames) - 1 and single["clean"].shape == (14, 96, 256)
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_worker, args=(2, 29517, ret, rec), nprocs=2, join=True)
        for rank in range(2):
            for key in want:
                assert ret[rank][key] == want[key], f"rank {rank}: {key} differs from the single-GPU result"
    # the command line: same CSV from one process and from two ranks (CLAHE + sharded NLM + pair loop)
    two, one = (open(os.path.join(rec + d, "velocity_estimates.csv")).read() for d in ("_out2", "_out1"))
    assert two == one, "two ranks:\n" + two[:400] + "\none process:\n" + one[:400]

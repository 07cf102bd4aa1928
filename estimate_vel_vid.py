#!/usr/bin/env python3
"""B200 drop-in for the reference's video pipeline entry point.

Same command line as the reference script of this name (its argparse block is at lines 34-69):

    estimate_vel_vid.py [first last] [-f] [-p DIR] [-c] [-h264]  (+ the options added below)

DIR holds the recording (``video.mkv``, else the first ``*.mkv`` / ``*.h264``), the first
``*tstamps.txt`` and optionally ``*encoder.csv``.  Output is ``velocity_estimates.csv`` with the
reference's header and row layout (its lines 410, 572, 633-636).  What changed is the engine: the
per-pair displacement is the CUDA block search instead of cv.findTransformECC, and the recording is
decoded sequentially into one array.  Like the reference's main() (its lines 697-709) the default run is
decode -> CLAHE -> non-local-means denoise -> pair loop; the two pre-processing stages (its lines 201-299, 99 %
of its run time) run on the GPU, bit-exact with the cv2 calls they replace.  --no-preprocess skips them (the
search then sees the decoded frames); --no-clahe / --no-denoise skip one.  The cleaned frames are cached as
DIR/images.h5, dataset 'clean_images', exactly as the reference writes them (its lines 287-297) when h5py is
importable, else as DIR/images.npy; -c reloads whichever exists (its lines 711-713).
"""
import argparse
import os
import sys
from pathlib import Path

import numpy as np


def parse(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("fileIndex", type=int, nargs="*", help="first and last frame index")
    ap.add_argument("-f", "--full", nargs="?", dest="useWholeVideo", default="False", const="True")
    ap.add_argument("-p", "--path", nargs="?", type=Path, dest="srcDirectory", default="/dev/shm/")
    ap.add_argument("-c", "--continue", nargs="?", dest="continuation", default="False", const="True",
                    help="reuse the cleaned frames of a previous run (DIR/images.h5 ['clean_images'], else DIR/images.npy)")
    ap.add_argument("-h264", nargs="?", dest="h264", default="False", const="True",
                    help="read the first *.h264 in DIR instead of the .mkv")
    # additions (the reference has no search parameters)
    ap.add_argument("--block", type=int, default=16)
    ap.add_argument("--range", type=int, default=64, dest="search_range",
                    help="search range in full-resolution pixels (default 64 = 32 of the reference's half-resolution pixels)")
    ap.add_argument("--range-y", type=int, default=0, dest="range_y",
                    help="vertical search range (default 0 = the same as --range); the recordings move along x, a small value is several times faster")
    ap.add_argument("--min-range", type=int, default=6, dest="min_range",
                    help="texture gate: a block takes part when max - min of its pixels reaches this many grey levels (default 6)")
    ap.add_argument("--units", choices=("reference", "fullres"), default="reference",
                    help="'reference' halves displacements like the reference's half-resolution ECC result")
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--out", type=Path, default=None, help="directory for the CSV (default: DIR, else the cwd)")
    ap.add_argument("--no-preprocess", action="store_true", help="skip CLAHE and denoising: search the decoded frames (fast path, not what the reference does)")
    ap.add_argument("--no-clahe", action="store_true", help="skip CLAHE(2.0,(8,8)) (reference :201-219; GPU kernels, bit-exact with cv2)")
    ap.add_argument("--no-denoise", action="store_true", help="skip the non-local-means denoise (reference :246-299: h=15, 19/41 windows, 1/5/9 frames; GPU kernel, bit-exact with cv2)")
    ap.add_argument("--preprocess", action="store_true", help="accepted for compatibility with earlier versions: pre-processing is the default now")
    a = ap.parse_args(argv)
    whole = a.useWholeVideo == "True"
    if not whole and len(a.fileIndex) != 2:
        ap.error("give FIRST LAST frame indices, or -f for the whole video")
    return a, whole


def load_cache(cache_h5, cache_npy):
    """The -c cache: images.h5['clean_images'] as the reference writes it (:287-297) when h5py can read it, else images.npy."""
    if os.path.exists(cache_h5):
        try:
            import h5py
            with h5py.File(cache_h5, "r") as f:
                return np.asarray(f["clean_images"][:], dtype=np.uint8)
        except ImportError:
            print(f"{cache_h5} exists but h5py is not installed; looking for {cache_npy}")
    if os.path.exists(cache_npy):
        return np.load(cache_npy)
    return None


def save_cache(frames, cache_h5, cache_npy):
    from motion_estimation_b200 import api
    if not api.write_clean_images_h5(frames, cache_h5):
        np.save(cache_npy, frames)


def ranks():
    """(rank, world) when started under torchrun (one process per GPU): the non-local-means stage -- nine tenths of the run time --
    then shards its frame ranges over the GPUs (api.denoise_hf5); everything else runs on rank 0.  (0, 1) otherwise."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world <= 1:
        return 0, 1
    import torch
    import torch.distributed as dist
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if not dist.is_initialized():
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return dist.get_rank(), dist.get_world_size()


def main(argv=None):
    a, whole = parse(argv)
    import motion_estimation_b200 as me
    from motion_estimation_b200 import api, ingest
    rank, world = ranks()
    if world > 1:
        a.device = int(os.environ.get("LOCAL_RANK", "0"))

    src = str(a.srcDirectory)
    first, last = (0, None) if whole else a.fileIndex
    cache_h5, cache_npy = os.path.join(src, "images.h5"), os.path.join(src, "images.npy")
    if a.continuation == "True":                      # reference :711-713
        frames = load_cache(cache_h5, cache_npy)
        if frames is None:
            sys.exit(f"-c given but neither {cache_h5} (needs h5py) nor {cache_npy} can be read")
    else:
        target = src
        if a.h264 == "True":
            hits = sorted(Path(src).glob("*.h264"))
            if not hits:
                sys.exit(f"no *.h264 in {src}")
            target = str(hits[0])
        frames = ingest.read_frames(target, first, last)
        api.set_defaults(device=a.device)
        do_clahe = not (a.no_preprocess or a.no_clahe)
        do_denoise = not (a.no_preprocess or a.no_denoise)
        if do_clahe:                                  # reference :201-219, on the GPU, bit-exact with cv2's CLAHE
            frames = api.adaptive_histogram_equalization(frames)
        if do_denoise:                                # reference :246-299: non-local means, bit-exact with cv2's
            frames = api.denoise_hf5(frames, n_list=len(frames) + 1)      # len(numberList) = frames + 1 (its line 73)
        if (do_clahe or do_denoise) and os.access(src, os.W_OK) and rank == 0:
            save_cache(frames, cache_h5, cache_npy)
    if rank != 0:                                     # the pair loop is sequential inside a recording: rank 0 carries on alone
        import torch.distributed as dist
        dist.barrier()
        return None
    if len(frames) < 2:
        sys.exit("need at least two frames")
    number_list = list(range(first, first + len(frames) + 1))
    cfg = me.SearchConfig(block=a.block, range=a.search_range, range_y=a.range_y, min_range=a.min_range)
    res = api.end_process(frames, imagePath=src, firstFrame=first, numberList=number_list, config=cfg,
                          device=a.device, units=a.units, return_details=True)
    out_dir = str(a.out) if a.out else (src if os.access(src, os.W_OK) else os.getcwd())
    os.makedirs(out_dir, exist_ok=True)
    from motion_estimation_b200 import postprocess
    postprocess.write_velocity_csv(out_dir, res.out_information)        # reference header + (vx, vy, t) rows (its lines 410, 572)
    # encoder ground truth and the resampled position series, when the recording ships an encoder log (reference 417-487, 589-642)
    import glob
    if glob.glob(os.path.join(src, "*encoder.csv")) and len(res.tsList) >= 3:
        # the reference indexes the timecode file with lastFrame (its line 429); the example files are
        # truncated, so fall back to the last line that still holds a value
        stamp_lines = open(ingest.find_timestamps(src)).read().split("\n")
        last_line = min(first + len(frames), max(i + 1 for i, v in enumerate(stamp_lines) if v.strip() and not v.startswith("#")))
        enc_pos, enc_ts = postprocess.encoder_velocity(src, last_line)
        opt, enc, grid, _, _ = postprocess.dataset_correlation(res.motion_list_opt, res.tsList, enc_pos, enc_ts)
        postprocess.write_positions_csv(out_dir, grid, opt, enc)
        print(f"wrote {os.path.join(out_dir, 'correlated_positions.csv')}")
    pos = res.motion_list_opt[-1] if res.motion_list_opt else 0.0
    print(f"\nfailed motion estimates: {res.failed_estimates} of {len(number_list)}")
    print(f"final position according to block search: {pos}")
    print(f"final position according to the dense block field: {res.final_position_dense}")     # the reference's Farneback figure (:582)
    # a search range that is too small cannot be seen in one pair (false matches look like any other minimum); over a
    # recording it shows as displacements crowding the range and a low share of blocks with a clean interior minimum
    scale = 0.5 if a.units == "reference" else 1.0
    near = [abs(d[0]) > 0.75 * scale * a.search_range for d in res.pair_displacements]
    if near and sum(near) > 0.1 * len(near):
        print(f"warning: {sum(near)} of {len(near)} accepted displacements exceed 75 % of the search range "
              f"(+-{a.search_range} px): rerun with a larger --range")
    print(f"accepted pairs: {len(res.accepted_pairs)}; searches run: {res.searches} in {res.launches} launches")
    print(f"wrote {os.path.join(out_dir, 'velocity_estimates.csv')}")
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    return res


if __name__ == "__main__":
    main()

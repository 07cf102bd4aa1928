#!/usr/bin/env python3
"""bench.py -- throughput of the displacement-search path on BASELINE.json's workloads.

    python bench.py --gpus N --steps K --warmup W [--workload NAME]      (N > 1: launched by torch.distributed.run)
    python bench.py --impl reference --gpus N --steps K --warmup W [--workload NAME]

Workloads (one JSON line each; `config.workload` names it):

  1080p     DEFAULT, BASELINE configs[2]: synthetic 1920x1080 uint8 luma video, 16x16 blocks, +-32 px full search, the
            estimate_vel_vid.py path.  One STEP = 64 consecutive frame pairs (65 frames, 134.8 MB > the 126 MB L2) per
            GPU: search + per-pair median on the GPU, the records of all ranks all-gathered from device memory (NCCL, side
            stream, overlapped with the next step) and reduced to a velocity series on rank 0.  Weak scaling.
  strong    the same search on ONE recording of 512 pairs split over the ranks (strong scaling).
  4k        configs[3]: 3840x2160, 8x8 blocks, +-64 px, 16 pairs per step per GPU, pairs sharded over the ranks.
  sweep64   configs[4]: 64 concurrent recordings shared out over the ranks, 720p / 1080p / 4K x range 8..128, one pair per
            recording per step; the line's value is pairs/s over the whole sweep, the table is in `sweep`.
  clip      the example recordings' shape (640x64, 180 fps) through api.end_process: the anchored loop on a 4096-frame
            synthetic recording; value = frames/s on decoded frames, e2e = frames/s with the reference's CLAHE +
            non-local-means pre-processing in front (its default pipeline).
  nlm       the non-local-means kernel alone (640x64, h=15, 19/41 windows, 9-frame window), frames/s.
  clahe     the CLAHE kernels alone (64 x 1080p per step), frames/s against the HBM roofline.

Every line carries: value (inputs resident in HBM, device-timed with CUDA events, max over ranks), e2e (host buffers in,
host results out, copies inside the timed region), roofline for the dominant kernel, cpu_baseline (a bounded sample on the
host cores, rank 0, N = 1 only), clocks sampled during the timed region, gpu_launches.  The e2e arm of the search workloads
uploads step k+1 while step k searches; its first step goes in four slices (ME_BENCH_FILL_PARTS) so that the search starts
after a quarter of an upload.

The reference arm runs the reference's own CPU code for the workload (its pyramid ECC per pair, restated in
oracle/ecc_port.py because /root/reference is Python and does not travel; cv2's CLAHE / fastNlMeansDenoisingMulti for
the two pre-processing workloads) on the host cores.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PIXELS_PER_MM, FRAME_DT_MS = 611.0, 5.531
WORKLOADS = ("1080p", "strong", "4k", "sweep64", "clip", "nlm", "clahe")
NCU_DRAM_BYTES_1080P = 137510656 + 8808704      # profiles/r02_1080p_ncu_summary.txt: sad_search_persistent<16,5,4,208>, 64 pairs of 1080p
SEARCH_SHAPES = {   # workload -> (H, W, block, range, pairs per step per GPU)
    "1080p": (1080, 1920, 16, 32, 64),
    "strong": (1080, 1920, 16, 32, 512),
    "4k": (2160, 3840, 8, 64, 16),
}
METRICS = {
    "1080p": "frame-pairs/sec at 1080p, 16x16 blocks, +-32 px full-search SAD",
    "strong": "frame-pairs/sec at 1080p, 16x16 blocks, +-32 px full-search SAD (one 512-pair recording, strong scaling)",
    "4k": "frame-pairs/sec at 4K, 8x8 blocks, +-64 px full-search SAD",
    "sweep64": "frame-pairs/sec over 64 concurrent recordings, 720p-4K, +-8..128 px",
    "clip": "frames/sec through end_process on a 640x64 recording",
    "nlm": "frames/sec of non-local-means denoising, 640x64, 9-frame window",
    "clahe": "frames/sec of CLAHE at 1080p",
}
NAMES = {
    "1080p": "configs[2]: synthetic 1080p uint8 luma video, 16x16 blocks, +-32 px full search, estimate_vel_vid.py path",
    "strong": "configs[2] shape, one recording of 512 frame pairs split over the GPUs",
    "4k": "configs[3]: synthetic 4K video, 8x8 blocks, +-64 px search, frame pairs sharded across the GPUs",
    "sweep64": "configs[4]: 64 concurrent 720p-4K recordings, search range 8-128 px",
    "clip": "configs[1] shape: 640x64 recording, anchored pair loop of estimate_vel_vid.py (end_process), with and without its pre-processing",
    "nlm": "pre-processing stage: cv.fastNlMeansDenoisingMulti(h=15, 19, 41, 9 frames) on 640x64 frames (estimate_vel_vid.py:246-299)",
    "clahe": "pre-processing stage: cv.createCLAHE(2.0, (8,8)).apply on 1080p frames (estimate_vel_vid.py:201-219)",
}


def read_measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons sampled with NVML while a timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.sm, self.reasons, self.max_mhz, self.err = index, False, [], set(), None, None

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {nv.nvmlClocksEventReasonHwSlowdown: "hw_slowdown",
                     nv.nvmlClocksEventReasonHwThermalSlowdown: "hw_thermal_slowdown",
                     nv.nvmlClocksEventReasonSwThermalSlowdown: "sw_thermal_slowdown",
                     nv.nvmlClocksEventReasonSwPowerCap: "sw_power_cap"}
            while not self.stop_flag:
                self.sm.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                bits = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                for bit, name in names.items():
                    if bits & bit:
                        self.reasons.add(name)
                time.sleep(0.02)
        except Exception as e:  # pragma: no cover
            self.err = repr(e)

    def result(self):
        self.stop_flag = True
        self.join(timeout=2)
        d = {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz,
             "reasons": sorted(self.reasons), "samples": len(self.sm)}
        if self.err:
            d["error"] = self.err
        return d


class Ranks:
    """torch.distributed plumbing of one bench process (one per GPU)."""

    def __init__(self, gpus):
        import torch
        self.torch = torch
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != gpus:
            raise SystemExit(f"--gpus {gpus} but WORLD_SIZE={self.world}: launch N>1 with torch.distributed.run")
        torch.cuda.set_device(self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max(self, x):
        if self.dist is None:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum(self, x):
        if self.dist is None:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return float(t.item())

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()


def alu_peak(me, local, peaks, clocks):
    """(peak cand*px/s of the VABSDIFF4 pipe, description, SM count, instr/clk/SM, f_max Hz): the issue rate is
    microbenchmarked in this run (the repo's own kernel; MEASURED_PEAKS.json has no integer entry), the clock is the
    driver-measured maximum SM clock."""
    ipc, mhz_seen, sms = me.microbench_vabsdiff4(local, 0)
    f_max = float(peaks.get("sm_max_mhz") or (clocks or {}).get("sm_max_mhz") or 1965.0) * 1e6
    return sms * ipc * 4.0 * f_max, sms, ipc, f_max


# ------------------------------------------------------------------------------- search workloads: 1080p, strong, 4k
def velocity_series(pdx):
    """Rank 0's reduction of gathered displacements to mm/min (estimate_vel_vid.py:556-560 arithmetic)."""
    gap_min = FRAME_DT_MS / 1000.0 / 60.0
    return (np.asarray(pdx, np.float64) / PIXELS_PER_MM) / gap_min


def run_search(args, rk, wl):
    import torch
    import motion_estimation_b200 as me
    from motion_estimation_b200 import shard, synth

    H, W, B, RNG, PAIRS = SEARCH_SHAPES[wl]
    strong = wl == "strong"
    if strong:                                         # one recording, the same on every rank; this rank's contiguous pair slice
        frames_all, shifts_all = synth.translating_frames(PAIRS + 1, H, W, RNG, seed_texture=3, seed_motion=4)
        lo, hi = shard.shard_range(PAIRS, rk.rank, rk.world)
        frames_np, shifts = frames_all[lo:hi + 1], shifts_all[lo:hi]
        del frames_all
    else:
        frames_np, shifts = synth.translating_frames(PAIRS + 1, H, W, RNG, seed_texture=10 * rk.rank, seed_motion=10 * rk.rank + 1)
    P = len(frames_np) - 1
    per = -(-PAIRS // rk.world) if strong else PAIRS    # records per rank in the gather
    cfg = me.SearchConfig(block=B, range=RNG)
    est = me.MotionEstimator(H, W, cfg, max_frames=2 * (P + 1), max_pairs=P, device=rk.local)
    geo = est.geometry
    pinned = torch.empty((P + 1, H, W), dtype=torch.uint8, pin_memory=True)
    pinned.numpy()[:] = frames_np
    host_frames = pinned.numpy()
    half = [[(k, k + 1) for k in range(P)], [(P + 1 + k, P + 2 + k) for k in range(P)]]
    gather = shard.RecordGather(per, me.PAIR_DTYPE) if rk.world > 1 else None
    counts = [shard.shard_range(PAIRS, r, rk.world)[1] - shard.shard_range(PAIRS, r, rk.world)[0] for r in range(rk.world)] if strong else None
    peaks, peak_src = read_measured_peaks()

    # ---- value: frames resident in HBM; the all_gather of every step's records is inside the timed region (N > 1)
    est.upload(host_frames, first_slot=0)
    est.sync()
    for _ in range(args.warmup):
        est.launch(half[0])
        if gather:
            gather.wait(gather.start(est, P))
    est.sync()
    rec = est.fetch_pairs(P)
    inside = np.all(np.abs(shifts) < RNG, axis=1)      # a shift of exactly +-R sits on the border of the search: never an interior minimum
    ok = bool(np.all(np.rint(rec["pdx"])[inside] == shifts[inside, 0]) and np.all(np.rint(rec["pdy"])[inside] == shifts[inside, 1])
              and rec["flag"][inside].all())
    rk.barrier()
    sampler = ClockSampler(rk.local)
    sampler.start()
    launches0 = est.launch_count()
    vel = None
    if gather:
        t0 = torch.cuda.Event(enable_timing=True)
        t0.record(gather.side)                         # both streams are idle after the barrier: this is the start of the region
        est.timer_start()
        prev = None
        for _ in range(args.steps):
            est.launch(half[0])
            ticket = gather.start(est, P)              # device -> device -> NCCL on the side stream; the next launch does not wait
            if prev is not None:
                vel = velocity_series(gather.wait(prev, counts)["pdx"])     # consume step k-1 while step k searches
            prev = ticket
        search_ms = est.timer_stop()
        all_rec = gather.wait(prev, counts)
        vel = velocity_series(all_rec["pdx"])
        dev_ms = t0.elapsed_time(gather.done[prev])    # first launch .. last gathered record on the host-visible buffer
    else:
        est.timer_start()
        for _ in range(args.steps):
            est.launch(half[0])
        dev_ms = search_ms = est.timer_stop()
        all_rec = est.fetch_pairs(P)
        vel = velocity_series(all_rec["pdx"])
    rk.barrier()
    launches = est.launch_count() - launches0
    clocks = sampler.result()
    ms_step = rk.max(dev_ms / args.steps)
    search_ms_step = rk.max(search_ms / args.steps)
    total_pairs = PAIRS if strong else rk.world * PAIRS
    value = total_pairs / (ms_step * 1e-3)
    kernel_ms = []
    for _ in range(3):
        est.launch(half[0])
        est.sync()
        kernel_ms.append(est.last_search_ms())

    # ---- e2e: host frames in, host records out, every step (pinned host -> device inside the region, overlapped)
    # (measured at N = 1, 10 steps: 1080p 8.575 -> 8.422 ms per step with the first step in four slices; the 512-pair recording of the
    # strong-scaling workload did not gain -- 68.0 -> 68.5 ms -- and keeps its first step whole)
    parts = max(1, min(int(os.environ.get("ME_BENCH_FILL_PARTS", "1" if strong else "4")), P))
    cut = [P * i // parts for i in range(parts + 1)]

    def first_step():
        """Step 0 in `parts` slices -- the search of a slice starts when ITS frames are on the device, so the pipeline fills after
        a quarter of an upload instead of a whole one.  Returns (ticket, pairs) of every slice (no gather)."""
        tickets = []
        for i in range(parts):
            a, b = cut[i], cut[i + 1]
            lo = 0 if i == 0 else a + 1
            est.upload(host_frames[lo:b + 1], first_slot=lo)
            est.launch(half[0][a:b])
            if gather:
                gather.stage(est, b - a, offset=a)
            else:
                tickets.append((est.last_ticket(), b - a))
        return tickets

    def e2e_steps(n):
        out, prev, head = None, None, []
        for s in range(n):
            if s == 0:
                tickets = first_step()
            else:
                est.launch(half[s & 1])
            est.upload(host_frames, first_slot=(P + 1) * ((s + 1) & 1))     # next step's input, overlaps this search
            if gather:
                ticket = gather.start(est, 0 if s == 0 else P)              # step 0's records were staged slice by slice
                if prev is not None:
                    out = gather.wait(prev, counts)
            else:
                ticket = est.last_ticket()
                if s == 0:                                                   # the record ring holds four searches: take the early slices now
                    head = [est.fetch_pairs(m, ticket=t) for t, m in tickets[:-1]]
                if prev is not None:
                    tail = est.fetch_pairs(P if s > 1 else cut[-1] - cut[-2], ticket=prev)   # step k-1's records while step k searches
                    out = np.concatenate(head + [tail]) if s == 1 else tail
            prev = ticket
        if gather:
            return gather.wait(prev, counts)
        tail = est.fetch_pairs(P if n > 1 else cut[-1] - cut[-2], ticket=prev)
        return np.concatenate(head + [tail]) if n == 1 else tail

    e2e_steps(max(args.warmup, 2))
    rk.barrier()
    sampler2 = ClockSampler(rk.local)
    sampler2.start()
    launches1 = est.launch_count()
    t0w = time.perf_counter()
    e2e_rec = e2e_steps(args.steps)
    rk.barrier()
    e2e_s = rk.max(time.perf_counter() - t0w)
    e2e_launches = est.launch_count() - launches1
    clocks2 = sampler2.result()
    e2e_value = total_pairs * args.steps / e2e_s
    mine = e2e_rec[sum(counts[:rk.rank]):sum(counts[:rk.rank]) + P] if strong and gather else e2e_rec[rk.rank * P:(rk.rank + 1) * P]
    ok = ok and bool(np.array_equal(mine["pdx"], rec["pdx"]))
    ok_all = rk.sum(0.0 if ok else 1.0) == 0.0

    peak_alu, sms, ipc, f_max = alu_peak(me, rk.local, peaks, clocks)
    if rk.rank != 0:
        est.close()
        return None
    k_ms = search_ms / args.steps                       # the kernel's average launch duration over the timed region (CUDA events, its stream)
    cand_valid, cand_nominal = geo["cand_px_per_pair_valid"] * P, geo["cand_px_per_pair"] * P
    achieved = cand_valid / (k_ms * 1e-3)               # the dominant kernel alone, this rank's launch
    alg_bytes = (P + 1) * H * W + P * geo["nbx"] * geo["nby"] * 17 + P * 32
    hbm_gbs = float(peaks["hbm_gbs"])
    kname = "sad_search_persistent" if geo.get("resident_kernel") else "sad_search_kernel"
    roofline = {
        "bound": "int_alu", "kernel": f"{kname}<{B},5,4,{geo['window_pitch']}>", "achieved": achieved / 1e9, "peak": peak_alu / 1e9,
        "unit": "Gcand*px/s", "frac": achieved / peak_alu,
        # dram__bytes_read.sum + dram__bytes_write.sum of one launch of this kernel on this workload, from the committed ncu --set full
        # capture (137.51 + 8.81 MB against 143.5 MB algorithmic: frames read once, records written once); other shapes: not captured
        "traffic": NCU_DRAM_BYTES_1080P if (wl == "1080p" and P == 64 and geo.get("resident_kernel")) else None,
        "traffic_note": "from profiles/r02_1080p_ncu_summary.txt (ncu --set full of the same kernel and step, a separate run)",
        "peak_is": f"{sms} SMs x {ipc:.2f} VABSDIFF4 thread-instr/clk/SM (this repo's microbenchmark, run here; MEASURED_PEAKS.json "
                   f"has no integer-ALU entry) x 4 px x {f_max / 1e6:.0f} MHz max SM clock; {peak_src}",
        "frac_at_observed_clock": achieved / (sms * ipc * 4.0 * (clocks["sm_mhz"] or f_max / 1e6) * 1e6),
        "achieved_nominal": cand_nominal / (k_ms * 1e-3) / 1e9,
        "kernel_ms_per_launch": k_ms, "kernel_share_of_step": k_ms / ms_step, "search_ms_per_step_max_over_ranks": search_ms_step,
        "kernel_ms_alone": float(np.mean(kernel_ms)),
        "hbm": {"achieved": alg_bytes / (k_ms * 1e-3) / 1e9, "peak": hbm_gbs, "unit": "GB/s",
                "frac": alg_bytes / (k_ms * 1e-3) / 1e9 / hbm_gbs, "algorithmic_bytes_per_launch": alg_bytes},
    }

    cpu = None
    if rk.world == 1:
        from oracle import sad as osad
        n, t_cpu = 0, time.perf_counter()
        while n < min(24, P) and (time.perf_counter() - t_cpu < 12.0 or n < 1):
            o = osad.search(frames_np[n], frames_np[n + 1], B, RNG)
            ok_all = ok_all and o["pdx"] == rec["pdx"][n] and o["pdy"] == rec["pdy"][n]     # the checker, not the product
            n += 1
        t_cpu = time.perf_counter() - t_cpu
        cpu = {"value": n / t_cpu, "unit": "frame-pairs/s", "cores": osad.max_threads(), "kind": "port",
               "sample": f"first {n} of the step's {P} pairs through oracle/sad_oracle.c (OpenMP over blocks, psadbw SAD), {t_cpu:.1f} s",
               "gcand_px_per_s": n * geo["cand_px_per_pair_valid"] / t_cpu / 1e9}

    line = {
        "metric": METRICS[wl], "value": value, "unit": "frame-pairs/s", "n_gpus": rk.world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if strong else "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": NAMES[wl], "height": H, "width": W, "block": B, "range": RNG,
                   "pairs_per_step_per_gpu": P, "frames_per_step_per_gpu": P + 1,
                   "l2": f"inputs larger than L2: {(P + 1) * H * W / 1e6:.1f} MB of frames per step per GPU vs 126 MB",
                   "parallelism": (f"frame pairs sharded over {rk.world} GPU(s); per step one NCCL all_gather of the 32-byte pair records, "
                                   "read from device memory on a side stream and overlapped with the next step's search; inside the timed region"
                                   if rk.world > 1 else "1 GPU, no exchange step"),
                   "tile_blocks_x": geo["tile_blocks_x"], "threads_per_cta": geo["threads_per_cta"],
                   "window_pitch": geo["window_pitch"], "candidate_slabs": [geo["slabs_x"], geo["slabs_y"]],
                   "smem_bytes": geo["smem_bytes"], "ctas_per_pair": geo["tiles_per_pair"]},
        "gcand_px_per_s": total_pairs * geo["cand_px_per_pair_valid"] / (ms_step * 1e-3) / 1e9,
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "frame-pairs/s", "h2d_bytes_per_step": (P + 1) * H * W,
                "d2h_bytes_per_step": total_pairs * 32 if gather else P * 32, "ms_per_step": e2e_s * 1e3 / args.steps,
                "clocks": clocks2, "gpu_launches": int(e2e_launches)},
        "gpu_launches": int(launches),
        "roofline": roofline,
        "cpu_baseline": cpu,
        "results_checked": bool(ok_all),
        "velocity_mm_min_first3": [float(v) for v in np.ravel(vel)[:3]],
    }
    est.close()
    return line


# ------------------------------------------------------------------------------------------------------- sweep64
def run_sweep(args, rk):
    """configs[4]: 64 concurrent recordings (64 / N per GPU), one frame pair per recording per step, every shape x range."""
    import torch
    import motion_estimation_b200 as me
    from motion_estimation_b200 import shard, synth
    RECS, B = 64, 16
    lo, hi = shard.shard_range(RECS, rk.rank, rk.world)
    P = hi - lo
    peaks, peak_src = read_measured_peaks()
    peak_alu, sms, ipc, f_max = alu_peak(me, rk.local, peaks, None)
    steps = max(2, min(args.steps, 5))
    table, tot_pairs, tot_ms, tot_e2e_s, launches, h2d = [], 0, 0.0, 0.0, 0, 0
    gather = shard.RecordGather(-(-RECS // rk.world), me.PAIR_DTYPE) if rk.world > 1 else None
    sampler = ClockSampler(rk.local)
    sampler.start()
    for name, (H, W) in (("720p", (720, 1280)), ("1080p", (1080, 1920)), ("4k", (2160, 3840))):
        frames, shifts = synth.translating_frames(2 * RECS, H, W, 14, seed_texture=7, seed_motion=8)
        mine = np.ascontiguousarray(frames[2 * lo:2 * hi])                 # recording i = frames (2i, 2i+1)
        pinned = torch.empty(mine.shape, dtype=torch.uint8, pin_memory=True)
        pinned.numpy()[:] = mine
        pairs = [(2 * i, 2 * i + 1) for i in range(P)]
        halves = [pairs, [(2 * P + a, 2 * P + b) for a, b in pairs]]       # two sets of ring slots: step k+1 is uploaded while step k searches
        for R in (8, 16, 32, 64, 128):
            est = me.MotionEstimator(H, W, me.SearchConfig(B, R), max_frames=4 * P, max_pairs=P, device=rk.local)
            est.upload(pinned.numpy())
            for _ in range(3):
                est.launch(pairs)
            est.sync()
            rk.barrier()
            l0 = est.launch_count()
            est.timer_start()
            for _ in range(steps):
                est.launch(pairs)
                if gather:
                    ticket = gather.start(est, P)
            ms = est.timer_stop() / steps
            if gather:
                gather.wait(ticket)
            launches += est.launch_count() - l0
            rec = est.fetch_pairs(P)
            exact = R >= 16 or bool(np.all(np.abs(shifts[2 * lo:2 * hi:2]) < R))
            good = (not exact) or bool(np.all(np.rint(rec["pdx"]) == shifts[2 * lo:2 * hi:2, 0]))
            rk.barrier()
            t0 = time.perf_counter()
            est.upload(pinned.numpy(), first_slot=0)                        # e2e: host frames in, host records out, every step
            prev, e2e_rec = None, None
            for s in range(steps):
                est.launch(halves[s & 1])
                est.upload(pinned.numpy(), first_slot=2 * P * ((s + 1) & 1))    # the next step's frames, while this one searches
                ticket = gather.start(est, P) if gather else est.last_ticket()
                if prev is not None:
                    e2e_rec = gather.wait(prev) if gather else est.fetch_pairs(P, ticket=prev)
                prev = ticket
            e2e_rec = gather.wait(prev) if gather else est.fetch_pairs(P, ticket=prev)
            e2e_s = rk.max(time.perf_counter() - t0) / steps
            if not gather:
                good = good and bool(np.array_equal(e2e_rec["pdx"], rec["pdx"]))
            ms = rk.max(ms)
            geo = est.geometry
            work = geo["cand_px_per_pair_valid"] * P
            table.append({"shape": name, "range": R, "pairs_per_s": RECS / (ms * 1e-3), "e2e_pairs_per_s": RECS / e2e_s,
                          "ms_per_step": ms, "frac_int_alu": work / (ms * 1e-3) / peak_alu, "resident_kernel": geo["resident_kernel"],
                          "tile": [geo["tile_blocks_x"], geo["tile_blocks_y"]], "slabs": [geo["slabs_x"], geo["slabs_y"]], "ok": good})
            tot_pairs += RECS * steps
            tot_ms += ms * steps
            tot_e2e_s += e2e_s * steps
            h2d += mine.nbytes
            est.close()
    clocks = sampler.result()
    if rk.rank != 0:
        return None
    worst = min(table, key=lambda r: r["frac_int_alu"])
    return {
        "metric": METRICS["sweep64"], "value": tot_pairs / (tot_ms * 1e-3), "unit": "frame-pairs/s", "n_gpus": rk.world,
        "steps": steps, "warmup": 3, "ms_per_step": tot_ms / (15 * steps), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": NAMES["sweep64"], "recordings": RECS, "recordings_per_gpu": P, "block": B,
                   "shapes": ["1280x720", "1920x1080", "3840x2160"], "ranges": [8, 16, 32, 64, 128],
                   "l2": "inputs larger than L2 for 1080p and 4K (265 MB / 1.06 GB of frames at N = 1); the 720p set (118 MB) is not",
                   "parallelism": f"64 recordings shared out over {rk.world} GPU(s); per step one all_gather of the pair records (N > 1)"},
        "clocks": clocks,
        "e2e": {"value": tot_pairs / tot_e2e_s, "unit": "frame-pairs/s", "h2d_bytes_per_step": h2d // 15, "d2h_bytes_per_step": RECS * 32},
        "gpu_launches": int(launches),
        "roofline": {"bound": "int_alu", "kernel": f"sad_search (worst shape of the sweep: {worst['shape']} +-{worst['range']})",
                     "achieved": worst["frac_int_alu"] * peak_alu / 1e9, "peak": peak_alu / 1e9, "unit": "Gcand*px/s",
                     "frac": worst["frac_int_alu"], "traffic": None,
                     "peak_is": f"{sms} SMs x {ipc:.2f} VABSDIFF4 thread-instr/clk/SM (microbenchmarked here) x 4 px x {f_max / 1e6:.0f} MHz; {peak_src}"},
        "cpu_baseline": None,
        "sweep": table,
        "results_checked": all(r["ok"] for r in table),
    }


# ---------------------------------------------------------------------------------------------------------- clip
def clip_recording(n=4096, seed=11):
    """A synthetic recording of the example clips' shape: 640x64, three quarters of it translating 0.5-6 px per frame
    (sub-pixel), then a static tail (the filament stops while the camera keeps running, as in the examples)."""
    from motion_estimation_b200 import synth
    moving = 3 * n // 4
    frames, x, _ = synth.subpixel_sequence(moving, 64, 640, seed, step_lo=0.5, step_hi=6.0)
    tail = np.repeat(frames[-1:], n - moving, axis=0)
    noise = np.random.default_rng(seed).integers(-1, 2, size=tail.shape)
    tail = np.clip(tail.astype(np.int16) + noise, 0, 255).astype(np.uint8)
    return np.concatenate([frames, tail]), x


def run_clip(args, rk):
    import motion_estimation_b200 as me
    from motion_estimation_b200 import api
    N = 4096
    frames, x = clip_recording(N, seed=11 + rk.rank)                    # one recording per GPU (recordings are what shards)
    stamps = [FRAME_DT_MS * k for k in range(N + 3)]
    cfg = api.default_config
    if os.environ.get("ME_BENCH_RANGE_Y"):              # experiments: a rectangular search window (the recordings move along x)
        cfg = me.SearchConfig(cfg.block, cfg.range, range_y=int(os.environ["ME_BENCH_RANGE_Y"]))
    peaks, peak_src = read_measured_peaks()
    peak_alu, sms, ipc, f_max = alu_peak(me, rk.local, peaks, None)
    api.set_defaults(config=cfg, device=rk.local)

    # ---- value: frames resident in the ring, the anchor walk inside the library (me_anchored_scan), device-timed
    est = me.MotionEstimator(64, 640, cfg, max_frames=N + 1, max_pairs=256, device=rk.local)
    est.upload(frames, first_slot=1)
    est.sync()
    for _ in range(max(args.warmup, 1)):
        est.anchored_scan(1, 2, N - 1)
    rk.barrier()
    sampler = ClockSampler(rk.local)
    sampler.start()
    l0 = est.launch_count()
    est.timer_start()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        rec, searches, launches = est.anchored_scan(1, 2, N - 1)
    wall_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    dev_ms = est.timer_stop() / args.steps
    n_launch = est.launch_count() - l0
    clocks = sampler.result()
    geo = est.geometry
    ms_step = rk.max(dev_ms)
    accepted = int(rec["accepted"].sum())
    est.close()

    # ---- e2e: host frames -> velocity series through api.end_process; with the reference's pre-processing in front, and without
    def pipeline(pre):
        f = frames
        if pre:
            f = api.denoise_hf5(api.adaptive_histogram_equalization(f), n_list=len(f) + 1, distributed=False)   # every rank has its own recording
        return api.end_process(f, firstFrame=0, timestamps_ms=stamps, return_details=True)

    pipeline(True)
    rk.barrier()
    reps = max(1, min(args.steps, 3))
    t0 = time.perf_counter()
    for _ in range(reps):
        res_pre = pipeline(True)
    rk.barrier()
    e2e_pre_s = rk.max(time.perf_counter() - t0) / reps
    t0 = time.perf_counter()
    for _ in range(reps):
        res_raw = pipeline(False)
    rk.barrier()
    e2e_raw_s = rk.max(time.perf_counter() - t0) / reps
    api.close_engines()
    ap = np.array(res_raw.accepted_pairs)
    moving = ap[:, 1] < 3 * N // 4
    d = np.array(res_raw.pair_displacements)[moving, 0]
    err = d - (x[ap[moving, 1]] - x[ap[moving, 0]]) / 2
    if rk.rank != 0:
        return None

    cpu = None
    if rk.world == 1:
        from motion_estimation_b200 import velocity
        from oracle import ecc_port
        n_cpu = 160
        lines = ["# timecode format v2"] + [repr(t) for t in stamps]

        def pair(a, k):
            pdx, pdy, flag = ecc_port.ecc_pair_video(frames[a], frames[k])
            return pdx, pdy, flag, ecc_port.dense_flow_mean(frames[a], frames[k])[0]
        t0 = time.perf_counter()
        velocity.anchored_pair_loop(n_cpu, pair, velocity.timecode_lookup(lines))
        dt = time.perf_counter() - t0
        import cv2
        cpu = {"value": n_cpu / dt, "unit": "frames/s", "cores": cv2.getNumThreads(), "kind": "port",
               "sample": f"the reference's pair loop (pyramid ECC + Farneback per frame, oracle/ecc_port.py on cv2 {cv2.__version__}) on the first "
                         f"{n_cpu} frames, no pre-processing, {dt:.1f} s; its non-local means adds 1.9 s per frame (SURVEY 6)"}
    work = searches * geo["cand_px_per_pair_valid"]
    return {
        "metric": METRICS["clip"], "value": rk.world * N / (ms_step * 1e-3), "unit": "frames/s", "n_gpus": rk.world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": NAMES["clip"], "height": 64, "width": 640, "frames": N, "block": cfg.block, "range": cfg.range,
                   "l2": f"one recording is {N * 64 * 640 / 1e6:.0f} MB of frames (larger than the 126 MB L2)",
                   "parallelism": f"one recording per GPU ({rk.world} in flight); the anchored loop is sequential inside a recording",
                   "searches_per_step": int(searches), "launches_per_step": int(launches), "accepted_pairs": accepted,
                   "host_wall_ms_per_step": wall_ms},
        "clocks": clocks,
        "e2e": {"value": rk.world * N / e2e_pre_s, "unit": "frames/s", "h2d_bytes_per_step": 3 * N * 64 * 640,
                "d2h_bytes_per_step": 2 * N * 64 * 640 + (N - 1) * 40, "ms_per_step": e2e_pre_s * 1e3,
                "what": "host frames -> CLAHE -> non-local means -> anchored pair loop -> velocity rows (the reference's default pipeline)",
                "without_preprocessing": {"value": rk.world * N / e2e_raw_s, "ms_per_step": e2e_raw_s * 1e3,
                                          "h2d_bytes_per_step": N * 64 * 640, "d2h_bytes_per_step": (N - 1) * 40}},
        "gpu_launches": int(n_launch),
        "roofline": {"bound": "int_alu", "kernel": f"sad_search_kernel<16,5,4,{geo['window_pitch']}> in batches of 2-256 pairs",
                     "achieved": work / (dev_ms * 1e-3) / 1e9, "peak": peak_alu / 1e9, "unit": "Gcand*px/s",
                     "frac": work / (dev_ms * 1e-3) / peak_alu, "traffic": None,
                     "peak_is": f"{sms} SMs x {ipc:.2f} VABSDIFF4 thread-instr/clk/SM (microbenchmarked here) x 4 px x {f_max / 1e6:.0f} MHz; {peak_src}",
                     "note": "latency-bound: every batch waits for the one before it (the anchor rule), so the device idles between launches"},
        "cpu_baseline": cpu,
        "subpixel_error_half_res_px": {"mean": float(err.mean()), "std": float(err.std()), "max_abs": float(np.abs(err).max())},
        "rows": {"with_preprocessing": len(res_pre.out_information), "without": len(res_raw.out_information)},
    }


# -------------------------------------------------------------------------------------------------- nlm / clahe
NLM_OPS_PER_OFFSET_PX = 10      # see DESIGN.md 3.4: |a-b|, square, 2+2 sliding-sum updates, shift, table read, 2 accumulates


def run_nlm(args, rk):
    import motion_estimation_b200 as me
    from motion_estimation_b200 import api
    H, W, N, HALO = 64, 640, 256, 4
    frames, _ = clip_recording(N + 2 * HALO, seed=21 + rk.rank)
    windows = [0] * HALO + [9] * N + [0] * HALO                         # halo frames are read only, like the chunks of denoise_hf5
    peaks, peak_src = read_measured_peaks()
    _, sms, ipc, f_max = alu_peak(me, rk.local, peaks, None)
    est = me.MotionEstimator(H, W, me.SearchConfig(16, 16), max_frames=N + 2 * HALO, max_pairs=1, device=rk.local)
    est.upload(frames)
    for _ in range(max(args.warmup, 1)):
        est.denoise(0, windows)
    est.sync()
    rk.barrier()
    sampler = ClockSampler(rk.local)
    sampler.start()
    l0 = est.launch_count()
    est.copy_timer_start()
    for _ in range(args.steps):
        est.denoise(0, windows)                                          # in place; run time does not depend on the content
    ms_step = rk.max(est.copy_timer_stop() / args.steps)
    launches = est.launch_count() - l0
    clocks = sampler.result()
    est.close()
    api.set_defaults(device=rk.local)
    api.denoise_hf5(frames, n_list=len(frames) + 1, distributed=False)          # every rank denoises its own frames (weak scaling)
    rk.barrier()
    t0 = time.perf_counter()
    reps = max(1, min(args.steps, 5))
    for _ in range(reps):
        clean = api.denoise_hf5(frames, n_list=len(frames) + 1, distributed=False)          # host frames in, host frames out
    rk.barrier()
    e2e_s = rk.max(time.perf_counter() - t0) / reps
    sharded = None
    if rk.world > 1:                                                     # strong scaling: ONE recording, its frame ranges (+ 4-frame halo) over the ranks
        same, _ = clip_recording(2048, seed=5)
        api.denoise_hf5(same[:64], n_list=65)                            # warm-up of the sharded path (NCCL buffers)
        rk.barrier()
        t0 = time.perf_counter()
        api.denoise_hf5(same, n_list=len(same) + 1)                      # distributed by default under torch.distributed
        rk.barrier()
        t_sh = rk.max(time.perf_counter() - t0)
        sharded = {"value": len(same) / t_sh, "unit": "frames/s", "frames": len(same), "ms": t_sh * 1e3,
                   "what": "api.denoise_hf5 on one 2048-frame recording, frame ranges sharded over the ranks, cleaned frames all-gathered; host in, host out"}
    api.close_engines()
    if rk.rank != 0:
        return None
    offsets = 9 * 41 * 41
    ops = N * H * W * offsets * NLM_OPS_PER_OFFSET_PX
    peak = sms * ipc * f_max                                             # integer-ALU thread-instructions / s
    cpu = None
    if rk.world == 1:
        import cv2
        n_cpu, t0 = 0, time.perf_counter()
        while n_cpu < 6 and time.perf_counter() - t0 < 12:
            want = cv2.fastNlMeansDenoisingMulti(srcImgs=list(frames[:16]), imgToDenoiseIndex=5 + n_cpu, temporalWindowSize=9, h=15,
                                                 templateWindowSize=19, searchWindowSize=41)
            n_cpu += 1
        dt = time.perf_counter() - t0
        assert np.array_equal(want, clean[5 + n_cpu - 1]), "NLM result differs from cv2"           # the checker, not the product
        cpu = {"value": n_cpu / dt, "unit": "frames/s", "cores": cv2.getNumThreads(), "kind": "reference",
               "sample": f"{n_cpu} frames through cv2.fastNlMeansDenoisingMulti (cv2 {cv2.__version__}), the call the reference makes "
                         f"(estimate_vel_vid.py:226-232), {dt:.1f} s"}
    return {
        "metric": METRICS["nlm"], "value": rk.world * N / (ms_step * 1e-3), "unit": "frames/s", "n_gpus": rk.world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": NAMES["nlm"], "height": H, "width": W, "frames_per_step_per_gpu": N, "h": 15, "template_window": 19,
                   "search_window": 41, "temporal_window": 9, "ms_per_frame": ms_step / N,
                   "l2": "10.8 MB of frames per step: L2-resident by nature of the stage (each frame is read 9 x 1681 times)",
                   "parallelism": f"frame ranges with a 4-frame halo per GPU ({rk.world})"},
        "clocks": clocks,
        "e2e": {"value": rk.world * len(frames) / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": int(frames.nbytes * 1.03),
                "d2h_bytes_per_step": int(frames.nbytes), "ms_per_step": e2e_s * 1e3, "what": "api.denoise_hf5: host frames in, host frames out",
                "one_recording_sharded": sharded},
        "gpu_launches": int(launches),
        "roofline": {"bound": "int_alu", "kernel": "nlm_kernel<19>", "achieved": ops / (ms_step * 1e-3) / 1e12, "peak": peak / 1e12,
                     "unit": "T thread-instr/s", "frac": ops / (ms_step * 1e-3) / peak, "traffic": None,
                     "algorithmic_ops_per_offset_px": NLM_OPS_PER_OFFSET_PX, "offset_px_per_s": N * H * W * offsets / (ms_step * 1e-3),
                     "peak_is": f"{sms} SMs x {ipc:.2f} integer-ALU thread-instr/clk/SM (the VABSDIFF4 microbenchmark of this run: the ALU pipe's "
                                f"issue rate) x {f_max / 1e6:.0f} MHz; {peak_src}"},
        "cpu_baseline": cpu,
    }


def run_clahe(args, rk):
    import motion_estimation_b200 as me
    from motion_estimation_b200 import api, synth
    H, W, N = 1080, 1920, 64
    frames, _ = synth.translating_frames(N, H, W, 16, seed_texture=31 + rk.rank, seed_motion=5)
    peaks, peak_src = read_measured_peaks()
    est = me.MotionEstimator(H, W, me.SearchConfig(16, 16), max_frames=N, max_pairs=1, device=rk.local)
    est.upload(frames)
    for _ in range(args.warmup):
        est.clahe(0, N)
    est.sync()
    rk.barrier()
    sampler = ClockSampler(rk.local)
    sampler.start()
    l0 = est.launch_count()
    reps = 50                                                            # a step is 0.2 ms: repeat it so that the clock sampler sees the region
    est.copy_timer_start()
    for _ in range(args.steps * reps):
        est.clahe(0, N)                                                  # 132.7 MB of frames per pass > L2
    ms_step = rk.max(est.copy_timer_stop() / (args.steps * reps))
    launches = est.launch_count() - l0
    clocks = sampler.result()
    est.close()
    api.set_defaults(device=rk.local)
    out = api.adaptive_histogram_equalization(frames)
    rk.barrier()
    t0 = time.perf_counter()
    reps = max(1, min(args.steps, 5))
    for _ in range(reps):
        out = api.adaptive_histogram_equalization(frames)
    rk.barrier()
    e2e_s = rk.max(time.perf_counter() - t0) / reps
    api.close_engines()
    if rk.rank != 0:
        return None
    alg = 3 * N * H * W + N * 64 * 256 * 2
    hbm = float(peaks["hbm_gbs"])
    cpu = None
    if rk.world == 1:
        import cv2
        cl = cv2.createCLAHE(clipLimit=2.0, tileGridSize=(8, 8))
        t0, n_cpu = time.perf_counter(), 0
        while time.perf_counter() - t0 < 10:
            want = cl.apply(frames[n_cpu % N])
            assert np.array_equal(want, out[n_cpu % N]), "CLAHE result differs from cv2"            # the checker, not the product
            n_cpu += 1
        dt = time.perf_counter() - t0
        cpu = {"value": n_cpu / dt, "unit": "frames/s", "cores": cv2.getNumThreads(), "kind": "reference",
               "sample": f"{n_cpu} frames through cv2.createCLAHE(2.0, (8,8)).apply (cv2 {cv2.__version__}), the call the reference makes "
                         f"(estimate_vel_vid.py:209-213), {dt:.1f} s"}
    return {
        "metric": METRICS["clahe"], "value": rk.world * N / (ms_step * 1e-3), "unit": "frames/s", "n_gpus": rk.world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": NAMES["clahe"], "height": H, "width": W, "frames_per_step_per_gpu": N, "clip_limit": 2.0, "tiles": [8, 8],
                   "l2": f"inputs larger than L2: {N * H * W / 1e6:.1f} MB of frames per step", "parallelism": f"frames per GPU ({rk.world})"},
        "clocks": clocks,
        "e2e": {"value": rk.world * N / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": N * H * W, "d2h_bytes_per_step": N * H * W,
                "ms_per_step": e2e_s * 1e3, "what": "api.adaptive_histogram_equalization: host frames in, host frames out"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "clahe_lut_kernel + clahe_apply_kernel", "achieved": alg / (ms_step * 1e-3) / 1e9, "peak": hbm,
                     "unit": "GB/s", "frac": alg / (ms_step * 1e-3) / 1e9 / hbm, "algorithmic_bytes_per_launch": alg,
                     # ncu --set full of the three kernels of one 64 x 1080p pass (profiles/r02_clahe_ncu_summary.txt): 132.73 + 3.90,
                     # 1.05, 143.39 + 73.83 MB -- below the algorithmic 3 B/px because part of the apply pass's re-read hits L2
                     "traffic": 354898432 if (H, W, N) == (1080, 1920, 64) else None,
                     "peak_is": f"copy bandwidth, {peak_src}"},
        "cpu_baseline": cpu,
    }


# ---------------------------------------------------------------------------------------- reference arm (host CPU)
def run_reference(args):
    """The reference's CPU implementation of the workload on the host cores, rank 0 only: its per-pair function (4-pass x
    4-level pyramid cv.findTransformECC, estimate_vel_vid.py:306-369) for the search workloads, its pair loop for `clip`,
    the cv2 calls it makes for `nlm` / `clahe`.  Each step is a bounded sample of the workload."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import cv2 as cv
    from motion_estimation_b200 import synth
    from oracle import ecc_port
    wl = args.workload
    cores = cv.getNumThreads()
    base = {"impl": "reference", "metric": METRICS[wl], "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic"}
    if wl in ("1080p", "strong", "4k", "sweep64"):
        H, W = (2160, 3840) if wl == "4k" else (1080, 1920)
        rng = 64 if wl == "4k" else 32
        frames, shifts = synth.translating_frames(9, H, W, rng, seed_texture=0, seed_motion=1)
        per_step = 1 if wl == "4k" else 2
        pairs = [(k % 8, k % 8 + 1) for k in range((args.warmup + args.steps) * per_step)]
        for a, b in pairs[:args.warmup * per_step]:
            ecc_port.ecc_pair_video(frames[a], frames[b])
        t0, err = time.perf_counter(), 0.0
        for a, b in pairs[args.warmup * per_step:]:
            pdx, pdy, flag = ecc_port.ecc_pair_video(frames[a], frames[b])
            if flag:
                err = max(err, abs(2 * float(pdx) - shifts[a][0]), abs(2 * float(pdy) - shifts[a][1]))
        dt = time.perf_counter() - t0
        v, unit = args.steps * per_step / dt, "frame-pairs/s"
        sample = (f"{per_step} pair(s) of {W}x{H} per step through oracle/ecc_port.py = the reference's calc_ECC_transform (cv2 {cv.__version__}, "
                  f"{cores} threads); max |2*ecc - true shift| = {err:.3f} px")
        cfg = {"workload": NAMES[wl], "height": H, "width": W,
               "note": "the reference has no block search; this is its own per-pair function on frames of the same shape"}
    elif wl == "clip":
        from motion_estimation_b200 import velocity
        frames, _ = clip_recording(4096, seed=11)
        n = 40
        stamps = ["# timecode format v2"] + [repr(FRAME_DT_MS * k) for k in range(n * (args.steps + args.warmup) + 8)]

        def pair(a, k):
            pdx, pdy, flag = ecc_port.ecc_pair_video(frames[a], frames[k])
            return pdx, pdy, flag, ecc_port.dense_flow_mean(frames[a], frames[k])[0]
        dt = 0.0
        for s in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            velocity.anchored_pair_loop(n, lambda a, k, o=s * n: pair(a + o, k + o), velocity.timecode_lookup(stamps))
            if s >= args.warmup:
                dt += time.perf_counter() - t0
        v, unit = args.steps * n / dt, "frames/s"
        sample = f"{n} frames per step through the reference's pair loop (ECC + Farneback per frame), no pre-processing, {cores} threads"
        cfg = {"workload": NAMES[wl], "height": 64, "width": 640}
    elif wl == "nlm":
        frames, _ = clip_recording(64, seed=21)
        dt = 0.0
        for s in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            cv.fastNlMeansDenoisingMulti(srcImgs=list(frames[:24]), imgToDenoiseIndex=4 + s % 16, temporalWindowSize=9, h=15,
                                         templateWindowSize=19, searchWindowSize=41)
            if s >= args.warmup:
                dt += time.perf_counter() - t0
        v, unit = args.steps / dt, "frames/s"
        sample = f"one 640x64 frame per step through cv2.fastNlMeansDenoisingMulti(h=15, 19, 41, 9 frames), {cores} threads"
        cfg = {"workload": NAMES[wl], "height": 64, "width": 640}
    else:
        frames, _ = synth.translating_frames(8, 1080, 1920, 16, seed_texture=31, seed_motion=5)
        cl = cv.createCLAHE(clipLimit=2.0, tileGridSize=(8, 8))
        dt = 0.0
        for s in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            for f in frames:
                cl.apply(f)
            if s >= args.warmup:
                dt += time.perf_counter() - t0
        v, unit = args.steps * len(frames) / dt, "frames/s"
        sample = f"8 frames of 1080p per step through cv2.createCLAHE(2.0, (8,8)).apply, {cores} threads"
        cfg = {"workload": NAMES[wl], "height": 1080, "width": 1920}
    base.update({"value": v, "unit": unit, "ms_per_step": dt * 1e3 / args.steps, "config": cfg,
                 "cpu_baseline": {"value": v, "unit": unit, "cores": cores, "kind": "port" if wl not in ("nlm", "clahe") else "reference",
                                  "sample": sample},
                 "e2e": {"value": v, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    print(json.dumps(base))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=("ours", "reference"), default="ours")
    ap.add_argument("--workload", choices=WORKLOADS, default="1080p")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    args.warmup = max(args.warmup, 3)
    rk = Ranks(args.gpus)
    if args.workload in SEARCH_SHAPES:
        line = run_search(args, rk, args.workload)
    else:
        line = {"sweep64": run_sweep, "clip": run_clip, "nlm": run_nlm, "clahe": run_clahe}[args.workload](args, rk)
    if line is not None:
        print(json.dumps(line))
    rk.barrier()
    rk.close()


if __name__ == "__main__":
    main()

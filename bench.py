#!/usr/bin/env python3
"""bench.py -- frame-pairs/s of the displacement search on BASELINE.json's headline workload.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched by torch.distributed.run)
    python bench.py --impl reference --gpus N --steps K --warmup W

Workload (configs[2]): synthetic 1920x1080 uint8 luma video, 16x16 blocks, +-32 px full search, the
estimate_vel_vid.py path.  One STEP = one batch of 64 consecutive frame pairs (65 frames, 134.8 MB --
larger than the 126 MB L2) per GPU: search + per-pair median on the GPU, the records of all ranks
gathered (NCCL) and reduced to a velocity series on rank 0.

  value   pairs/s with the frames already resident in HBM (device time, CUDA events, max over ranks)
  e2e     pairs/s through the public API with HOST frames: pinned host -> device copy of the step's 65
          frames and the device -> host read of its records inside the timed region (copy of step k+1
          overlaps the search of step k on the copy stream)
  roofline        the search kernel against the measured VABSDIFF4 issue rate (integer ALU), and HBM
  cpu_baseline    the C oracle (oracle/sad_oracle.c, OpenMP) on a bounded sample of the same pairs

The reference arm runs the reference's own per-pair algorithm (pyramid ECC on cv2, restated in
oracle/ecc_port.py because /root/reference is Python and does not travel) on the host cores.
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

H, W, BLOCK, RANGE, PAIRS = 1080, 1920, 16, 32, 64
PIXELS_PER_MM, FRAME_DT_MS = 611.0, 5.531
METRIC = "frame-pairs/sec at 1080p, 16x16 blocks, +-32 px full-search SAD"
WORKLOAD = "configs[2]: synthetic 1080p uint8 luma video, 16x16 blocks, +-32 px full search, estimate_vel_vid.py path"


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


def make_frames(seed):
    from motion_estimation_b200 import synth
    return synth.translating_frames(PAIRS + 1, H, W, RANGE, seed_texture=seed, seed_motion=seed + 1)


def velocity_series(records, n_ranks):
    """Rank 0's reduction of the gathered records to mm/min (estimate_vel_vid.py:556-560 arithmetic)."""
    gap_min = FRAME_DT_MS / 1000.0 / 60.0
    pdx = records["pdx"].astype(np.float64).reshape(n_ranks, -1)
    return (pdx / PIXELS_PER_MM) / gap_min


# --------------------------------------------------------------------------------------------- ours
def run_ours(args):
    import torch
    import motion_estimation_b200 as me
    from motion_estimation_b200 import shard

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch N>1 with torch.distributed.run")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    cfg = me.SearchConfig(block=BLOCK, range=RANGE)
    est = me.MotionEstimator(H, W, cfg, max_frames=2 * (PAIRS + 1), max_pairs=PAIRS, device=local)
    geo = est.geometry
    frames_np, shifts = make_frames(seed=10 * rank)
    pinned = torch.empty((PAIRS + 1, H, W), dtype=torch.uint8, pin_memory=True)
    pinned.numpy()[:] = frames_np
    host_frames = pinned.numpy()
    half = [[(k, k + 1) for k in range(PAIRS)], [(PAIRS + 1 + k, PAIRS + 2 + k) for k in range(PAIRS)]]
    gather_buf = [torch.empty((PAIRS, shard.PAIR_FIELDS), dtype=torch.int32, device="cuda") for _ in range(world)]

    def gather(rec):
        """all ranks' records -> every rank (the path's one exchange step); rank 0 reduces to velocity."""
        if dist is None:
            return rec
        t = torch.from_numpy(shard.records_to_int32(rec)).cuda()
        dist.all_gather(gather_buf, t)
        return shard.int32_to_records(torch.stack(gather_buf).cpu().numpy(), me.PAIR_DTYPE)

    # measured integer-SAD issue rate for the roofline denominator
    ipc, mhz_seen, sms = me.microbench_vabsdiff4(local, 0)

    # ---- value: frames resident in HBM
    est.upload(host_frames, first_slot=0)
    est.sync()
    for _ in range(args.warmup):
        est.launch(half[0])
    est.sync()
    rec = est.fetch_pairs(PAIRS)
    ok = bool(np.all(np.rint(rec["pdx"]) == shifts[:, 0]) and np.all(np.rint(rec["pdy"]) == shifts[:, 1]) and rec["flag"].all())
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = est.launch_count()
    t0 = time.perf_counter()
    est.timer_start()
    for _ in range(args.steps):
        est.launch(half[0])
    dev_ms = est.timer_stop()
    rec = est.fetch_pairs(PAIRS)
    all_rec = gather(rec)
    barrier()
    wall_s = time.perf_counter() - t0
    launches = est.launch_count() - launches0
    clocks = sampler.result()
    ms_step = max_over_ranks(dev_ms / args.steps)
    wall_ms_step = max_over_ranks(wall_s * 1e3 / args.steps)
    value = world * PAIRS / (ms_step * 1e-3)
    kernel_ms = []
    for _ in range(3):
        est.launch(half[0])
        est.sync()
        kernel_ms.append(est.last_search_ms())
    vel = velocity_series(all_rec, world) if rank == 0 else None

    # ---- e2e: host frames in, host records out, every step
    def e2e_steps(n):
        est.upload(host_frames, first_slot=0)
        out = None
        for s in range(n):
            est.launch(half[s & 1])
            est.upload(host_frames, first_slot=(PAIRS + 1) * ((s + 1) & 1))   # next step's input, overlaps this search
            out = gather(est.fetch_pairs(PAIRS))
        return out

    e2e_steps(max(args.warmup, 2))
    barrier()
    sampler2 = ClockSampler(local)
    sampler2.start()
    launches1 = est.launch_count()
    t0 = time.perf_counter()
    e2e_rec = e2e_steps(args.steps)
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e_launches = est.launch_count() - launches1
    clocks2 = sampler2.result()
    e2e_value = world * PAIRS * args.steps / e2e_s
    ok = ok and bool(np.array_equal(e2e_rec["pdx"][:PAIRS], rec["pdx"]))

    if rank != 0:
        est.close()
        if dist is not None:
            dist.destroy_process_group()
        return

    peaks, peak_src = read_measured_peaks()
    f_max = float(peaks.get("sm_max_mhz") or clocks.get("sm_max_mhz") or 1965.0) * 1e6
    k_ms = float(np.mean(kernel_ms))
    cand_valid, cand_nominal = geo["cand_px_per_pair_valid"] * PAIRS, geo["cand_px_per_pair"] * PAIRS
    peak_alu = sms * ipc * 4.0 * f_max                         # cand*px/s: one VABSDIFF4 = 4 |a-b|
    achieved = cand_valid / (ms_step * 1e-3)
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get("dram_bytes_per_launch")
    except Exception:
        pass
    alg_bytes = (PAIRS + 1) * H * W + PAIRS * geo["nbx"] * geo["nby"] * 17 + PAIRS * 24
    hbm_gbs = float(peaks["hbm_gbs"])
    roofline = {
        "bound": "int_alu", "kernel": (f"sad_search_persistent<16,5,4,{geo['window_pitch']}>" if geo.get("resident_kernel") else f"sad_search_kernel<16,5,4,{geo['window_pitch']}>"), "achieved": achieved / 1e9, "peak": peak_alu / 1e9,
        "unit": "Gcand*px/s", "frac": achieved / peak_alu, "traffic": traffic,
        "peak_is": f"{sms} SMs x {ipc:.2f} VABSDIFF4 thread-instr/clk/SM (microbenchmarked in this run) x 4 px x "
                   f"{f_max / 1e6:.0f} MHz max SM clock; {peak_src}",
        "frac_at_observed_clock": achieved / (sms * ipc * 4.0 * (clocks["sm_mhz"] or f_max / 1e6) * 1e6),
        "achieved_nominal": cand_nominal / (ms_step * 1e-3) / 1e9,
        "kernel_ms_per_launch": k_ms, "kernel_share_of_step": k_ms / ms_step,
        "hbm": {"achieved": alg_bytes / (ms_step * 1e-3) / 1e9, "peak": hbm_gbs, "unit": "GB/s",
                "frac": alg_bytes / (ms_step * 1e-3) / 1e9 / hbm_gbs, "algorithmic_bytes_per_launch": alg_bytes},
    }

    cpu = None
    if world == 1:
        from oracle import sad as osad
        n, t_cpu = 0, time.perf_counter()
        while n < 24 and (time.perf_counter() - t_cpu < 12.0 or n < 2):
            o = osad.search(frames_np[n], frames_np[n + 1], BLOCK, RANGE)
            ok = ok and o["pdx"] == rec["pdx"][n] and o["pdy"] == rec["pdy"][n]     # the checker, not the product
            n += 1
        t_cpu = time.perf_counter() - t_cpu
        cpu = {"value": n / t_cpu, "unit": "frame-pairs/s", "cores": osad.max_threads(), "kind": "port",
               "sample": f"first {n} of the step's {PAIRS} pairs through oracle/sad_oracle.c (OpenMP), {t_cpu:.1f} s",
               "gcand_px_per_s": n * geo["cand_px_per_pair_valid"] / t_cpu / 1e9}

    line = {
        "metric": METRIC, "value": value, "unit": "frame-pairs/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "height": H, "width": W, "block": BLOCK, "range": RANGE,
                   "pairs_per_step_per_gpu": PAIRS, "frames_per_step_per_gpu": PAIRS + 1,
                   "l2": f"inputs larger than L2: {(PAIRS + 1) * H * W / 1e6:.1f} MB of frames per step vs 126 MB",
                   "parallelism": f"frame pairs sharded over {world} GPU(s), one all_gather of pair records per step",
                   "tile_blocks_x": geo["tile_blocks_x"], "threads_per_cta": geo["threads_per_cta"],
                   "window_pitch": geo["window_pitch"], "candidate_slabs": [geo["slabs_x"], geo["slabs_y"]],
                   "smem_bytes": geo["smem_bytes"], "ctas_per_pair": geo["tiles_per_pair"]},
        "gcand_px_per_s": achieved / 1e9,
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "frame-pairs/s", "h2d_bytes_per_step": (PAIRS + 1) * H * W,
                "d2h_bytes_per_step": PAIRS * 24, "ms_per_step": e2e_s * 1e3 / args.steps, "clocks": clocks2,
                "gpu_launches": int(e2e_launches)},
        "gpu_launches": int(launches),
        "wall_ms_per_step": wall_ms_step,
        "roofline": roofline,
        "cpu_baseline": cpu,
        "results_checked": bool(ok),
        "velocity_mm_min_first3": [float(v) for v in vel[0][:3]],
    }
    print(json.dumps(line))
    est.close()
    if dist is not None:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------- reference
def run_reference(args):
    """The reference's CPU implementation of the path: pyramid ECC per pair (estimate_vel_vid.py:306-369),
    run on the host cores with cv2's own thread pool.  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import cv2 as cv
    from oracle import ecc_port
    frames, shifts = make_frames(seed=0)
    per_step = 2
    pairs = [(k % PAIRS, k % PAIRS + 1) for k in range((args.warmup + args.steps) * per_step)]
    for a, b in pairs[:args.warmup * per_step]:
        ecc_port.ecc_pair_video(frames[a], frames[b])
    t0 = time.perf_counter()
    err = 0.0
    for a, b in pairs[args.warmup * per_step:]:
        pdx, pdy, flag = ecc_port.ecc_pair_video(frames[a], frames[b])
        if flag:
            err = max(err, abs(2 * float(pdx) - shifts[a][0]), abs(2 * float(pdy) - shifts[a][1]))
    dt = time.perf_counter() - t0
    v = args.steps * per_step / dt
    cores = cv.getNumThreads()
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "frame-pairs/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3 / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "height": H, "width": W,
                   "note": "the reference has no block search; this is its own per-pair function (4-pass x 4-level "
                           "pyramid cv.findTransformECC) on the same frames"},
        "cpu_baseline": {"value": v, "unit": "frame-pairs/s", "cores": cores, "kind": "port",
                         "sample": f"{per_step} pairs per step through oracle/ecc_port.py (cv2 {cv.__version__}, "
                                   f"{cores} threads); max |2*ecc - true shift| = {err:.3f} px"},
        "e2e": {"value": v, "unit": "frame-pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=("ours", "reference"), default="ours")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

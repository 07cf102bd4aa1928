"""First-contact GPU script: microbenchmark, parity against the oracle, a timing probe.
Run on the GPU box:  python tools/gpu_check.py  (writes gpurun_out/gpu_check.log when redirected)."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import motion_estimation_b200 as me
from motion_estimation_b200 import synth
from oracle import sad as osad


def parity(H, W, B, R, n=3, shift=None, **kw):
    frames, shifts = synth.translating_frames(n, H, W, shift if shift is not None else R)
    cfg = me.SearchConfig(block=B, range=R, **kw)
    with me.MotionEstimator(H, W, cfg, max_frames=n, max_pairs=n) as est:
        est.upload(frames)
        pairs = [(k, k + 1) for k in range(n - 1)] + [(n - 1, 0)]
        res = est.search(pairs)
        bad = 0
        for i, (a, c) in enumerate(pairs):
            o = osad.search(frames[a], frames[c], B, R, min_range=cfg.min_range, subpixel=cfg.subpixel,
                            min_valid=cfg.min_valid, min_curv=cfg.min_curv)
            g = est.fetch_blocks(i)
            for k in ("mv", "sad", "sub", "flags"):
                if not np.array_equal(o[k], g[k]):
                    bad += 1
                    d = np.argwhere(o[k] != g[k])
                    print(f"  MISMATCH {k} pair {i}: {len(d)} entries, first {d[:3].tolist()} oracle {o[k][tuple(d[0])]} gpu {g[k][tuple(d[0])]}")
            r = res[i]
            if not (r["pdx"] == o["pdx"] and r["pdy"] == o["pdy"] and r["flag"] == o["flag"]
                    and r["nvalid_x"] == o["nvalid_x"] and r["nvalid_y"] == o["nvalid_y"]):
                bad += 1
                print(f"  MISMATCH pair record {i}: gpu {r} oracle {o['pdx']},{o['pdy']},{o['flag']},{o['nvalid_x']},{o['nvalid_y']}")
        print(f"parity H={H} W={W} B={B} R={R} tma={est.geometry['uses_tma']} tbx={est.geometry['tile_blocks_x']} "
              f"thr={est.geometry['threads_per_cta']} smem={est.geometry['smem_bytes']}: {'OK' if not bad else 'FAIL'}  "
              f"(pdx={res[0]['pdx']:.3f} true {shifts[0].tolist()})", flush=True)
        return bad == 0


def timing(H, W, B, R, P=16, reps=5):
    frames, _ = synth.translating_frames(P + 1, H, W, R)
    cfg = me.SearchConfig(block=B, range=R)
    with me.MotionEstimator(H, W, cfg, max_frames=P + 1, max_pairs=P) as est:
        est.upload(frames)
        pairs = [(k, k + 1) for k in range(P)]
        est.search(pairs)
        ms = []
        for _ in range(reps):
            est.launch(pairs); est.sync(); ms.append(est.last_search_ms())
        g = est.geometry
        best = min(ms)
        print(f"timing {W}x{H} B={B} R={R} P={P}: kernel ms {['%.3f' % m for m in ms]}  pairs/s {P / best * 1e3:.1f}  "
              f"Gcand.px/s nominal {g['cand_px_per_pair'] * P / best / 1e6:.1f} valid {g['cand_px_per_pair_valid'] * P / best / 1e6:.1f} "
              f"geom {g}", flush=True)


if __name__ == "__main__":
    print(me._lib.lib().me_version().decode())
    for mode in range(2):
        ipc, mhz, sms = me.microbench_vabsdiff4(0, mode)
        print(f"microbench mode {mode}: {ipc:.2f} VABSDIFF4 thread-instr/clk/SM at {mhz:.0f} MHz on {sms} SMs", flush=True)
    ok = True
    os.environ["ME_NO_TMA"] = "1"
    ok &= parity(64, 96, 16, 8)
    ok &= parity(64, 640, 16, 32)
    os.environ["ME_NO_TMA"] = "0"
    os.environ.pop("ME_NO_TMA")
    ok &= parity(64, 96, 16, 8)
    ok &= parity(64, 640, 16, 32)
    ok &= parity(70, 100, 16, 8)          # ragged frame
    ok &= parity(72, 200, 8, 12)
    ok &= parity(96, 160, 32, 16)
    ok &= parity(270, 480, 16, 32)
    ok &= parity(136, 256, 8, 64, shift=20)
    print("ALL PARITY", "OK" if ok else "FAILED", flush=True)
    timing(1080, 1920, 16, 32)
    timing(2160, 3840, 8, 64, P=2, reps=3)
    timing(64, 640, 16, 32, P=64)

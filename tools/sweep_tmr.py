"""Kernel-only time over block rows per tile (ME_TMR) for small ranges: python tools/sweep_tmr.py"""
import os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for shape in ("1080 1920 16 8 64", "1080 1920 16 16 32", "1080 1920 16 24 32", "1080 1920 8 8 32", "1080 1920 8 16 16", "1080 1920 32 8 64",
              "1080 1920 32 16 32", "720 1280 16 8 64", "720 1280 16 16 32", "2160 3840 16 8 16", "2160 3840 16 16 16", "1080 1920 8 32 16"):
    for tmr in ("", "1", "2", "3", "4", "6", "8"):
        env = dict(os.environ)
        if tmr:
            env["ME_TMR"] = tmr
        r = subprocess.run([sys.executable, os.path.join(root, "tools", "run_one.py")] + shape.split(), env=env, capture_output=True, text=True)
        line = (r.stdout.strip().split("\n") or [""])[-1]
        import re
        m = re.search(r"ms=([\d.]+) frac=([\d.]+).*'tile_blocks_x': (\d+), 'threads_per_cta': (\d+), 'smem_bytes': (\d+).*'window_pitch': (\d+).*'tile_blocks_y': (\d+)", line)
        print(shape, "| ME_TMR=%-2s" % (tmr or "-"), "|", ("ms=%s frac=%s tbx=%s thr=%s smem=%s pitch=%s tby=%s" % m.groups()) if m else (r.stdout[-200:] + r.stderr[-300:]), flush=True)

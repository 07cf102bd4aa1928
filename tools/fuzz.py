"""More seeds of tests/test_gpu_parity.py::test_random_geometries_equal_oracle than the suite runs:
    python tools/fuzz.py [first_seed] [last_seed]      (GPU box; prints every failing seed)"""
import os, sys, importlib.util
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)
import numpy as np
import motion_estimation_b200 as me
from motion_estimation_b200 import synth
spec = importlib.util.spec_from_file_location("parity", os.path.join(root, "tests", "test_gpu_parity.py"))
parity = importlib.util.module_from_spec(spec)
spec.loader.exec_module(parity)

lo = int(sys.argv[1]) if len(sys.argv) > 1 else 40
hi = int(sys.argv[2]) if len(sys.argv) > 2 else 400
bad = []
for seed in range(lo, hi):
    H, W, B, R, n_pairs, mode, kw = parity._fuzz_case(seed)
    os.environ.pop("ME_PERSISTENT", None)
    if mode:
        os.environ["ME_PERSISTENT"] = "2" if mode == 1 else "0"
    n = 4
    frames, _ = synth.translating_frames(n, H, W, min(R, 20), seed_texture=seed, seed_motion=seed + 1)
    cfg = me.SearchConfig(block=B, range=R, **kw)
    rng = np.random.default_rng(seed)
    pairs = [tuple(int(v) for v in rng.integers(0, n, 2)) for _ in range(n_pairs)]
    try:
        with me.MotionEstimator(H, W, cfg, max_frames=n, max_pairs=max(n_pairs, 2)) as est:
            est.upload(frames)
            res = est.search(pairs)
            for i in sorted(set(int(v) for v in rng.integers(0, n_pairs, 3))):
                a, c = pairs[i]
                parity.check_pair(est, i, frames[a], frames[c], cfg, res[i])
    except Exception as e:                                   # noqa: BLE001  (report and go on)
        bad.append(seed)
        print(f"seed {seed}: {H}x{W} B={B} R={R} pairs={n_pairs} mode={mode} {kw}: {type(e).__name__}: {str(e)[:200]}", flush=True)
print(f"fuzz seeds {lo}..{hi - 1}: {len(bad)} failing {bad}")
sys.exit(1 if bad else 0)

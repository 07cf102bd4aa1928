#!/bin/bash
# GPU box: the whole GPU suite, then the anchored-walk bench with the phase trace (square and rectangular window; look-ahead fitted to the
# grid and not) and the headline bench.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for fit in 1 0; do
for ry in "" 8; do
    ME_SCAN_FIT=$fit ME_BENCH_RANGE_Y=$ry ME_SCAN_TRACE=86 python bench.py --workload clip --steps 3 --warmup 1 2>gpurun_out/clip_ry$ry.err | python -c "
import sys,json;d=json.loads(sys.stdin.read().strip().split(chr(10))[-1]);print('clip fit=$fit ry=$ry',round(d['value']),d['ms_per_step'])"
    grep "scan trace" gpurun_out/clip_ry$ry.err | tail -3
done
done
python bench.py --steps 10 --warmup 3 2>/dev/null | python -c "
import sys,json;d=json.loads(sys.stdin.read().strip().split(chr(10))[-1]);print('1080p',round(d['value']),d['roofline']['frac'],round(d['e2e']['value']))"

#!/bin/bash
# GPU box: the tests that walk the anchor rule, then the anchored-walk bench with the phase trace (square and rectangular window).
mkdir -p gpurun_out
python -m pytest tests/test_gpu_api.py tests/test_gpu_recordings.py tests/test_gpu_cli.py -x -q 2>&1 | tail -2
for ry in "" 8; do
    ME_BENCH_RANGE_Y=$ry ME_SCAN_TRACE=${1:-86} python bench.py --workload clip --steps 3 --warmup 1 2>gpurun_out/clip_ry$ry.err | python -c "
import sys,json;d=json.loads(sys.stdin.read().strip().split(chr(10))[-1]);print('clip ry=$ry',round(d['value']),d['ms_per_step'])"
    grep "scan trace" gpurun_out/clip_ry$ry.err | tail -3
done

#!/bin/bash
# GPU box: every bench workload at N = 1 (JSON lines into gpurun_out/), after the GPU tests.
# usage: tools/gpu_bench_all.sh [tag]   -- run through gpurun from the repository root
tag=${1:-r02}
out=gpurun_out
mkdir -p $out
python -m pytest tests -m gpu -q > $out/${tag}_gputest.log 2>&1
tail -5 $out/${tag}_gputest.log
python bench.py --steps 20 --warmup 3 > $out/${tag}_bench_1080p_1gpu.json 2> $out/${tag}_bench_1080p.err
for w in strong 4k sweep64 clip nlm clahe; do
  python bench.py --steps 10 --warmup 3 --workload $w > $out/${tag}_bench_${w}_1gpu.json 2> $out/${tag}_bench_${w}.err
  echo "$w rc=$?"; tail -3 $out/${tag}_bench_${w}.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/*_bench_*_1gpu.json')):
    try:
        d=json.loads(open(f).read().strip().split('\n')[-1])
        print(f.split('/')[-1], round(d['value'],1), d['unit'], 'e2e', round(d['e2e']['value'],1), 'frac', round(d['roofline']['frac'],3), 'ms', round(d['ms_per_step'],3), d.get('results_checked'), d['clocks']['sm_mhz'], d['clocks']['reasons'])
    except Exception as e:
        print(f, 'unreadable', e)
PY

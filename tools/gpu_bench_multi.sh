#!/bin/bash
# GPU box with N GPUs: bench workloads under torch.distributed.run, JSON lines into gpurun_out/.
# usage: tools/gpu_bench_multi.sh N tag workload...
n=$1; tag=$2; shift 2
out=gpurun_out; mkdir -p $out
port=29700
for w in "$@"; do
  port=$((port+1))
  steps=10; [ "$w" = "1080p" ] && steps=20
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --steps $steps --warmup 3 --workload $w \
      > $out/${tag}_bench_${w}_${n}gpu.json 2> $out/${tag}_bench_${w}_${n}gpu.err
  echo "$w N=$n rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*\|NCCL version" $out/${tag}_bench_${w}_${n}gpu.err | tail -3
  python - <<PY
import json
try:
    d=json.loads(open('$out/${tag}_bench_${w}_${n}gpu.json').read().strip().split('\n')[-1])
    print('   value',round(d['value'],1),d['unit'],'e2e',round(d['e2e']['value'],1),'ms',round(d['ms_per_step'],3),'checked',d.get('results_checked'),'clocks',d['clocks']['sm_mhz'],d['clocks']['reasons'])
except Exception as e:
    print('   unreadable',e)
PY
done

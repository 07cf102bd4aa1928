#!/bin/bash
# Round-end evidence run on the GPU box: tests, bench (both arms), ncu launch list + full capture of the bench command,
# kernel-only sweep, pre-processing timings.  Everything lands in gpurun_out/; summaries are copied to profiles/ by hand.
set -u
cd "$(dirname "$0")/.."
O=gpurun_out
python -m pytest tests -x -q -m gpu > $O/pt.log 2>&1; tail -2 $O/pt.log
python bench.py --steps 10 --warmup 3 > $O/bench.json 2> $O/bench.err; tail -c 600 $O/bench.json
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
if [ -s $O/bench.json ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 > $O/ncu_launches.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:sad_search -c 1 -s 4 -o $O/prof_final -f python bench.py --steps 2 --warmup 3 > $O/ncu_full.log 2>&1
fi
python tools/sweep.py > $O/sweep.csv 2> $O/sweep.err
python tools/time_ingest.py > $O/ingest.log 2>&1
python tools/time_nlm.py 64 640 64 9 smooth >> $O/ingest.log 2>&1
python tools/time_nlm.py 1080 1920 4 1 smooth >> $O/ingest.log 2>&1
cat $O/ingest.log

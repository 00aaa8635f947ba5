#!/bin/bash
# tests + bench + per-launch list of one step.  Usage: gpurun --timeout 900 -- tools/gpu_quick.sh <tag> [pytest -k expr]
tag=${1:-q}; out=gpurun_out; mkdir -p $out
if [ -n "$2" ]; then python -m pytest tests -m gpu -x -q -k "$2" 2>&1 | tail -3; else python -m pytest tests -m gpu -x -q 2>&1 | tail -3; fi
python bench.py --steps 5 --warmup 3 > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench exit $?"
python -c "
import json; d=json.load(open('$out/bench_$tag.json')); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['kernel_ms_per_step'], d['clocks'])"
CMD="python bench.py --profile-range"
$CMD > $out/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 4000 --csv --log-file $out/launches_$tag.csv $CMD > $out/ncu1_$tag.log 2>&1
echo "ncu launches exit $?"
python tools/launch_summary.py $out/launches_$tag.csv | head -12

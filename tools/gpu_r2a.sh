#!/bin/bash
# round 2, first GPU call: all GPU tests (with output of the new BASELINE-shape tests), smoke, bench, launch list, memory-kernel ncu pass
tag=${1:-r2a}; out=gpurun_out; mkdir -p $out
python -m pytest tests -m gpu -x -q -s > $out/pytest_$tag.log 2>&1; echo "pytest exit $?"; tail -4 $out/pytest_$tag.log
grep -E "C2|C3|C4|C5|un-forced|pad-leak|engineered|spread" $out/pytest_$tag.log | head -40
( time python bench.py --steps 10 --warmup 3 > $out/bench_$tag.json 2> $out/bench_$tag.err ) 2>&1 | grep real; tail -c 1500 $out/bench_$tag.err
python -c "
import json; d=json.load(open('$out/bench_$tag.json')); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['kernel_ms_per_step'], d['clocks']); print(d['cpu_baseline']); print(d['gpu_eager_baseline']); print(d['latency_p50_ms'])"
CMD="python bench.py --profile-range"
$CMD > $out/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 4000 --csv --log-file $out/launches_$tag.csv $CMD > $out/ncu1_$tag.log 2>&1
echo "ncu launches exit $?"
python tools/launch_summary.py $out/launches_$tag.csv | head -14
tools/gpu_ncu_mem.sh $tag

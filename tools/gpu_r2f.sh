#!/bin/bash
tag=${1:-r2f}
tools/gpu_tests.sh $tag > gpurun_out/tests_$tag.txt 2>&1; grep -E "passed|failed|FAILED|^E  " gpurun_out/pytest_$tag.log | tail -12
grep -E "^golden|^C3 rows|^C5|^smoke" gpurun_out/pytest_$tag.log | head -12
python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench exit $?"; tail -c 500 gpurun_out/bench_$tag.err
MASG_LN_FUSE=0 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${tag}_off.json 2>/dev/null
python -c "
import json
for t in ('$tag','${tag}_off'):
    d=json.load(open('gpurun_out/bench_%s.json'%t)); print(t, d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['kernel_ms_per_step'], d['gpu_launches'], d['clocks'], d['latency_p50_ms']['value'])"
CMD="python bench.py --profile-range"
$CMD > gpurun_out/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 4000 --csv --log-file gpurun_out/launches_$tag.csv $CMD > gpurun_out/ncu1_$tag.log 2>&1
python tools/launch_summary.py gpurun_out/launches_$tag.csv x | sed -n 1,14p
python tools/launch_summary.py gpurun_out/launches_$tag.csv x | grep -E "^(4[0-9]|5[0-9]) "

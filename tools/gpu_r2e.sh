#!/bin/bash
# tests, C2 latency (PDL auto for sub-wave launches), bench + launch list (resident weights across N tiles)
tools/gpu_tests.sh r2e > gpurun_out/tests_r2e.txt 2>&1; grep -E "passed|failed|FAILED" gpurun_out/pytest_r2e.log | tail -5
python tools/latency_probe.py 2>&1 | tail -1
MASG_TC_WRESIDENT=0 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r2e_off.json 2>/dev/null
python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r2e.json 2> gpurun_out/bench_r2e.err; echo "bench exit $?"
python -c "
import json
for t in ('r2e_off','r2e'):
    d=json.load(open('gpurun_out/bench_%s.json'%t)); print(t, d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['kernel_ms_per_step'], d['clocks'], d['latency_p50_ms']['value'])"
CMD="python bench.py --profile-range"
$CMD > gpurun_out/plain_r2e.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 4000 --csv --log-file gpurun_out/launches_r2e.csv $CMD > gpurun_out/ncu1_r2e.log 2>&1
python tools/launch_summary.py gpurun_out/launches_r2e.csv x | sed -n 1,16p
python tools/launch_summary.py gpurun_out/launches_r2e.csv x | grep -E "^(5[4-9]|6[0-9]|7[0-2]) "

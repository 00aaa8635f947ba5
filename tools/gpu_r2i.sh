#!/bin/bash
tools/gpu_tests.sh r2i > gpurun_out/tests_r2i.txt 2>&1; grep -E "passed|failed|FAILED|^E  " gpurun_out/pytest_r2i.log | tail -12
grep -E "^golden.*x3" gpurun_out/pytest_r2i.log
python tools/latency_probe.py 2>&1 | tail -1
python bench.py --steps 8 --warmup 3 --no-cpu-baseline --precision bf16x3 > gpurun_out/bench_r2i_x3.json 2> gpurun_out/bench_r2i_x3.err; echo "x3 bench exit $?"; tail -c 300 gpurun_out/bench_r2i_x3.err
python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r2i.json 2> gpurun_out/bench_r2i.err; echo "bench exit $?"
python -c "
import json
for t in ('r2i_x3','r2i'):
    d=json.load(open('gpurun_out/bench_%s.json'%t)); print(t, d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['kernel_ms_per_step'], d['gpu_launches'], d['clocks'], d['latency_p50_ms']['value'])"

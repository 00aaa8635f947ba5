#!/bin/bash
tools/gpu_tests.sh r2j > gpurun_out/tests_r2j.txt 2>&1; grep -E "passed|failed|FAILED|^E  " gpurun_out/pytest_r2j.log | tail -12
python tools/latency_probe.py 2>&1 | tail -1
MASG_TC_SMALL=0 python tools/latency_probe.py 2>&1 | tail -1
MASG_TC_SMALL_TILES=148 python tools/latency_probe.py 2>&1 | tail -1
MASG_TC_SMALL_TILES=24 python tools/latency_probe.py 2>&1 | tail -1
PROBE_B=2 PROBE_L=1000 python tools/latency_probe.py 2>&1 | tail -1
PROBE_B=2 PROBE_L=1000 MASG_TC_SMALL=0 python tools/latency_probe.py 2>&1 | tail -1
PROBE_B=2 PROBE_L=1000 MASG_TC_SMALL_TILES=148 python tools/latency_probe.py 2>&1 | tail -1
python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r2j.json 2> gpurun_out/bench_r2j.err; echo "bench exit $?"; tail -c 300 gpurun_out/bench_r2j.err
python -c "
import json
d=json.load(open('gpurun_out/bench_r2j.json')); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['frac'], d['gpu_launches'], d['clocks'], d['latency_p50_ms']['value'])"

#!/bin/bash
tools/gpu_tests.sh r2d
tools/gpu_latency.sh
python bench.py --steps 8 --warmup 3 > gpurun_out/bench_r2d.json 2> gpurun_out/bench_r2d.err; echo "bench exit $?"; tail -c 600 gpurun_out/bench_r2d.err
python -c "
import json; d=json.load(open('gpurun_out/bench_r2d.json')); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['kernel_ms_per_step'], d['clocks']); print(d['cpu_baseline']); print(d['gpu_eager_baseline']); print(d['latency_p50_ms'])"

#!/bin/bash
# C4 batch-shape sweep on one GPU.  Usage: gpurun -- tools/gpu_c4sweep.sh
out=gpurun_out; mkdir -p $out
for cfg in "64 40000" "128 40000" "256 40000" "512 40000" "256 80000" "512 80000"; do set -- $cfg
  python bench.py --workload c4 --steps 2 --warmup 3 --max-batch $1 --frame-budget $2 > $out/bench_c4_mb$1_fb$2.json 2> $out/bench_c4_mb$1_fb$2.err || tail -3 $out/bench_c4_mb$1_fb$2.err
  python -c "
import json; d=json.loads(open('$out/bench_c4_mb$1_fb$2.json').read().strip().splitlines()[-1]); print('mb $1 fb $2', round(d['ms_per_step'],1), round(d['value']), round(d['e2e']['value']), round(d['roofline']['frac'],3), d['config']['batches'], d['gpu_launches'])"
done

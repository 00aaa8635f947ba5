#!/bin/bash
# A/B of an environment switch on one box: tests (default setting), then bench with VAR=0 / VAR=1 twice.  Usage: gpurun -- tools/gpu_ab.sh VAR [tag]
var=${1:-MASG_PDL}; tag=${2:-ab}; out=gpurun_out; mkdir -p $out
python -m pytest tests -m gpu -x -q 2>&1 | tail -1
for rep in 1 2; do for v in 0 1; do
  env $var=$v python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $out/bench_${tag}_${v}_$rep.json 2> $out/bench_${tag}_${v}_$rep.err || tail -5 $out/bench_${tag}_${v}_$rep.err
  python -c "
import json; d=json.load(open('$out/bench_${tag}_${v}_$rep.json')); print('$var=$v', round(d['ms_per_step'],3), round(d['value']), round(d['e2e']['value']), round(d['roofline']['frac'],3), d['roofline']['step_ms'], d['clocks']['sm_mhz'])"
done; done

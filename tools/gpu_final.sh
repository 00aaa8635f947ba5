#!/bin/bash
# the driver's round-end sequence on one box: GPU tests, smoke(), the reference arm, the default bench line
out=gpurun_out; mkdir -p $out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
( time python bench.py --impl reference --steps 3 --warmup 1 > $out/bench_final_ref.json 2> $out/bench_final_ref.err ) 2>&1 | grep real
( time python bench.py > $out/bench_final.json 2> $out/bench_final.err ) 2>&1 | grep real
python -c "
import json
r=json.load(open('$out/bench_final_ref.json')); d=json.load(open('$out/bench_final.json'))
print('ref', round(r['value'],2), r['cpu_baseline']['cores'], 'ours', round(d['ms_per_step'],2), round(d['value']), 'e2e', round(d['e2e']['value']), 'frac', round(d['roofline']['frac'],3), 'share', round(d['roofline']['share_of_step'],3), 'lat', round(d['latency_p50_ms']['value'],2), d['clocks'], d['cpu_baseline']['value'])"

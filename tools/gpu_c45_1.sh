#!/bin/bash
out=gpurun_out
python bench.py --workload c5 --steps 8 --warmup 3 > $out/bench_c5_r2_1gpu.json 2> $out/bench_c5_r2_1gpu.err; echo "c5 exit $?"
python bench.py --workload c4 --steps 3 --warmup 3 > $out/bench_c4_r2_1gpu.json 2> $out/bench_c4_r2_1gpu.err; echo "c4 exit $?"
for w in c5 c4; do python - <<PY
import json
d = json.loads(open('$out/bench_${w}_r2_1gpu.json').read().strip().splitlines()[-1])
print('$w', d['n_gpus'], 'ms', round(d['ms_per_step'], 2), 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'frac', round(d['roofline']['frac'], 3), 'launches', d['gpu_launches'], d['clocks'])
PY
done

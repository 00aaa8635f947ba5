#!/bin/bash
# C3 (weak; plain and with the NCCL waveform gather) / C5 / C4 (strong) bench lines at N GPUs of one box (no tests).
# Usage: gpurun --gpus N --timeout 900 -- tools/gpu_scale.sh <tag> N
tag=${1:-sc}; n=${2:-2}; out=gpurun_out; mkdir -p $out
python bench.py --gpus $n --steps 8 --warmup 3 --no-cpu-baseline > $out/bench_c3_$tag.json 2> $out/bench_c3_$tag.err; echo "c3 exit $?"
if [ "$n" -gt 1 ]; then python bench.py --gpus $n --steps 8 --warmup 3 --no-cpu-baseline --gather > $out/bench_c3g_$tag.json 2> $out/bench_c3g_$tag.err; echo "c3 gather exit $?"; fi
python bench.py --gpus $n --workload c5 --steps 8 --warmup 3 > $out/bench_c5_$tag.json 2> $out/bench_c5_$tag.err; echo "c5 exit $?"
python bench.py --gpus $n --workload c4 --steps 3 --warmup 3 > $out/bench_c4_$tag.json 2> $out/bench_c4_$tag.err; echo "c4 exit $?"
for w in c3 c3g c5 c4; do [ -f $out/bench_${w}_$tag.json ] || continue; tail -c 300 $out/bench_${w}_$tag.err; python - <<PY
import json
try:
    d = json.loads(open('$out/bench_${w}_$tag.json').read().strip().splitlines()[-1])
    print('$w', d['n_gpus'], 'ms', round(d['ms_per_step'], 2), 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'frac', round(d['roofline']['frac'], 3), 'launches', d['gpu_launches'], d['clocks'], d.get('gather') and 'gather')
except Exception as e: print('$w', 'no line', e)
PY
done

#!/bin/bash
# One gpurun call: GPU parity tests, a plain bench run, then the ncu passes of the same command
# (launch list; full-set capture of every k_conv_tc / k_respair_tc launch of one step exported as CSV; a 4-launch
# capture with source of the fused pair kernel).  Usage: gpurun --timeout 1500 -- tools/gpu_profile.sh [tag]
tag=${1:-cur}
out=gpurun_out
mkdir -p $out
python -m pytest tests -m gpu -x -q > $out/pytest_$tag.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest_$tag.log
tail -3 $out/pytest_$tag.log
python bench.py --steps 5 --warmup 3 > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench exit $?"
cat $out/bench_$tag.json
CMD="python bench.py --profile-range"
$CMD > $out/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 4000 --csv --log-file $out/launches_$tag.csv $CMD > $out/ncu1_$tag.log 2>&1
echo "ncu launches exit $?"
$CMD > $out/plain_$tag.log 2>&1 &&
ncu --set full --clock-control none --profile-from-start off -k 'regex:k_conv_tc|k_respair' -c 130 -o $out/prof_all_$tag -f $CMD > $out/ncu2_$tag.log 2>&1
echo "ncu full exit $?"
ncu -i $out/prof_all_$tag.ncu-rep --page raw --csv > $out/prof_all_$tag.csv 2>/dev/null
ls -la $out/prof_all_$tag.ncu-rep; rm -f $out/prof_all_$tag.ncu-rep
$CMD > $out/plain_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:k_respair_tc -s 9 -c 4 -o $out/prof_src_$tag -f $CMD > $out/ncu3_$tag.log 2>&1
echo "ncu src exit $?"
ls -la $out

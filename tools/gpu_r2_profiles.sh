#!/bin/bash
# round-2 evidence on one box: launch list, DRAM/tensor metrics of every dense launch, memory-kernel pass, one --set full capture (with source) of the top kernel
tag=${1:-r2p}; out=gpurun_out; mkdir -p $out
CMD="python bench.py --profile-range"
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench exit $?"
$CMD > $out/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 4000 --csv --log-file $out/launches_$tag.csv $CMD > $out/ncu1_$tag.log 2>&1
echo "launches exit $?"; python tools/launch_summary.py $out/launches_$tag.csv | head -20
tools/gpu_ncu_dram.sh $tag
python tools/ncu_summary.py $out/ncu_dram_$tag.csv $out/ncu_dram_summary_$tag.csv
tools/gpu_ncu_mem.sh $tag
# top kernel by time: k_conv_tc<128, planar> — full section set + source for 6 launches of stage 1/2 (skip the first 60 = FastSpeech2)
$CMD > $out/plain_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on --profile-from-start off -k 'regex:k_conv_tc' -s 70 -c 6 -o $out/prof_top_$tag -f $CMD > $out/ncu3_$tag.log 2>&1
echo "ncu full exit $?"
ncu -i $out/prof_top_$tag.ncu-rep --page raw --csv > $out/prof_top_$tag.csv 2>/dev/null
ncu -i $out/prof_top_$tag.ncu-rep --page details --csv > $out/prof_top_details_$tag.csv 2>/dev/null
ls -la $out/prof_top_$tag.*

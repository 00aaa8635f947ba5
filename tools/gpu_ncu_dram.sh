#!/bin/bash
# DRAM bytes, duration and tensor-pipe activity of every k_conv_tc / k_respair_tc launch of one C3 step (a few replay passes per kernel,
# against ~40 for --set full).  Usage: gpurun -- tools/gpu_ncu_dram.sh <tag>
tag=${1:-z}; out=gpurun_out; mkdir -p $out
CMD="python bench.py --profile-range"
$CMD > $out/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,launch__shared_mem_per_block_dynamic,launch__registers_per_thread \
    --clock-control none --profile-from-start off -k 'regex:k_conv_tc|k_respair' -c 130 --page raw --csv --log-file $out/ncu_dram_$tag.csv $CMD > $out/ncu_dram_$tag.log 2>&1
echo "ncu exit $?"; grep -v "^==" $out/ncu_dram_$tag.csv | head -3 | cut -c1-300

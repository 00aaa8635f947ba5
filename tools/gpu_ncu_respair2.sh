#!/bin/bash
# `ncu --set full` (with source) of the CTA-pair kernels in one C3 step: three k_respair2_tc launches (C = 128 k = 3, C = 64 k = 7, k = 11)
# and two pair-mode k_conv_tc launches.  Usage: gpurun -- tools/gpu_ncu_respair2.sh <tag>
tag=${1:-rp2}; out=gpurun_out; mkdir -p $out
CMD="python bench.py --profile-range"
$CMD > $out/plain_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on --profile-from-start off -k 'regex:k_respair2_tc' -c 9 -o $out/prof_rp2_$tag -f $CMD > $out/ncu_rp2_$tag.log 2>&1
echo "ncu exit $?"
ncu -i $out/prof_rp2_$tag.ncu-rep --page raw --csv > $out/prof_rp2_$tag.csv 2>/dev/null
python - <<PY
import csv
rows = list(csv.reader(open('$out/prof_rp2_$tag.csv')))
hdr = rows[0]
want = ['Kernel Name', 'gpu__time_duration.sum', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'smsp__inst_executed.sum', 'launch__cluster_size',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct', 'launch__registers_per_thread', 'launch__shared_mem_per_block_dynamic']
idx = [hdr.index(w) for w in want if w in hdr]
print(','.join(hdr[i] for i in idx))
for r in rows[2:]:
    print(','.join(r[i].replace(',', '') for i in idx))
PY

#!/bin/bash
# ncu pass over the memory-bound kernels of one C3 step (LayerNorm, bucketize+embedding, LengthRegulator, embedding, conv_post,
# GMM sampling, masks ...): duration, DRAM bytes, DRAM throughput %, L2 bytes.  Usage: gpurun -- tools/gpu_ncu_mem.sh <tag>
tag=${1:-m}; out=gpurun_out; mkdir -p $out
CMD="python bench.py --profile-range"
$CMD > $out/plain_mem_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed,lts__t_bytes.sum,launch__registers_per_thread \
    --clock-control none --profile-from-start off \
    -k 'regex:k_layernorm|k_ln_|k_bucket_add|k_length_regulate|k_conv_post|k_gmm_|k_embed|k_add_pos|k_add_rowvec|k_rowdot|k_duration|k_mel_lens|k_mask|k_compact' \
    -c 200 --page raw --csv --log-file $out/ncu_mem_$tag.csv $CMD > $out/ncu_mem_$tag.log 2>&1
echo "ncu mem exit $?"
python tools/ncu_mem_summary.py $out/ncu_mem_$tag.csv $out/ncu_mem_summary_$tag.csv 2>&1 | tail -30

#!/bin/bash
# what-if timing of the fused-LayerNorm epilogue: 0 = as is, 256 = no residual loads, 512 = no global stores, 768 = neither
for dbg in 0 256 512 768; do
  MASG_TC_DEBUG=$dbg ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 4000 --csv --log-file gpurun_out/launches_ln$dbg.csv python bench.py --profile-range > gpurun_out/ncu_ln$dbg.log 2>&1
  echo "== MASG_TC_DEBUG=$dbg"; python tools/launch_summary.py gpurun_out/launches_ln$dbg.csv x | grep -E "k_conv_tc<256" | awk '{n[$2]++; t[$2]+=$NF} END {for (k in n) print k, n[k], t[k]}'
  python tools/launch_summary.py gpurun_out/launches_ln$dbg.csv x | grep -E "^(4[3-5]|1[0-2]) "
done

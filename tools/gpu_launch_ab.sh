#!/bin/bash
# ncu launch list of one C3 step under VAR=0 and VAR=1 (per-launch times of the same launches side by side).  Usage: gpurun -- tools/gpu_launch_ab.sh VAR
var=${1:-MASG_TC_PAIR_KW128}; out=gpurun_out; mkdir -p $out
for v in 0 1; do
  env $var=$v ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 4000 --csv --log-file $out/launches_ab_$v.csv python bench.py --profile-range > $out/ncu_ab_$v.log 2>&1
  echo "$var=$v exit $?"
done
paste <(python tools/launch_summary.py $out/launches_ab_0.csv x | grep -E "^[0-9]+ " | awk '{print $1, $2, $3, $NF}') <(python tools/launch_summary.py $out/launches_ab_1.csv x | grep -E "^[0-9]+ " | awk '{print $NF}') | awk '{d=$5-$4; s0+=$4; s1+=$5; if (d>0.004||d<-0.004) printf "%s %s %s %.3f %.3f %+.3f\n",$1,$2,$3,$4,$5,d} END {printf "total %.3f %.3f\n", s0, s1}'

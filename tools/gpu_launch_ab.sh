#!/bin/bash
# ncu launch list of one C3 step under VAR=V0 and VAR=V1 (per-launch times of the same launches side by side; differences above THR ms
# are listed).  Usage: gpurun -- tools/gpu_launch_ab.sh VAR [V0 V1 [THR]]
var=${1:-MASG_TC_PAIR_KW128}; v0=${2:-0}; v1=${3:-1}; thr=${4:-0.004}; out=gpurun_out; mkdir -p $out
for v in $v0 $v1; do
  env $var=$v ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 4000 --csv --log-file $out/launches_ab_$v.csv python bench.py --profile-range > $out/ncu_ab_$v.log 2>&1
  echo "$var=$v exit $?"
done
paste <(python tools/launch_summary.py $out/launches_ab_$v0.csv x | grep -E "^[0-9]+ " | awk '{print $1, $2, $3, $NF}') <(python tools/launch_summary.py $out/launches_ab_$v1.csv x | grep -E "^[0-9]+ " | awk '{print $NF}') | awk -v thr=$thr '{d=$5-$4; s0+=$4; s1+=$5; if (d>thr||d<-thr) printf "%s %s %s %.3f %.3f %+.3f\n",$1,$2,$3,$4,$5,d} END {printf "total %.3f %.3f\n", s0, s1}'

#!/bin/bash
# what-if switches of k_conv_tc on the C3 layer shapes + the decoder-plain-bf16 experiment
out=gpurun_out; mkdir -p $out
for d in 0 16 2 4 8 6; do echo "== dbg $d"; MASG_TC_DEBUG=$d python tools/conv_probe.py --reps 5 2>&1 | tail -30; done > $out/whatif.log 2>&1
echo "== MASG_DEC_SPLIT=0 parity" > $out/decsplit.log
MASG_DEC_SPLIT=0 python -m pytest tests -m gpu -x -q -k "fs2 or full_size or end_to_end" >> $out/decsplit.log 2>&1
MASG_DEC_SPLIT=0 python bench.py --steps 5 --warmup 3 --no-cpu-baseline >> $out/decsplit.log 2>&1
tail -5 $out/decsplit.log

#!/bin/bash
# what-if switches on the HBM-side convs: stage-2 k = 3 first conv (case 5), upsampler 2 (case 9)
for dbg in 0 4 2 8 16; do echo "== MASG_TC_DEBUG=$dbg"; for c in 5 9; do MASG_TC_DEBUG=$dbg python tools/conv_probe.py --only $c 2>&1 | tail -2; done; done

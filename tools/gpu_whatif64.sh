#!/bin/bash
# residual epilogue of k_conv_tc with the residual in its own buffer (DRAM reads, as in the generator) against the same conv without it
for dbg in 0 16 64; do echo "== MASG_TC_DEBUG=$dbg"; for c in 4 22 21 6 23 19 24 5; do MASG_TC_DEBUG=$dbg python tools/conv_probe.py --only $c 2>&1 | tail -2; done; done

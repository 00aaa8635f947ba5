#!/bin/bash
for v in 0 1; do echo "== MASG_TC_MSUB_K1=$v"; for c in 15 16 17 11 12; do MASG_TC_MSUB_K1=$v python tools/conv_probe.py --only $c 2>&1 | tail -1; done; done

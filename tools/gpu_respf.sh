#!/bin/bash
# residual L2 prefetch (MASG_TC_RESPF) off / on: probes with the residual in its own buffer, then tests and an A/B of the bench
for v in 0 1; do echo "== MASG_TC_RESPF=$v"; for c in 22 23 24; do MASG_TC_RESPF=$v python tools/conv_probe.py --only $c 2>&1 | tail -1; done; done
tools/gpu_ab.sh MASG_TC_RESPF respf

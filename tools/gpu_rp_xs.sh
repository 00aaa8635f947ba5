#!/bin/bash
# fused pair kernel: probe of every case (C = 64 k = 3 with 3 + 2 and 4 + 1 slots), then tests + bench + launch list
python tools/respair_probe.py 2>&1 | head -8
MASG_RP_C64K3=41 python tools/respair_probe.py --only 5 2>&1 | tail -1
MASG_RP_C64K3=41 python tools/respair_probe.py --only 6 2>&1 | tail -1
tools/gpu_quick.sh x

#!/bin/bash
# fused pair kernel at k = 3: what-if switches (2 no residual loads/stores, 4 no stores, 128 no leaky-relu pass) and the timeline of CTA 0
for dbg in 0 2 4 128; do echo "== MASG_TC_DEBUG=$dbg"; for c in 0 5; do MASG_TC_DEBUG=$dbg python tools/respair_probe.py --only $c 2>&1 | tail -1; done; done
echo "== timeline (MASG_TC_DEBUG=32), case 0"; MASG_TC_DEBUG=32 python tools/respair_probe.py --only 0 --reps 1 2>&1 | head -70

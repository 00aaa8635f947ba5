#!/bin/bash
# fused pair kernel, C = 32: 3 against 6 activation slots; timeline of CTA 0 with 6
for v in 3 6; do echo "== MASG_RP_XS=$v"; MASG_RP_XS=$v python tools/respair_probe.py 2>&1 | head -5; done
echo "== timeline (MASG_TC_DEBUG=32), case 0, 6 slots"; MASG_TC_DEBUG=32 python tools/respair_probe.py --only 0 --reps 1 2>&1 | sed -n 1,9p
echo "== timeline, case 3 (k11 d1)"; MASG_TC_DEBUG=32 python tools/respair_probe.py --only 3 --reps 1 2>&1 | sed -n 1,9p
python -m pytest tests -m gpu -x -q -k "respair or resblock or hifigan" 2>&1 | tail -2

#!/bin/bash
echo "== default"; python tools/debug_c3.py 2>&1 | tail -22
echo "== L=50"; DBG_L=50 python tools/debug_c3.py 2>&1 | tail -22 | head -8
echo "== L=100"; DBG_L=100 python tools/debug_c3.py 2>&1 | tail -22 | head -8
echo "== MASG_ATTN_PLANAR=0"; MASG_ATTN_PLANAR=0 python tools/debug_c3.py 2>&1 | tail -22 | head -6
echo "== MASG_DEC_MODE=1 (split)"; MASG_DEC_MODE=1 python tools/debug_c3.py 2>&1 | tail -22 | head -6
echo "== MASG_F16_PLANAR=0"; MASG_F16_PLANAR=0 python tools/debug_c3.py 2>&1 | tail -22 | head -6
echo "== MASG_LN_PLANAR=0"; MASG_LN_PLANAR=0 python tools/debug_c3.py 2>&1 | tail -22 | head -6
echo "== fp32"; DBG_PREC=fp32 python tools/debug_c3.py 2>&1 | tail -22 | head -6

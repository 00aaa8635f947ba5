#!/bin/sh
# rebuild libmasg.so in-tree and print register/spill lines of the tcgen05 kernel
cd "$(dirname "$0")/.." && python -c "
import importlib,sys
sys.path.insert(0,'.')
b=importlib.import_module('mid-attribute-speaker-generation_b200.build'); b.build(verbose=False)" 2>&1 | grep -v "^ptxas info\|^$" | head -30
grep -A2 "k_conv_tc" mid-attribute-speaker-generation_b200/csrc/build/conv_tc.ptxas.log | grep -E "registers|spill" | sort | uniq -c

#!/bin/bash
# all GPU tests with the printed parity numbers.  Usage: gpurun -- tools/gpu_tests.sh <tag> [pytest -k expr]
tag=${1:-t}; out=gpurun_out; mkdir -p $out
if [ -n "$2" ]; then python -m pytest tests -m gpu -q -s -k "$2" > $out/pytest_$tag.log 2>&1; else python -m pytest tests -m gpu -q -s > $out/pytest_$tag.log 2>&1; fi
echo "pytest exit $?"; tail -15 $out/pytest_$tag.log
grep -E "^(C2|C3|C4|C5|un-forced|  pad-leak|engineered|spread|golden|e2e|smoke)" $out/pytest_$tag.log | head -120

#!/bin/bash
# Round 2, GPU call 29: the own-history parity test of the stand-alone ReSTIR DI reuse passes.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
( time timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -s -k "own_history" ) > $O/r29_tests.log 2>&1
grep -n "share of\|passed\|failed\|Error\|assert" $O/r29_tests.log | cut -c1-600 | tail -12

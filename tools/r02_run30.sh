#!/bin/bash
# Round 2, GPU call 30: the C++ host GPUBaker through its C handles.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
( time timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_bake.py -x -q -m gpu -s -k "host_layer_gpubaker or own_history" ) > $O/r30_tests.log 2>&1
grep -n "share of\|passed\|failed\|Error\|assert" $O/r30_tests.log | cut -c1-600 | tail -12

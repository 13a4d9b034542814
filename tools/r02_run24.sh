#!/bin/bash
# Round 2, GPU call 24: the LUT bakers and the texture import on the GPU (new tests), full-size bake timings.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
( time timeout 900 python -m pytest tests/test_bake.py tests/test_scene_import.py -x -q -m gpu -s ) > $O/r24_tests.log 2>&1
tail -25 $O/r24_tests.log | cut -c1-1500

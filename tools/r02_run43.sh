#!/bin/bash
# Round 2, GPU call 43: the whole GPU suite and smoke() at the final HEAD (default record layout).
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
( time timeout 160 python -m pytest tests -x -q -m gpu ) > gpurun_out/r43_tests.log 2>&1; tail -5 gpurun_out/r43_tests.log | cut -c1-200

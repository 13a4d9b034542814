#!/bin/bash
# Round 2, GPU call 34: bake tests and timings with balanced chunk plans (conductor 11 threads per texel, thin glass 3).
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
( time timeout 900 python -m pytest tests/test_bake.py -x -q -m gpu ) > $O/r34_tests.log 2>&1; tail -4 $O/r34_tests.log | cut -c1-300
timeout 600 python tools/bake_tables.py --tables conductor,glossy,glass,thin_glass > $O/r34_bake_times.json 2> $O/r34_bake_times.err; cat $O/r34_bake_times.json; tail -3 $O/r34_bake_times.err

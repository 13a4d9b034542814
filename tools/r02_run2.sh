#!/bin/bash
# Round 2, GPU call 2: the no-allocation entry point's test alone, then the whole GPU suite (it ends with that test), smoke, default bench.
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
( time timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu ) > gpurun_out/r2_noalloc_alone.log 2>&1
tail -4 gpurun_out/r2_noalloc_alone.log
( time timeout 2400 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest.log 2>&1
tail -6 gpurun_out/r2_pytest.log
timeout 600 python __graft_entry__.py smoke > gpurun_out/r2_smoke.log 2>&1; tail -2 gpurun_out/r2_smoke.log
timeout 900 python bench.py --steps 8 > gpurun_out/r2_bench_c3.json 2> gpurun_out/r2_bench_c3.err; tail -c 1500 gpurun_out/r2_bench_c3.json

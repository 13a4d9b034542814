#!/bin/bash
# NOTE: compute-sanitizer is closed on the GPU pool of this project (it prints "closed on this pool and stays closed" and exits 86): kept for pools where it runs.
# compute-sanitizer pass over every kernel family: memcheck (out-of-bounds / misaligned accesses, leaks at exit) on the
# whole tour, racecheck (shared-memory hazards: the traversal kernel's shared stacks, the scan kernels) and initcheck on the Cornell part.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
python tools/sanitizer_workload.py > $O/san_plain.log 2>&1 || { tail -20 $O/san_plain.log; exit 1; }
timeout 2400 compute-sanitizer --tool memcheck --leak-check full --print-limit 30 --log-file $O/san_memcheck.txt python tools/sanitizer_workload.py > $O/san_memcheck_run.log 2>&1
echo "memcheck exit $?"; tail -4 $O/san_memcheck.txt
timeout 1800 compute-sanitizer --tool racecheck --print-limit 30 --log-file $O/san_racecheck.txt python tools/sanitizer_workload.py > $O/san_racecheck_run.log 2>&1
echo "racecheck exit $?"; tail -4 $O/san_racecheck.txt

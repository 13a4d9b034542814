#!/bin/bash
# Sweep of the scheduling knobs of the ray-queue traversal kernel (trace_queue.cuh) on C3 and C2, single path lane so that the
# per-stage times are meaningful; the pixel checksum must not move. Run on the GPU box: bash tools/trace_sweep.sh
cd "$(dirname "$0")/.."
export B200PT_PATH_LANES=1
for w in synthetic cornell; do
  echo "== $w: refill idle lanes (default 4)"
  for r in 1 2 4 8 16; do echo -n "refill=$r  "; B200PT_TRACE_REFILL_IDLE=$r python tools/stage_times.py $w 1920 1080 8 | grep -v '^$'; done
  echo "== $w: triangle postpone threshold (default 8)"
  for t in 1 4 8 16 24; do echo -n "postpone=$t  "; B200PT_TRACE_TRI_POSTPONE=$t python tools/stage_times.py $w 1920 1080 8 | grep -v '^$'; done
  echo "== $w: trace grid, % of 7-8 resident blocks per SM (default 100)"
  for g in 50 75 100 200; do echo -n "grid=$g%  "; B200PT_TRACE_GRID_PCT=$g python tools/stage_times.py $w 1920 1080 8 | grep -v '^$'; done
done

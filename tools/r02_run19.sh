#!/bin/bash
# Round 2, GPU call 19: SURFACE kernels at 7 blocks per SM (72 registers) + queue grids of 24 blocks per SM by default; grid-share sweep.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
( time timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "sample_batches or path_modes or golden_image" ) > $O/r19_tests.log 2>&1
tail -4 $O/r19_tests.log
run() { name=$1; wl=$2; shift 2; env "$@" timeout 600 python bench.py --workload $wl --steps 2 --no-cpu-baseline > $O/r19_$name.json 2> $O/r19_$name.err; python - $O/r19_$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); r=d["roofline"]
    print(sys.argv[1], round(d["value"],1), "Mrays/s", round(d["ms_per_step"],2), "ms/step", {k:round(v,2) for k,v in r["stage_ms"].items()}, "lanes", d["config"]["path_lanes"])
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-800:])
PY
}
run c3_default synthetic B200PT_PATH_LANES=2
run c3_share1 synthetic B200PT_PATH_LANES=2 B200PT_GRID_SHARE=1
run c3_share1_q16 synthetic B200PT_PATH_LANES=2 B200PT_GRID_SHARE=1 B200PT_QUEUE_BLOCKS_PER_SM=16
run c3_lanes3_q36 synthetic B200PT_PATH_LANES=3 B200PT_QUEUE_BLOCKS_PER_SM=36
run c3_lanes3_q24 synthetic B200PT_PATH_LANES=3
run c2_default cornell B200PT_PATH_LANES=2
run c2_q16 cornell B200PT_PATH_LANES=2 B200PT_QUEUE_BLOCKS_PER_SM=16
run c5_default adaptive_full B200PT_PATH_LANES=2

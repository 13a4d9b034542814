#!/bin/bash
# Round 2, GPU call 27: k_wf_surface / k_wf_camera_surface compiled for 4 / 5 / 6 resident blocks per SM (128 / 96 / 80 registers) instead of 167 registers.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
run() { name=$1; wl=$2; shift 2; env "$@" timeout 600 python bench.py --workload $wl --steps 3 --no-cpu-baseline > $O/r27_$name.json 2> $O/r27_$name.err; python - $O/r27_$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); r=d["roofline"]
    print(sys.argv[1], round(d["value"],1), "Mrays/s", round(d["ms_per_step"],2), "ms/step", {k:round(v,2) for k,v in r["stage_ms"].items()})
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-800:])
PY
}
for n in 4 5; do run c3_sb$n synthetic B200PT_TRACE_PIPELINE=2 B200PT_LIB_DIR=$GRAFT_REPO_ROOT/gpurun_alt/sb$n; done
run c3_base synthetic B200PT_TRACE_PIPELINE=2

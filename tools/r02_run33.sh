#!/bin/bash
# Round 2, GPU call 33: triangle fetch policy of the ray-queue traversal (PT_TRACE_TRI_LOAD = 1 streaming loads, 2 = L1 no-allocate; triangle prefetch into L2 only).
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
run() { name=$1; wl=$2; shift 2; env "$@" timeout 600 python bench.py --workload $wl --steps 3 --no-cpu-baseline > $O/r33_$name.json 2> $O/r33_$name.err; python - $O/r33_$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); r=d["roofline"]
    print(sys.argv[1], round(d["value"],1), "Mrays/s", round(d["ms_per_step"],2), "ms/step", {k:round(v,2) for k,v in r["stage_ms"].items()})
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-800:])
PY
}
for n in ${ALTS:-1 2}; do run c3_tl$n synthetic B200PT_LIB_DIR=$GRAFT_REPO_ROOT/gpurun_alt/tl$n; done
run c3_base synthetic

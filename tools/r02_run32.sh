#!/bin/bash
# Round 2, GPU call 32: the phase kernels P1-P5 compiled for 8 / 10 resident blocks of 64 threads per SM (128 / 102 registers: no / fewer spills) instead of 12 (80 registers).
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
run() { name=$1; wl=$2; shift 2; env "$@" timeout 600 python bench.py --workload $wl --steps 3 --no-cpu-baseline > $O/r32_$name.json 2> $O/r32_$name.err; python - $O/r32_$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); r=d["roofline"]
    print(sys.argv[1], round(d["value"],1), "Mrays/s", round(d["ms_per_step"],2), "ms/step", {k:round(v,2) for k,v in r["stage_ms"].items()})
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-800:])
PY
}
for n in ${ALTS:-8 10}; do run c3_pb$n synthetic B200PT_LIB_DIR=$GRAFT_REPO_ROOT/gpurun_alt/pb$n; done
run c3_base synthetic

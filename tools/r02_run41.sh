#!/bin/bash
# Round 2, GPU call 41: C5 (4K adaptive, default steps) by record layout: which of the late layout changes costs it 2.7 %.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
run() { name=$1; shift; env "$@" timeout 600 python bench.py --workload adaptive_full --width 3840 --height 2160 --samples-per-frame 8 --no-cpu-baseline > $O/r41_$name.json 2> $O/r41_$name.err; python - $O/r41_$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); r=d["roofline"]
    print(sys.argv[1], round(d["value"],1), "Mrays/s", round(d["ms_per_step"],2), "ms/step", {k:round(v,2) for k,v in r["stage_ms"].items()})
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-800:])
PY
}
run all B200PT_PATH_RECORDS=1
run path_only B200PT_PATH_RECORDS=2
run arrays B200PT_PATH_RECORDS=0
run arrays_noshading B200PT_PATH_RECORDS=0 B200PT_SHADING_RECORDS=0

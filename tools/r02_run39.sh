#!/bin/bash
# Round 2, GPU call 39: NEE records (64 B) and hit records (32 B) grouped as well (B200PT_PATH_RECORDS=1) against separate arrays (=0): bit-identity tests, C3 / C2 / C4.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
( time timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "path_record or shading_records or sample_batches or path_modes or path_lanes or ris_phases or restir_phases" ) > $O/r39_tests.log 2>&1; tail -4 $O/r39_tests.log | cut -c1-400
run() { name=$1; wl=$2; shift 2; env "$@" timeout 600 python bench.py --workload $wl --steps 3 --no-cpu-baseline > $O/r39_$name.json 2> $O/r39_$name.err; python - $O/r39_$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); r=d["roofline"]
    print(sys.argv[1], round(d["value"],1), "Mrays/s", round(d["ms_per_step"],2), "ms/step", {k:round(v,2) for k,v in r["stage_ms"].items()})
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-800:])
PY
}
run c3_records synthetic B200PT_PATH_RECORDS=1
run c3_arrays synthetic B200PT_PATH_RECORDS=0
run c2_records cornell B200PT_PATH_RECORDS=1
run c4_records many_lights B200PT_PATH_RECORDS=1

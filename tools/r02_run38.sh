#!/bin/bash
# Round 2, GPU call 38: grouped path records (ray 32 B, hit 64 B, radiance 32 B; B200PT_PATH_RECORDS=1) against seven separate arrays (=0): bit-identity tests, then C3 / C2 / C4 / C5.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
( time timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "path_record or shading_records or sample_batches or path_modes or path_lanes or ris_phases" ) > $O/r38_tests.log 2>&1; tail -4 $O/r38_tests.log | cut -c1-400
run() { name=$1; wl=$2; shift 2; extra=""; if [ "$wl" = "adaptive_full" ]; then extra="--width 3840 --height 2160 --samples-per-frame 8"; fi; env "$@" timeout 600 python bench.py --workload $wl $extra --steps 3 --no-cpu-baseline > $O/r38_$name.json 2> $O/r38_$name.err; python - $O/r38_$name.json <<'PY'
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
run c2_arrays cornell B200PT_PATH_RECORDS=0
run c4_records many_lights B200PT_PATH_RECORDS=1
run c4_arrays many_lights B200PT_PATH_RECORDS=0
run c5_records adaptive_full B200PT_PATH_RECORDS=1
run c5_arrays adaptive_full B200PT_PATH_RECORDS=0

#!/bin/bash
# Round 2, GPU call 23: RIS light candidates with their triangles fetched together -- bit-identity tests, C4 bench, launch list.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
( time timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_function_kats.py -x -q -m gpu -k "ris or restir or path_modes or permutation or golden or kats" ) > $O/r23_tests.log 2>&1
tail -5 $O/r23_tests.log | cut -c1-300
run() { name=$1; wl=$2; shift 2; env "$@" timeout 600 python bench.py --workload $wl --steps 3 --no-cpu-baseline > $O/r23_$name.json 2> $O/r23_$name.err; python - $O/r23_$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); r=d["roofline"]
    print(sys.argv[1], round(d["value"],1), "Mrays/s", round(d["ms_per_step"],2), "ms/step", {k:round(v,2) for k,v in r["stage_ms"].items()}, "lanes", d["config"]["path_lanes"])
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-800:])
PY
}
run c4 many_lights B200PT_PATH_LANES=2
run c3 synthetic B200PT_PATH_LANES=2
W=many_lights
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/p23_${W}_launches.csv python bench.py --profile --steps 1 --samples-per-frame 1 --workload $W > $O/p23_${W}_ncu_launches.log 2>&1
python tools/ncu_launches.py $O/p23_${W}_launches.csv --sum > $O/p23_${W}_launches_summary.txt 2>&1
tail -18 $O/p23_${W}_launches_summary.txt

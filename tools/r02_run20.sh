#!/bin/bash
# Round 2, GPU call 20: fused ReSTIR DI pass with SEARCH + per-neighbour SITES kernels (bit-identity, C4 bench on / off); path-lane sweep at 12 queue blocks per lane.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
( time timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "restir or ris_phases or sample_batches or path_modes" ) > $O/r20_tests.log 2>&1
tail -6 $O/r20_tests.log | cut -c1-300
run() { name=$1; wl=$2; shift 2; env "$@" timeout 600 python bench.py --workload $wl --steps 2 --no-cpu-baseline > $O/r20_$name.json 2> $O/r20_$name.err; python - $O/r20_$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); r=d["roofline"]
    print(sys.argv[1], round(d["value"],1), "Mrays/s", round(d["ms_per_step"],2), "ms/step", {k:round(v,2) for k,v in r["stage_ms"].items()}, "lanes", d["config"]["path_lanes"])
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-800:])
PY
}
run c4_sites1 many_lights B200PT_RESTIR_SITE_PARALLEL=1
run c4_sites0 many_lights B200PT_RESTIR_SITE_PARALLEL=0
run c4_sites1_lanes3 many_lights B200PT_RESTIR_SITE_PARALLEL=1 B200PT_PATH_LANES=3
run c3_lanes2 synthetic B200PT_PATH_LANES=2
run c3_lanes3 synthetic B200PT_PATH_LANES=3
run c3_lanes4 synthetic B200PT_PATH_LANES=4
run c2_lanes3 cornell B200PT_PATH_LANES=3
run c2_lanes4 cornell B200PT_PATH_LANES=4
W=many_lights
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/p20_${W}_launches.csv python bench.py --profile --steps 1 --samples-per-frame 1 --workload $W > $O/p20_${W}_ncu_launches.log 2>&1
python tools/ncu_launches.py $O/p20_${W}_launches.csv --sum > $O/p20_${W}_launches_summary.txt 2>&1
grep -i "restir" $O/p20_${W}_launches_summary.txt | tail -8

#!/bin/bash
# Round 2, GPU call 13: phased ReSTIR DI passes -- bit-identity against the single kernels, ReSTIR parity tests, C4 bench phased / unphased, launch list.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
( time timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "restir or ris_phases or permutation" ) > $O/r13_tests.log 2>&1
tail -25 $O/r13_tests.log
for on in 1 0; do
  B200PT_RESTIR_PHASES=$on timeout 900 python bench.py --workload many_lights --steps 3 --samples-per-frame 8 --no-cpu-baseline > $O/r13_c4_phases$on.json 2> $O/r13_c4_phases$on.err
done
for f in $O/r13_c4_phases*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); r=d["roofline"]
    print(sys.argv[1], round(d["value"],1), "Mrays/s", round(d["ms_per_step"],2), "ms/step", {k:round(v,2) for k,v in r["stage_ms"].items()}, "launches", d["gpu_launches"])
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
done
W=many_lights
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/p13_${W}_launches.csv python bench.py --profile --steps 1 --samples-per-frame 1 --workload $W > $O/p13_${W}_ncu_launches.log 2>&1
python tools/ncu_launches.py $O/p13_${W}_launches.csv --sum > $O/p13_${W}_launches_summary.txt 2>&1
tail -32 $O/p13_${W}_launches_summary.txt

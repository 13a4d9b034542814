#!/bin/bash
# Round 2, GPU call 14: every ReSTIR DI pass phased (bit-identity incl. the stand-alone temporal / spatial passes), C4 bench with the register-capped
# collect kernels, then the compute-sanitizer pass (tools/r02_sanitizer.sh).
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
( time timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "restir or ris_phases or permutation" ) > $O/r14_tests.log 2>&1
tail -12 $O/r14_tests.log
timeout 900 python bench.py --workload many_lights --steps 3 --samples-per-frame 8 --no-cpu-baseline > $O/r14_c4.json 2> $O/r14_c4.err
python - $O/r14_c4.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); r=d["roofline"]
    print(sys.argv[1], round(d["value"],1), "Mrays/s", round(d["ms_per_step"],2), "ms/step", {k:round(v,2) for k,v in r["stage_ms"].items()}, "launches", d["gpu_launches"], "frac", round(r["frac"],3))
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
W=many_lights
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/p14_${W}_launches.csv python bench.py --profile --steps 1 --samples-per-frame 1 --workload $W > $O/p14_${W}_ncu_launches.log 2>&1
python tools/ncu_launches.py $O/p14_${W}_launches.csv --sum > $O/p14_${W}_launches_summary.txt 2>&1
tail -22 $O/p14_${W}_launches_summary.txt
bash tools/r02_sanitizer.sh

#!/bin/bash
# Round 2, GPU call 11: phased RIS -- bit-identity against megakernel / split, ReSTIR + RIS parity, C4 and RIS-Cornell bench (phases on / off).
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "ris_phases or path_modes or restir or golden_image or permutation" ) > gpurun_out/r11_tests.log 2>&1
tail -25 gpurun_out/r11_tests.log
for on in 1 0; do
  B200PT_RIS_PHASES=$on timeout 900 python bench.py --workload many_lights --steps 3 --samples-per-frame 8 --no-cpu-baseline > gpurun_out/r11_c4_risphases$on.json 2> gpurun_out/r11_c4_risphases$on.err
done
for f in gpurun_out/r11_c4_risphases*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); r=d["roofline"]
    print(sys.argv[1], round(d["value"],1), "Mrays/s", round(d["ms_per_step"],2), "ms/step", {k:round(v,2) for k,v in r["stage_ms"].items()})
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
done

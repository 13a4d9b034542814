#!/bin/bash
# Round 2, GPU call 8: per-function KATs, C5 as specified (parity + 4K bench).
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests/test_gpu_function_kats.py tests/test_gpu_parity.py -x -q -m gpu -k "kats or c5_as_specified" ) > gpurun_out/r8_tests.log 2>&1
tail -25 gpurun_out/r8_tests.log
timeout 1200 python bench.py --workload adaptive_full --width 3840 --height 2160 --steps 4 > gpurun_out/r8_c5_full_4k.json 2> gpurun_out/r8_c5_full_4k.err
timeout 900 python bench.py --workload many_lights --steps 4 > gpurun_out/r8_c4.json 2> gpurun_out/r8_c4.err
for f in gpurun_out/r8_c5_full_4k.json gpurun_out/r8_c4.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); r=d["roofline"]
    print(sys.argv[1], round(d["value"],1), "Mrays/s e2e", round(d["e2e"]["value"],1), round(d["ms_per_step"],3), "ms/step", {k:round(v,3) for k,v in r["stage_ms"].items()}, "cpu", d.get("cpu_baseline",{}).get("value"), d.get("cpu_baseline",{}).get("cores"))
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
done

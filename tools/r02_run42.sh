#!/bin/bash
# Round 2, GPU call 42: final default layout (path records grouped, NEE / hit records as arrays): layout tests (0 / 1 / 2) and the default C3 line.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
( time timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "path_record or shading_records" ) > $O/r42_tests.log 2>&1; grep "passed\|failed" $O/r42_tests.log | cut -c1-200
timeout 300 python bench.py --no-cpu-baseline > $O/r42_c3.json 2> $O/r42_c3.err; python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/r42_c3.json") if l.startswith("{")][-1]); r=d["roofline"]
print("c3 default", round(d["value"],1), "Mrays/s e2e", round(d["e2e"]["value"],1), round(d["ms_per_step"],2), "ms/step", {k:round(v,2) for k,v in r["stage_ms"].items()})
PY

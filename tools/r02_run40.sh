#!/bin/bash
# Round 2, GPU call 40: validation of HEAD after the LUT bakers / texture import / trace flavour 2 -- whole GPU suite, smoke, the default bench lines of C3 / C2 / C4 / C5 (4K), the reference arm.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
( time timeout 2400 python -m pytest tests -x -q -m gpu ) > $O/r40_tests.log 2>&1
tail -6 $O/r40_tests.log | cut -c1-300
( time python -c "import __graft_entry__ as g; g.smoke()" ) > $O/r40_smoke.log 2>&1; tail -3 $O/r40_smoke.log | cut -c1-300
timeout 900 python bench.py > $O/r40_c3.json 2> $O/r40_c3.err
timeout 900 python bench.py --workload cornell --no-cpu-baseline > $O/r40_c2.json 2> $O/r40_c2.err
timeout 900 python bench.py --workload many_lights --no-cpu-baseline > $O/r40_c4.json 2> $O/r40_c4.err
timeout 900 python bench.py --workload adaptive_full --width 3840 --height 2160 --samples-per-frame 8 --no-cpu-baseline > $O/r40_c5.json 2> $O/r40_c5.err
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $O/r40_reference_arm.json 2> $O/r40_reference_arm.err
for f in $O/r40_c?.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); r=d["roofline"]
    print(sys.argv[1], round(d["value"],1), "Mrays/s", round(d["ms_per_step"],2), "ms/step e2e", round(d["e2e"]["value"],1), "launches", d["gpu_launches"], {k:round(v,2) for k,v in r["stage_ms"].items()}, "frac", round(r["frac"],3), "l2_frac", r.get("l2_frac"), d["clocks"], "cpu", (d.get("cpu_baseline") or {}).get("value"))
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
done
tail -c 600 $O/r40_reference_arm.json

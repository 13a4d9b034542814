#!/bin/bash
# Round 2, GPU call 26: distance cull at pop time in the ray-queue traversal: hit-id tests (Cornell batches, the 10 M-triangle tree, the static entry point),
# then C3 with the cull (default) and without (B200PT_TRACE_PIPELINE=2), C2 likewise.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
( time timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q -m gpu -k "ray_batches or hit_ids or flavours or no_alloc or static or golden_ray or c2_size" ) > $O/r26_tests.log 2>&1
tail -5 $O/r26_tests.log | cut -c1-300
run() { name=$1; wl=$2; shift 2; env "$@" timeout 600 python bench.py --workload $wl --steps 3 --no-cpu-baseline > $O/r26_$name.json 2> $O/r26_$name.err; python - $O/r26_$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); r=d["roofline"]
    print(sys.argv[1], round(d["value"],1), "Mrays/s", round(d["ms_per_step"],2), "ms/step", {k:round(v,2) for k,v in r["stage_ms"].items()}, "frac", round(r["frac"],3), {k:v for k,v in r.items() if "per_ray" in k or "nodes" in k})
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-800:])
PY
}
run c3_cull synthetic B200PT_TRACE_PIPELINE=0
run c3_nocull synthetic B200PT_TRACE_PIPELINE=2
run c2_cull cornell B200PT_TRACE_PIPELINE=0
run c2_nocull cornell B200PT_TRACE_PIPELINE=2

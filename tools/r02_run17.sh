#!/bin/bash
# Round 2, GPU call 17: device BVH builder with PLOC clustering -- hit-id tests on both algorithms (incl. the 10 M-triangle scene), C3 bench on the
# device-built trees (PLOC radius 8 / 16 / 32, radix tree) against the host-built tree.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
( time timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_comm.py -x -q -m gpu -s -k "device_built or hit_ids or comm" ) > $O/r17_tests.log 2>&1
grep -E "device build|passed|failed|Error|error" $O/r17_tests.log | cut -c1-400 | tail -12
run() { name=$1; shift; env "$@" timeout 600 python bench.py --steps 2 --no-cpu-baseline > $O/r17_$name.json 2> $O/r17_$name.err; python - $O/r17_$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); r=d["roofline"]; b=d["config"]["bvh"]
    print(sys.argv[1], round(d["value"],1), "Mrays/s", {k:round(v,2) for k,v in r["stage_ms"].items()}, "nodes/ray", round(r["nodes_per_ray"],2), "tris/ray", round(r["triangles_per_ray"],2), "sah", round(b["sah_cost"],2), "fill", round(b["node_fill"],3), "build s", round(b["build_seconds"],3), "device s", round(b["device_build_seconds"],4), "depth", b["max_depth"])
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
}
run host B200PT_BVH_BUILD_DEVICE=0
run ploc16 B200PT_BVH_BUILD_DEVICE=1
run ploc8 B200PT_BVH_BUILD_DEVICE=1 B200PT_PLOC_RADIUS=8
run ploc32 B200PT_BVH_BUILD_DEVICE=1 B200PT_PLOC_RADIUS=32
run lbvh B200PT_BVH_BUILD_DEVICE=1 B200PT_BVH_DEVICE_ALGORITHM=lbvh

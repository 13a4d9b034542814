#!/bin/bash
# Round 2, GPU call 7: whole GPU suite at HEAD (12 permutations, device builder, communicator), C3 on the device-built tree.
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
( time timeout 2700 python -m pytest tests -m gpu -x -q ) > gpurun_out/r7_pytest.log 2>&1
tail -15 gpurun_out/r7_pytest.log
B200PT_BVH_BUILD_DEVICE=1 timeout 900 python bench.py --workload synthetic --steps 4 --no-cpu-baseline > gpurun_out/r7_c3_devicebvh.json 2> gpurun_out/r7_c3_devicebvh.err
python - gpurun_out/r7_c3_devicebvh.json <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); r=d["roofline"]
    print(round(d["value"],1), "Mrays/s", round(d["ms_per_step"],3), "ms/step", {k:round(v,3) for k,v in r["stage_ms"].items()}, "nodes/ray", round(r["nodes_per_ray"],2), "tris/ray", round(r["triangles_per_ray"],2), d["config"]["bvh"])
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY

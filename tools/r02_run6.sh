#!/bin/bash
# Round 2, GPU call 6: device BVH8 builder + C-ABI communicator tests, C3 bench on the device-built tree.
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests/test_gpu_comm.py tests/test_gpu_parity.py -x -q -m gpu -s -k "communicator or device_built or c3_hit_ids" ) > gpurun_out/r6_tests.log 2>&1
grep -E "build:|passed|failed|Error|error" gpurun_out/r6_tests.log | head -20
B200PT_BVH_BUILD_DEVICE=1 timeout 900 python bench.py --workload synthetic --steps 4 --no-cpu-baseline > gpurun_out/r6_c3_devicebvh.json 2> gpurun_out/r6_c3_devicebvh.err
B200PT_BVH_BUILD_DEVICE=1 timeout 900 python bench.py --workload cornell --steps 8 --no-cpu-baseline > gpurun_out/r6_c2_devicebvh.json 2> gpurun_out/r6_c2_devicebvh.err
for f in gpurun_out/r6_c*_devicebvh.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); r=d["roofline"]
    print(round(d["value"],1), "Mrays/s", round(d["ms_per_step"],3), "ms/step", {k:round(v,3) for k,v in r["stage_ms"].items()}, "nodes/ray", round(r["nodes_per_ray"],2), "tris/ray", round(r["triangles_per_ray"],2), d["config"]["bvh"])
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
done

#!/bin/bash
# Round 2, GPU call 1: GPU test suite at HEAD + the pipelined ray-queue loop against the classic one on C3 and C2.
set -x
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/r1_smi.txt
nproc >> gpurun_out/r1_smi.txt
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/r1_pytest.log 2>&1
tail -5 gpurun_out/r1_pytest.log
for f in 0 6 5 7; do
  B200PT_TRACE_PIPELINE=$f timeout 600 python bench.py --workload synthetic --steps 8 --no-cpu-baseline > gpurun_out/r1_c3_pipe$f.json 2> gpurun_out/r1_c3_pipe$f.err
  B200PT_TRACE_PIPELINE=$f timeout 600 python bench.py --workload cornell --steps 16 --no-cpu-baseline > gpurun_out/r1_c2_pipe$f.json 2> gpurun_out/r1_c2_pipe$f.err
done
for f in gpurun_out/r1_c*_pipe*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); r=d["roofline"]
    print(round(d["value"],1), "Mrays/s", round(d["ms_per_step"],3), "ms", {k:round(v,3) for k,v in r["stage_ms"].items()}, "nodes/ray", round(r["nodes_per_ray"],2), "fill", d["config"]["bvh"].get("node_fill"), "depth", d["config"]["bvh"].get("max_depth"))
except Exception as e:
    print("failed", e)
PY
done

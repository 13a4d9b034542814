#!/bin/bash
# Round 2, GPU call 3: fused sample batches -- bit-identity test, whole suite, C3 / C2 bench by passes in flight.
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "sample_batches or path_modes or golden_image or frame_parity" ) > gpurun_out/r3_batches.log 2>&1
tail -8 gpurun_out/r3_batches.log
( time timeout 2400 python -m pytest tests -m gpu -x -q ) > gpurun_out/r3_pytest.log 2>&1
tail -6 gpurun_out/r3_pytest.log
for n in 0 1 4 8 16; do
  B200PT_SAMPLES_IN_FLIGHT=$n timeout 900 python bench.py --workload synthetic --steps 4 --samples-per-frame 16 --no-cpu-baseline > gpurun_out/r3_c3_inflight$n.json 2> gpurun_out/r3_c3_inflight$n.err
done
for n in 0 8; do
  B200PT_SAMPLES_IN_FLIGHT=$n timeout 900 python bench.py --workload cornell --steps 4 --samples-per-frame 16 --no-cpu-baseline > gpurun_out/r3_c2_inflight$n.json 2> gpurun_out/r3_c2_inflight$n.err
done
for f in gpurun_out/r3_c*_inflight*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); r=d["roofline"]
    print(round(d["value"],1), "Mrays/s e2e", round(d["e2e"]["value"],1), round(d["ms_per_step"],3), "ms/step", {k:round(v,3) for k,v in r["stage_ms"].items()}, "launches", d["gpu_launches"])
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
done

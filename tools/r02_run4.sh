#!/bin/bash
# Round 2, GPU call 4: ray binning before traversal -- bit-identity test, then C3 / C2 by cells per axis.
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "sample_batches" ) > gpurun_out/r4_bins_test.log 2>&1
tail -8 gpurun_out/r4_bins_test.log
for b in 0 4 5; do
  B200PT_RAY_BIN_BITS=$b timeout 900 python bench.py --workload synthetic --steps 4 --samples-per-frame 8 --no-cpu-baseline > gpurun_out/r4_c3_bins$b.json 2> gpurun_out/r4_c3_bins$b.err
done
B200PT_RAY_BIN_BITS_EXT=5 B200PT_RAY_BIN_BITS_NEE=0 timeout 900 python bench.py --workload synthetic --steps 4 --samples-per-frame 8 --no-cpu-baseline > gpurun_out/r4_c3_binsE5N0.json 2> gpurun_out/r4_c3_binsE5N0.err
B200PT_RAY_BIN_BITS_EXT=0 B200PT_RAY_BIN_BITS_NEE=5 timeout 900 python bench.py --workload synthetic --steps 4 --samples-per-frame 8 --no-cpu-baseline > gpurun_out/r4_c3_binsE0N5.json 2> gpurun_out/r4_c3_binsE0N5.err
for b in 0 4 5; do
  B200PT_RAY_BIN_BITS=$b timeout 900 python bench.py --workload cornell --steps 8 --samples-per-frame 8 --no-cpu-baseline > gpurun_out/r4_c2_bins$b.json 2> gpurun_out/r4_c2_bins$b.err
done
for f in gpurun_out/r4_c*_bins*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); r=d["roofline"]
    print(round(d["value"],1), "Mrays/s e2e", round(d["e2e"]["value"],1), round(d["ms_per_step"],3), "ms/step", {k:round(v,3) for k,v in r["stage_ms"].items()})
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
done

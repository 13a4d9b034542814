#!/bin/bash
# Round 2, multi-GPU: bench.py under torch.distributed.run on N GPUs of one box (tile sharding of C3, sample-pass sharding of C2, row tiles +
# ReSTIR halo of C4), the data-path collectives through the C ABI (include/b200pt_comm.h).
N=${1:-2}
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
run() {  # name, args...
  name=$1; shift
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/scale${N}_$name.json 2> gpurun_out/scale${N}_$name.err
  grep -v "^NaN at" gpurun_out/scale${N}_$name.json | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('$name', 'N=$N', round(d['value'],1), 'Mrays/s e2e', round(d['e2e']['value'],1), round(d['ms_per_step'],2), 'ms/step', d['scaling'], d['config']['parallelism'][:60], 'collective ms', round(d['config']['allreduce_ms'],3))
except Exception as e:
    print('$name failed', e); print(open('gpurun_out/scale${N}_$name.err').read()[-1200:])
"
}
run c3_columns --workload synthetic --steps 4 --no-cpu-baseline
run c3_samples --workload synthetic --steps 4 --no-cpu-baseline --sharding samples
run c2_samples --workload cornell --steps 8 --no-cpu-baseline
run c4_tiles --workload many_lights --steps 4 --no-cpu-baseline

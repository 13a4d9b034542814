#!/bin/bash
# Round 2, multi-GPU at the final defaults: C3 column tiles (balanced, default), C3 sample passes, C2 sample passes (default for cornell).
N=${1:-8}
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
run() {  # name, args...
  name=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/final${N}_$name.json 2> gpurun_out/final${N}_$name.err
  grep "^{" gpurun_out/final${N}_$name.json | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('$name', 'N=$N', round(d['value'],1), 'Mrays/s e2e', round(d['e2e']['value'],1), round(d['ms_per_step'],2), 'ms/step', d['scaling'], d['config'].get('column_balance'))
except Exception as e:
    print('$name failed', e); print(open('gpurun_out/final${N}_$name.err').read()[-1500:])
"
}
run c3_default --no-cpu-baseline
run c3_samples --no-cpu-baseline --sharding samples
run c2_default --workload cornell --no-cpu-baseline

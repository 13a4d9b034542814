#!/bin/bash
# Round 2, GPU call 31 (2 GPUs): the driver's N = 2 launch of the default bench at HEAD (column tiles of C3 over the C-ABI communicator) and the comm test.
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --no-cpu-baseline > gpurun_out/r31_c3_n2.json 2> gpurun_out/r31_c3_n2.err
grep "^{" gpurun_out/r31_c3_n2.json | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('c3 N=2', round(d['value'],1), 'Mrays/s e2e', round(d['e2e']['value'],1), round(d['ms_per_step'],2), 'ms/step', d['scaling'], d['config'].get('column_balance'), d['config'].get('sharding'))
except Exception as e:
    print('failed', e); print(open('gpurun_out/r31_c3_n2.err').read()[-1500:])
"
( timeout 600 python -m pytest tests/test_gpu_comm.py -x -q -m gpu ) 2>&1 | tail -3

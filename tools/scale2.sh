# two-rank smoke of the three sharding shapes (run with: gpurun --gpus 2 -- bash tools/scale2.sh)
N=${N:-2}
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 8 --warmup 3 "${@:2}" 2>&1 | tail -3; }
run 29511 --workload cornell
run 29512 --workload many_lights
run 29513 --workload synthetic

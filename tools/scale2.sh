# two-rank run of the sharding shapes (run with: gpurun --gpus 2 -- bash tools/scale2.sh): sample passes (C2 and C5 at 4K, weak),
# row tiles + ReSTIR halo (C4, strong), column tiles and row tiles (C3, strong)
N=${N:-2}
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 8 --warmup 3 "${@:2}" 2>&1 | grep '^{' | tail -1; }
run 29511 --workload cornell
run 29512 --workload many_lights
run 29513 --workload synthetic
run 29514 --workload synthetic --sharding tiles
run 29515 --workload adaptive --width 3840 --height 2160

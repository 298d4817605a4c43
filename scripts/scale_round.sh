#!/bin/bash
# Strong-scaling points of the two sharded workloads on N GPUs of one box: scripts/scale_round.sh N
N=$1
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $RUN --master-port 29531 bench.py --gpus $N --workload ring --steps 1 --warmup 1 2> gpurun_out/ring_${N}.err | grep '^{' > gpurun_out/ring512_${N}gpu.json
echo "ring exit $?"; cut -c1-260 gpurun_out/ring512_${N}gpu.json
timeout 600 $RUN --master-port 29532 bench.py --gpus $N --workload scene --steps 1 --warmup 1 2> gpurun_out/scene_${N}.err | grep '^{' > gpurun_out/scene512_${N}gpu.json
echo "scene exit $?"; cut -c1-260 gpurun_out/scene512_${N}gpu.json
tail -n 2 gpurun_out/ring_${N}.err | cut -c1-200; tail -n 2 gpurun_out/scene_${N}.err | cut -c1-200

#!/bin/bash
# 8-GPU pass: the default bench line at N = 8 (weak agg32 + strong-scaling sharded ring + DP train24), as the driver's SCALE run does
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
N=${NGPU:-8}
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_${N}gpu.log 2> gpurun_out/bench_${N}gpu.err; echo "bench$N exit $?"
tail -3 gpurun_out/bench_${N}gpu.err | cut -c1-300; cut -c1-300 gpurun_out/bench_${N}gpu.log

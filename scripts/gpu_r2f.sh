#!/bin/bash
# 2-GPU pass: the NCCL pytest cases that a 1-GPU box skips + the default bench line at N = 2 (sharded ring, DP train24)
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus2.txt
timeout 1500 python -m pytest tests/test_ring_gpu.py tests/test_sharded_scene_gpu.py "tests/test_train_gpu.py::test_training_step_two_processes_nccl" -m gpu -q --timeout 900 > gpurun_out/pytest_gpu_2gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu_2gpu.log
tail -5 gpurun_out/pytest_gpu_2gpu.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_2gpu.log 2> gpurun_out/bench_2gpu.err; echo "bench2 exit $?"
tail -5 gpurun_out/bench_2gpu.err; cut -c1-400 gpurun_out/bench_2gpu.log

#!/bin/bash
# Training-step pass on the GPU box: parity of the training kernels / step, then the configs[4] bench (1 GPU, or torchrun when NGPU > 1).
# EXTRA="--checkpoint" times the reference's checkpointed schedule instead of the keep-activations default.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
N=${NGPU:-1}
TAG=${N}gpu${EXTRA:+_ckpt}
if [ -z "$SKIP_TESTS" ]; then
timeout 900 python -m pytest tests/test_train_gpu.py -m gpu -q --timeout 600 > gpurun_out/pytest_train.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_train.log
tail -40 gpurun_out/pytest_train.log
fi
if [ "$N" = "1" ]; then
timeout 900 python bench.py --workload train --steps ${BENCH_STEPS:-3} --warmup 3 $EXTRA > gpurun_out/bench_train_$TAG.json 2> gpurun_out/bench_train_$TAG.err
else
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --workload train --gpus $N --steps ${BENCH_STEPS:-3} --warmup 3 $EXTRA > gpurun_out/bench_train_$TAG.json 2> gpurun_out/bench_train_$TAG.err
fi
echo "bench exit $?"; tail -5 gpurun_out/bench_train_$TAG.err; cat gpurun_out/bench_train_$TAG.json

#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_train_gpu.py tests/test_train_reference_gpu.py -m gpu -q --timeout 1000 > gpurun_out/pytest_t.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_t.log; tail -6 gpurun_out/pytest_t.log
timeout 900 python bench.py --workload train --steps 3 --warmup 2 > gpurun_out/bench_train.log 2> gpurun_out/bench_train.err; echo "train exit $?"; cut -c1-300 gpurun_out/bench_train.log; tail -3 gpurun_out/bench_train.err

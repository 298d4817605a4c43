#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_loss_gpu.py -m gpu -q --timeout 800 > gpurun_out/pytest_loss.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_loss.log; tail -30 gpurun_out/pytest_loss.log
timeout 600 python bench.py --workload loss > gpurun_out/bench_loss.log 2> gpurun_out/bench_loss.err; echo "loss exit $?"; cat gpurun_out/bench_loss.log; tail -5 gpurun_out/bench_loss.err

#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/test_overlay_gpu.py tests/test_train_gpu.py -m gpu -q -s --timeout 1500 > gpurun_out/pytest_ov.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_ov.log
grep -E "parity|passed|failed|Error|exit|assert|FAILED" gpurun_out/pytest_ov.log | cut -c1-900 | tail -20
timeout 900 python bench.py --workload train --steps 3 --warmup 2 > gpurun_out/bench_train.log 2> gpurun_out/bench_train.err; echo "train exit $?"
tail -3 gpurun_out/bench_train.err; cut -c1-300 gpurun_out/bench_train.log

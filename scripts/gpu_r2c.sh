#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 2400 python -m pytest tests/test_dpt_gpu.py tests/test_geometry_gpu.py tests/test_kernels_gpu.py tests/test_modules_gpu.py tests/test_overlay_gpu.py tests/test_train_reference_gpu.py tests/test_train_gpu.py -m gpu -q -s --timeout 1400 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "parity|passed|failed|Error|exit|assert|FAILED" gpurun_out/pytest_gpu.log | cut -c1-300 | tail -40
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?" >> gpurun_out/bench.err
tail -5 gpurun_out/bench.err; cut -c1-1500 gpurun_out/bench.log

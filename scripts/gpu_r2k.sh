#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_dpt_gpu.py tests/test_overlay_gpu.py -m gpu -q -s --timeout 1400 > gpurun_out/pytest_dpt.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_dpt.log
grep -E "parity|passed|failed|Error|exit|assert|FAILED" gpurun_out/pytest_dpt.log | cut -c1-300 | tail -20
timeout 900 python bench.py --workload dense --steps 3 --warmup 2 > gpurun_out/bench_dense.log 2> gpurun_out/bench_dense.err; echo "dense exit $?"
tail -5 gpurun_out/bench_dense.err; cut -c1-300 gpurun_out/bench_dense.log

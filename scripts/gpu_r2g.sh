#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 2400 python -m pytest tests/test_kernels_gpu.py tests/test_train_gpu.py tests/test_train_reference_gpu.py tests/test_modules_gpu.py tests/test_reference_gpu.py -m gpu -q -s --timeout 1400 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "passed|failed|Error|exit|assert|FAILED" gpurun_out/pytest_gpu.log | cut -c1-300 | tail -30
timeout 900 python bench.py --workload reloc --anchors 100 --queries 40 --steps 2 --warmup 3 > gpurun_out/bench_reloc.log 2> gpurun_out/bench_reloc.err; echo "reloc exit $?"
tail -3 gpurun_out/bench_reloc.err; cut -c1-400 gpurun_out/bench_reloc.log
timeout 900 python bench.py --workload train --steps 3 --warmup 2 > gpurun_out/bench_train.log 2> gpurun_out/bench_train.err; echo "train exit $?"
tail -3 gpurun_out/bench_train.err; cut -c1-600 gpurun_out/bench_train.log

#!/bin/bash
# round 2, first pass: live-reference parity at real size + default bench line
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt
timeout 1200 python -m pytest tests/test_reference_gpu.py -m gpu -q -x -s --timeout 900 > gpurun_out/pytest_ref.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_ref.log
grep -E "parity|passed|failed|Error|exit" gpurun_out/pytest_ref.log | tail -20
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?" >> gpurun_out/bench.err
tail -5 gpurun_out/bench.err; cat gpurun_out/bench.log

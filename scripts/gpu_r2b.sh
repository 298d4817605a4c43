#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -s --timeout 1400 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "parity|passed|failed|Error|exit|assert|FAILED" gpurun_out/pytest_gpu.log | tail -60
timeout 600 python bench.py --workload reloc --anchors 100 --queries 20 --steps 2 --warmup 3 > gpurun_out/bench_reloc.log 2> gpurun_out/bench_reloc.err; echo "reloc exit $?"
tail -3 gpurun_out/bench_reloc.err; cat gpurun_out/bench_reloc.log

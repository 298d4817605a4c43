#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 1700 python -m pytest tests/test_reference_gpu.py tests/test_overlay_gpu.py "tests/test_modules_gpu.py::test_forward_with_cache_matches_reference_golden" "tests/test_kernels_gpu.py" -m gpu -q -s --timeout 1400 > gpurun_out/pytest_ref.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_ref.log
grep -E "parity|passed|failed|Error|exit|assert" gpurun_out/pytest_ref.log | tail -40

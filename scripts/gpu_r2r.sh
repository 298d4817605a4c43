#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 1700 python -m pytest tests/test_overlay_gpu.py -m gpu -q --timeout 1600 -x -s -k "our_camera_head" > gpurun_out/pytest_ov.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_ov.log; tail -30 gpurun_out/pytest_ov.log

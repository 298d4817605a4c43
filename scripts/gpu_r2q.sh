#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_camera_head_train_gpu.py tests/test_pose_bwd_gpu.py tests/test_geometry_gpu.py -m gpu -q --timeout 800 -x -s > gpurun_out/pytest_cht.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_cht.log; tail -40 gpurun_out/pytest_cht.log

#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
python scripts/div_probe.py > gpurun_out/div_probe.log 2>&1; cat gpurun_out/div_probe.log

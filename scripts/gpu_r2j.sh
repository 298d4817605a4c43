#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 600 python scripts/attn_bwd_bench.py --poly > gpurun_out/attn_bwd_poly.log 2>&1; echo "bwd poly exit $?"; cat gpurun_out/attn_bwd_poly.log

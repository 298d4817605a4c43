#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 1500 python bench.py --workload trainstep --steps 3 --warmup 3 > gpurun_out/bench_trainstep.log 2> gpurun_out/bench_trainstep.err; echo "trainstep exit $?"; cut -c1-3000 gpurun_out/bench_trainstep.log; tail -15 gpurun_out/bench_trainstep.err

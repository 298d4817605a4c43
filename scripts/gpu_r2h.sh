#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 600 python scripts/attn_sweep.py --interleaved > gpurun_out/attn_sweep_interleaved.log 2>&1; echo "sweep exit $?"; cat gpurun_out/attn_sweep_interleaved.log
timeout 900 python bench.py --workload dense --steps 3 --warmup 2 > gpurun_out/bench_dense.log 2> gpurun_out/bench_dense.err; echo "dense exit $?"
tail -5 gpurun_out/bench_dense.err; cut -c1-300 gpurun_out/bench_dense.log

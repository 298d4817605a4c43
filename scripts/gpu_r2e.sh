#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 600 python scripts/attn_sweep.py --check > gpurun_out/attn_sweep.log 2>&1; echo "sweep exit $?"
cat gpurun_out/attn_sweep.log
timeout 900 python bench.py --workload reloc --anchors 100 --queries 40 --steps 2 --warmup 3 > gpurun_out/bench_reloc.log 2> gpurun_out/bench_reloc.err; echo "reloc exit $?"
tail -3 gpurun_out/bench_reloc.err; cut -c1-600 gpurun_out/bench_reloc.log

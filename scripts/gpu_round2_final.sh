#!/bin/bash
# What the driver runs at round end, on one B200: pytest -m gpu, smoke(), bench default, bench --impl reference
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q --timeout 1400 > gpurun_out/pytest_gpu_all.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu_all.log
tail -6 gpurun_out/pytest_gpu_all.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"
tail -3 gpurun_out/bench.err; cut -c1-250 gpurun_out/bench.log
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err; echo "ref exit $?"
tail -3 gpurun_out/bench_ref.err; cut -c1-400 gpurun_out/bench_ref.log

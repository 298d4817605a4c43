#!/bin/bash
# One GPU-box pass: parity tests, smoke, short bench, then (only if the plain bench exited 0) the ncu launch list and one
# full capture of the attention + GEMM kernels (B200_PROFILING.md recipe).  Logs under gpurun_out/.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
timeout 900 python bench.py --steps ${BENCH_STEPS:-3} --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?" >> gpurun_out/bench.err
tail -5 gpurun_out/bench.err; cat gpurun_out/bench.log
if [ -n "$SKIP_NCU" ]; then exit 0; fi
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/ncu_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_tc -s 24 -c 3 -o gpurun_out/prof_attn_tc -f $CMD > gpurun_out/ncu_full_attn.log 2>&1
echo "full attn exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 120 -c 8 -o gpurun_out/prof_gemm_tc -f $CMD > gpurun_out/ncu_full_gemm.log 2>&1
echo "full gemm exit $?"
tail -3 gpurun_out/ncu_plain.log; tail -3 gpurun_out/ncu_list.log; tail -3 gpurun_out/ncu_full_attn.log

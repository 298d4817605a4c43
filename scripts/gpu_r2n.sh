#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_train_gpu.py tests/test_kernels_gpu.py -m gpu -q --timeout 800 > gpurun_out/pytest_k.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_k.log; tail -4 gpurun_out/pytest_k.log
timeout 900 python bench.py --workload train --steps 3 --warmup 2 > gpurun_out/bench_train.log 2> gpurun_out/bench_train.err; echo "train exit $?"; cut -c1-250 gpurun_out/bench_train.log
D="python bench.py --workload dense --steps 1 --warmup 1"
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:gemm_tc_kernel<\(int\)[0-9]+, \(int\)[0-9]+, \(bool\)[01], \(int\)2, \(int\)0>' -s 20 -c 3 -o gpurun_out/prof_conv3x3 -f $D > gpurun_out/ncu_b.log 2>&1; echo "conv exit $?"
T="python bench.py --workload train --steps 1 --warmup 1 --only-timed"
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:gemm_tc_kernel<\(int\)[0-9]+, \(int\)[0-9]+, \(bool\)[01], \(int\)1, \(int\)1>' -s 30 -c 3 -o gpurun_out/prof_wgrad -f $T > gpurun_out/ncu_e.log 2>&1; echo "wgrad exit $?"
ls -la gpurun_out/prof_conv3x3.ncu-rep gpurun_out/prof_wgrad.ncu-rep

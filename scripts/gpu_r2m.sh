#!/bin/bash
# round-2 ncu full captures of the kernels added / changed this round (each command ran plain in earlier calls of this round)
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
A="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extras"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_tc4 -s 120 -c 3 -o gpurun_out/prof_attn_layer0 -f $A > gpurun_out/ncu_a.log 2>&1; echo "attn exit $?"
D="python bench.py --workload dense --steps 1 --warmup 1"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:gemm_tc_kernel<[0-9]+, [0-9]+, [0-9]+, 2, 0>" -s 20 -c 3 -o gpurun_out/prof_conv3x3 -f $D > gpurun_out/ncu_b.log 2>&1; echo "conv exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:split_pad -s 10 -c 2 -o gpurun_out/prof_split_pad -f $D > gpurun_out/ncu_c.log 2>&1; echo "split exit $?"
R="python bench.py --workload reloc --anchors 20 --queries 4 --steps 1 --warmup 1"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_f32_tiny -s 40 -c 4 -o gpurun_out/prof_gemv -f $R > gpurun_out/ncu_d.log 2>&1; echo "gemv exit $?"
T="python bench.py --workload train --steps 1 --warmup 1 --only-timed"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:gemm_tc_kernel<[0-9]+, [0-9]+, [0-9]+, 1, 1>" -s 30 -c 3 -o gpurun_out/prof_wgrad -f $T > gpurun_out/ncu_e.log 2>&1; echo "wgrad exit $?"
ls -la gpurun_out/*.ncu-rep; tail -2 gpurun_out/ncu_a.log gpurun_out/ncu_b.log gpurun_out/ncu_d.log gpurun_out/ncu_e.log | cut -c1-200

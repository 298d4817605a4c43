#!/bin/bash
# round 2 profiles: ncu launch list of the agg32 step + full captures of the attention kernel (global / frame / reloc launches)
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extras"
timeout 600 $CMD > gpurun_out/ncu_plain.log 2>&1 || { echo "plain bench failed"; tail -5 gpurun_out/ncu_plain.log; exit 1; }
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "list exit $?"
# the first aggregator layer's attention launches come after the 24 DINOv2 blocks of the warm-up step: skip into the timed step
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:attn_tc4 -s 100 -c 4 -o gpurun_out/prof_attn_tc -f $CMD > gpurun_out/ncu_full_attn.log 2>&1
echo "full attn exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 500 -c 6 -o gpurun_out/prof_gemm_tc -f $CMD > gpurun_out/ncu_full_gemm.log 2>&1
echo "full gemm exit $?"
tail -2 gpurun_out/ncu_list.log; tail -2 gpurun_out/ncu_full_attn.log
ls -la gpurun_out/*.ncu-rep

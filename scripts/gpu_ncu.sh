#!/bin/bash
# Launch list of one bench step + one full capture of the attention kernel (B200_PROFILING.md recipe).
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s ${NCU_SKIP:-930} -c ${NCU_COUNT:-930} --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "list exit $?"
ncu --set full --clock-control none --import-source on -k regex:${NCU_KERNEL:-attn_tc} -s ${NCU_KSKIP:-100} -c 3 -o gpurun_out/prof_${NCU_KERNEL:-attn_tc} -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "full exit $?"
tail -3 gpurun_out/ncu_plain.log; tail -3 gpurun_out/ncu_list.log; tail -3 gpurun_out/ncu_full.log

#!/bin/bash
# ncu evidence for the training step (configs[4]): launch list of one step and a full capture of the attention-backward kernel.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
CMD="python bench.py --workload train --steps 1 --warmup 1 --only-timed"
timeout 600 $CMD > gpurun_out/ncu_train_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_train_plain.log; exit 1; }
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 30000 --csv --log-file gpurun_out/launches_train.csv $CMD > gpurun_out/ncu_train_list.log 2>&1
echo "list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_bwd_kernel -s 12 -c 6 -o gpurun_out/prof_attn_bwd -f $CMD > gpurun_out/ncu_full_attn_bwd.log 2>&1
echo "full attn_bwd exit $?"
tail -2 gpurun_out/ncu_train_plain.log; tail -2 gpurun_out/ncu_train_list.log; tail -2 gpurun_out/ncu_full_attn_bwd.log; ls -la gpurun_out | head -30

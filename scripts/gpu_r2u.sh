#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
L="python bench.py --workload loss --steps 20 --warmup 3"
timeout 600 $L > gpurun_out/bench_loss2.log 2>&1; echo "loss exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:loss_residual_hist_kernel|loss_lookup_bwd_kernel|cdf_tables_kernel' -s 30 -c 3 -o gpurun_out/prof_loss -f $L > gpurun_out/ncu_loss.log 2>&1; echo "ncu loss exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:loss_|cdf_|pair_setup' -s 80 -c 16 --csv --log-file gpurun_out/launches_loss.csv $L > gpurun_out/ncu_loss_list.log 2>&1; echo "ncu list exit $?"
ls -la gpurun_out/prof_loss.ncu-rep gpurun_out/launches_loss.csv

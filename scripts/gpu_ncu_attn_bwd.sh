#!/bin/bash
# Full ncu capture of the attention-backward kernel on the isolated shapes of scripts/attn_bwd_bench.py (cheap: ~40 launches).
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 200 python scripts/attn_bwd_bench.py > gpurun_out/attn_bwd_plain.log 2>&1 || { tail -3 gpurun_out/attn_bwd_plain.log; exit 1; }
cat gpurun_out/attn_bwd_plain.log
# launches per shape: 13 x (delta + mode 0 + mode 1 [+ prefix mode 1]); skip the warm-up of frame24, take both passes of it and of global12
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_bwd_kernel -s 6 -c 2 -o gpurun_out/prof_attn_bwd_frame -f python scripts/attn_bwd_bench.py > gpurun_out/ncu_attn_bwd_frame.log 2>&1
echo "frame exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_bwd_kernel -s 32 -c 2 -o gpurun_out/prof_attn_bwd_global -f python scripts/attn_bwd_bench.py > gpurun_out/ncu_attn_bwd_global.log 2>&1
echo "global exit $?"

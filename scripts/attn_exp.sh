#!/bin/bash
# timing experiments on the attention kernel (outputs are garbage in the DBG / POLY=8 modes)
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
{
echo "== baseline"; python scripts/attn_sweep.py --check
for v in ${ATTN_VARIANTS:-"SR_ATTN_POLY=8" "SR_ATTN_DBG=3" "SR_ATTN_POLY=2" "SR_ATTN_POLY=3" "SR_ATTN_POLY=4"}; do echo "== $v"; env $v python scripts/attn_sweep.py; done
python scripts/attn_trace.py | tail -12
} > gpurun_out/attn_exp.log 2>&1
cat gpurun_out/attn_exp.log

#!/bin/bash
# Timing experiments on the attention-backward kernel (results are NOT valid attention gradients when the knob is set):
# SR_ATTN_BWD_DBG=1 skips the exp / dS arithmetic, =3 also skips the TMEM loads of S / dP.
# BASE_LIB (optional) = a second build of the library to time next to the in-tree one.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
{
if [ -n "$BASE_LIB" ]; then
for d in 0 1 3; do
  echo "== base lib, SR_ATTN_BWD_DBG=$d"
  SAILRECON_B200_LIB=$BASE_LIB SR_ATTN_BWD_DBG=$d timeout 300 python scripts/attn_bwd_bench.py
done
fi
for d in 0 1 3; do
  echo "== in-tree lib, SR_ATTN_BWD_DBG=$d"
  SR_ATTN_BWD_DBG=$d timeout 300 python scripts/attn_bwd_bench.py
done
} 2>&1 | tee gpurun_out/attn_bwd_exp.txt
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_train_gpu.py -m gpu -q -k "bwd or train or block or layer" --timeout 600 2>&1 | tail -5

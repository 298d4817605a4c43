cd $GRAFT_REPO_ROOT
nvidia-smi --query-gpu=name,memory.total --format=csv
for k in layernorm adaln gather_copy gemm_bf16_bias gelu_and_ls qkv_epilogue gemm_f32 attention_bf16_segments attention_bf16_prefix carry attention_f32 errors; do
  echo "=== $k"; timeout 300 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k "$k" -x 2>&1 | tail -15
done > gpurun_out/kernels1.log 2>&1
tail -100 gpurun_out/kernels1.log

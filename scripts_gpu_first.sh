#!/bin/bash
# first GPU contact: every kernel test in its own process (a CUDA fault must not poison the rest)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv > gpurun_out/gpu.txt
for t in test_pc_updates test_pc_philox test_stft_compress test_expand_istft test_stft_roundtrip test_groupnorm test_fir_resample test_conv_direct test_conv_tc test_attention test_temb_and_output test_cpu_tensors; do
  echo "=== $t" >> gpurun_out/kernels.log
  timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k $t -x --timeout 300 2>&1 | tail -40 >> gpurun_out/kernels.log
done
echo "=== parity" >> gpurun_out/kernels.log
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -q -s --timeout 600 2>&1 | tail -60 > gpurun_out/parity.log

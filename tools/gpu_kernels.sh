#!/bin/bash
# first-light GPU run: kernel-level parity in separate processes so that one fault does not hide the rest
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
rm -f gpurun_out/kernel_parity.jsonl
for sel in "not conv and not wgrad" "conv_fwd" "conv_dgrad" "wgrad"; do
  tag=$(echo "$sel" | tr ' ' '_')
  timeout -k 10 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "$sel" --timeout 240 --timeout-method=thread --tb=short \
     > gpurun_out/k_${tag}.log 2>&1
  echo "== $sel -> exit $?"; tail -n 25 gpurun_out/k_${tag}.log
done

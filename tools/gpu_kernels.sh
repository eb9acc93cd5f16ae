#!/bin/bash
# kernel-level parity in separate processes so that one fault does not hide the rest
mkdir -p gpurun_out
rm -f gpurun_out/kernel_parity.jsonl
for sel in "tma or reflect_pad" "conv_fwd and not tma" "conv_dgrad and not tma" "wgrad and not tma" "not conv and not wgrad and not reflect_pad"; do
  tag=$(echo "$sel" | tr ' ' '_')
  timeout -k 10 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "$sel" --timeout 240 --timeout-method=thread --tb=short \
     > gpurun_out/k_${tag}.log 2>&1
  echo "== $sel -> exit $?"; tail -n 25 gpurun_out/k_${tag}.log | cut -c1-300
done

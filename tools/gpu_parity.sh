#!/bin/bash
# all GPU parity tests + smoke (separate processes per group so one fault does not hide the rest)
mkdir -p gpurun_out
rm -f gpurun_out/kernel_parity.jsonl gpurun_out/golden_parity.jsonl
timeout -k 10 900 python -m pytest tests -m gpu -q --timeout 600 --timeout-method=thread --tb=short > gpurun_out/gpu_tests.log 2>&1
echo "== gpu tests exit $?"; tail -n 30 gpurun_out/gpu_tests.log
timeout -k 10 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
echo "== smoke exit $?"; tail -n 8 gpurun_out/smoke.log

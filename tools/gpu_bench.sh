#!/bin/bash
# full GPU test suite + smoke + bench with per-geometry breakdown
mkdir -p gpurun_out
rm -f gpurun_out/kernel_parity.jsonl gpurun_out/golden_parity.jsonl
timeout -k 10 900 python -m pytest tests -m gpu -q --timeout 600 --timeout-method=thread --tb=short > gpurun_out/gpu_tests.log 2>&1
echo "== gpu tests exit $?"; tail -n 15 gpurun_out/gpu_tests.log | cut -c1-300
timeout -k 10 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
echo "== smoke exit $?"; tail -n 3 gpurun_out/smoke.log
timeout -k 10 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --breakdown gpurun_out/conv_breakdown.json > gpurun_out/bench.log 2>&1
echo "== bench exit $?"; tail -n 1 gpurun_out/bench.log | cut -c1-1500

#!/bin/bash
# first end-to-end GPU run: module/model golden parity, smoke, short bench
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
rm -f gpurun_out/golden_parity.jsonl
timeout -k 10 900 python -m pytest tests/test_gpu_golden.py -m gpu -q --timeout 600 --timeout-method=thread --tb=short > gpurun_out/golden.log 2>&1
echo "== golden exit $?"; tail -n 40 gpurun_out/golden.log
timeout -k 10 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
echo "== smoke exit $?"; tail -n 15 gpurun_out/smoke.log
timeout -k 10 900 python bench.py --steps 2 --warmup 3 --batch 8 --no-cpu-baseline > gpurun_out/bench_b8.log 2>&1
echo "== bench b8 exit $?"; tail -n 12 gpurun_out/bench_b8.log
timeout -k 10 1200 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_b32.log 2>&1
echo "== bench b32 exit $?"; tail -n 12 gpurun_out/bench_b32.log

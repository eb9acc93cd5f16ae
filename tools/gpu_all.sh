#!/bin/bash
# full GPU test suite + smoke + bench (headline config with breakdown, then the other configs) + per-kernel table
mkdir -p gpurun_out
rm -f gpurun_out/*_parity.jsonl
timeout -k 10 1500 python -m pytest tests -m gpu -q --timeout 900 --timeout-method=thread --tb=short -x > gpurun_out/gpu_tests.log 2>&1
echo "== gpu tests exit $?"; tail -n 25 gpurun_out/gpu_tests.log | cut -c1-400
timeout -k 10 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
echo "== smoke exit $?"; tail -n 3 gpurun_out/smoke.log | cut -c1-400
timeout -k 10 900 python bench.py --steps 5 --warmup 3 --breakdown gpurun_out/breakdown_c2.json > gpurun_out/bench_c2.log 2>&1
echo "== bench c2 exit $?"; tail -n 1 gpurun_out/bench_c2.log | cut -c1-3000
for cfg in ${CONFIGS:-c1 wgan stargan inference}; do
  timeout -k 10 600 python bench.py --config $cfg --steps 5 --warmup 3 --no-cpu-baseline --no-incumbent --breakdown gpurun_out/breakdown_$cfg.json > gpurun_out/bench_$cfg.log 2>&1
  echo "== bench $cfg exit $?"; tail -n 1 gpurun_out/bench_$cfg.log | cut -c1-1200
done
timeout -k 10 300 python tools/kernel_times.py > gpurun_out/kernel_times.txt 2>&1
echo "== kernel_times exit $?"; grep -A30 "^kernels" gpurun_out/kernel_times.txt | cut -c1-150

#!/bin/bash
# bench with per-geometry breakdown, then the ncu launch list and one full capture of the top kernels (B200_PROFILING.md)
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --skip-e2e"
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --breakdown gpurun_out/conv_breakdown.json > gpurun_out/bench_breakdown.log 2>&1
echo "== bench exit $?"; tail -n 2 gpurun_out/bench_breakdown.log | cut -c1-600
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 30000 -c 12000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "== ncu list exit $?"; tail -n 3 gpurun_out/ncu_list.log
ncu --set full --clock-control none --import-source on -k regex:'conv_(igemm|wgrad)_kernel' -s 1500 -c 6 -o gpurun_out/prof_conv $CMD > gpurun_out/ncu_full.log 2>&1
echo "== ncu full exit $?"; tail -n 3 gpurun_out/ncu_full.log
ls -la gpurun_out

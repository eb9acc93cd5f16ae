#!/bin/bash
# ncu launch list of the last (timed) step of a short bench run (B200_PROFILING.md recipe); run only after the plain command exits 0
mkdir -p gpurun_out
CMD="env DEI2I_CUDA_GRAPH=0 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-incumbent --skip-e2e"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -n 5 gpurun_out/plain.log; exit 1; }
tail -n 1 gpurun_out/plain.log | cut -c1-300
L=$(python -c "import json;print(json.loads(open('gpurun_out/plain.log').read().strip().splitlines()[-1])['gpu_launches'])")
echo "launches per step: $L"
timeout -k 10 ${NCU_TIMEOUT:-600} ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "== ncu list exit $?"; tail -n 3 gpurun_out/ncu_list.log | cut -c1-300
python tools/summarize_launches.py gpurun_out/launches.csv -1 > gpurun_out/launch_summary.txt 2>&1
head -n 60 gpurun_out/launch_summary.txt

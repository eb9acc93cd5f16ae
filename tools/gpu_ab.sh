#!/bin/bash
# A/B of env switches on the headline bench (device-timed value only); usage: gpu_ab.sh "VAR=a VAR=b ..." (each token = one run)
mkdir -p gpurun_out
for kv in "$@"; do
  env $kv python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-incumbent --skip-e2e > gpurun_out/ab_$(echo $kv | tr '= ' '__').log 2>&1
  echo "== $kv: $(tail -n 1 gpurun_out/ab_$(echo $kv | tr '= ' '__').log | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print(round(d['ms_per_step'],2),'ms', round(d['value'],1),'img/s conv',round(r['all_conv_tflops'],1),'TF/s  stream', round(r['streaming_gbs']),'GB/s share', round(r['streaming_share_of_step'],3), 'launches', d['gpu_launches'])")"
done

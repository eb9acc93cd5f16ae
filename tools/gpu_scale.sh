#!/bin/bash
# Multi-GPU evidence, run as: gpurun --gpus N -- 'bash tools/gpu_scale.sh N'
# 1. isolated all-reduce timing with the NCCL_DEBUG algorithm lines, 2. bench.py at N ranks: default (one all-reduce after backward),
# bucketed overlap, overlap with a capped NCCL footprint, and replicas without any exchange (the max-over-ranks spread alone).
N=${1:-2}
STEPS=${STEPS:-20}
OUT=gpurun_out/scale_n$N
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,COLL,TUNING timeout 300 $TR --master-port 29511 tools/allreduce_probe.py > $OUT/allreduce_probe.log 2>&1
grep -E '^\{' $OUT/allreduce_probe.log > $OUT/allreduce_probe.jsonl
grep -E 'NVLS|channels,|nChannels|Init COMPLETE' $OUT/allreduce_probe.log | grep -v 'via P2P' | head -40 > $OUT/nccl_debug_lines.txt
grep -E 'Bytes -> Algo' $OUT/allreduce_probe.log | sort | uniq -c >> $OUT/nccl_debug_lines.txt
wc -l $OUT/allreduce_probe.log | cat
[ "$N" -le 2 ] && NCCL_MAX_CTAS=4 timeout 300 $TR --master-port 29512 tools/allreduce_probe.py >> $OUT/allreduce_probe.jsonl 2> $OUT/allreduce_probe_ctas4.err
NCCL_MAX_CTAS=8 timeout 300 $TR --master-port 29513 tools/allreduce_probe.py >> $OUT/allreduce_probe.jsonl 2> $OUT/allreduce_probe_ctas8.err
sed -i '/^{/!d' $OUT/allreduce_probe.jsonl
cat $OUT/allreduce_probe.jsonl
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 600 $TR --master-port 29520 bench.py --gpus $N --steps $STEPS --warmup 5 --no-cpu-baseline --no-incumbent > $OUT/bench_$name.json 2> $OUT/bench_$name.err
  tail -1 $OUT/bench_$name.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$name', d['value'], d['ms_per_step'], d['e2e']['value'], d.get('clocks'))"
}
run default DEI2I_DDP_OVERLAP=0
run overlap DEI2I_DDP_OVERLAP=1
[ "$N" -le 2 ] && run overlap_ctas4 DEI2I_DDP_OVERLAP=1 NCCL_MAX_CTAS=4
run overlap_ctas8 DEI2I_DDP_OVERLAP=1 NCCL_MAX_CTAS=8
[ "$N" -le 2 ] && run default_ctas8 DEI2I_DDP_OVERLAP=0 NCCL_MAX_CTAS=8
run no_exchange DEI2I_DDP_SKIP_ALLREDUCE=1
python bench.py --gpus 1 --steps $STEPS --warmup 5 --no-cpu-baseline --no-incumbent > $OUT/bench_n1.json 2> $OUT/bench_n1.err
tail -1 $OUT/bench_n1.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('n1', d['value'], d['ms_per_step'], d['e2e']['value'])"

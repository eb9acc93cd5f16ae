#!/bin/bash
# ncu --set full of the weak kernels in isolation (one launch each), summaries via tools/ncu_brief.py
mkdir -p gpurun_out
python tools/weak_ops.py fwd64 > gpurun_out/weak_ops_plain.txt 2>&1 || { echo plain run failed; tail -n 5 gpurun_out/weak_ops_plain.txt; exit 1; }
cat gpurun_out/weak_ops_plain.txt
run() {  # name, ops, kernel regex, skip
  timeout -k 10 600 ncu --set full --clock-control none --import-source on -k "regex:$3" --launch-skip $4 --launch-count 1 -f -o gpurun_out/ncu_$1 python tools/weak_ops.py $2 > gpurun_out/ncu_$1.log 2>&1
  echo "== ncu $1 exit $?"
  python tools/ncu_brief.py gpurun_out/ncu_$1.ncu-rep > gpurun_out/ncu_$1.txt 2>&1; cat gpurun_out/ncu_$1.txt | cut -c1-160
}
run fwd64 fwd64 "conv_tma_kernel" 2
run wgrad64 wgrad64 "wgrad_tma_kernel" 2
run c256fwd c256 "conv_tma_kernel" 2
run c256wgrad c256 "wgrad_tma_kernel" 2
run stemfwd stem "conv_tma_kernel" 2
run stemwgrad stem "wgrad_tma_kernel" 2
run spade_bwd1 spade_bwd "spade_cls_bwd1_kernel" 2
run spade_bwd2 spade_bwd "spade_cls_bwd2_kernel" 2
run spade_fwd spade_bwd "spade_cls_fwd_kernel" 3
run heads_dgrad heads "conv_c8_fwd_kernel" 2

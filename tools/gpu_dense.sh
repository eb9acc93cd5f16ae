#!/bin/bash
# ncu --set full of the dense tcgen05 kernels (fwd, dgrad, wgrad) on the dominant geometry
mkdir -p gpurun_out
cd tools
python conv_micro.py prof > ../gpurun_out/dense_plain.log 2>&1 || { tail -5 ../gpurun_out/dense_plain.log; exit 1; }
cat ../gpurun_out/dense_plain.log
ncu --set full --clock-control none --import-source on -k regex:'conv_tma_kernel' -s 9 -c 1 -o ../gpurun_out/prof_dense_fwd python conv_micro.py prof > ../gpurun_out/ncu_dense.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'conv_tma_kernel' -s 22 -c 1 -o ../gpurun_out/prof_dense_dgrad python conv_micro.py prof >> ../gpurun_out/ncu_dense.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'wgrad_tma_kernel' -s 9 -c 1 -o ../gpurun_out/prof_dense_wgrad python conv_micro.py prof >> ../gpurun_out/ncu_dense.log 2>&1
echo "ncu exit $?"; grep -c "==PROF== Report" ../gpurun_out/ncu_dense.log

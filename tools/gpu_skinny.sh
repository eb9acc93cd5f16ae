#!/bin/bash
mkdir -p gpurun_out
cd tools
python skinny_micro.py ${1:-all} > ../gpurun_out/skinny.log 2>&1 || { tail -5 ../gpurun_out/skinny.log; exit 1; }
cat ../gpurun_out/skinny.log
if [ -n "$2" ]; then
python skinny_micro.py ${1:-all} > ../gpurun_out/skinny_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"$2" -s ${3:-0} -c ${4:-3} -o ../gpurun_out/prof_skinny python skinny_micro.py ${1:-all} > ../gpurun_out/ncu_skinny.log 2>&1
echo "ncu exit $?"; tail -3 ../gpurun_out/ncu_skinny.log
fi

#!/bin/bash
# compute-sanitizer memcheck + racecheck over the kernel parity tests of the hand-rolled pipelines (SURVEY.md section 5 aux row)
mkdir -p gpurun_out
SEL='tma or spade_cls or heads or epilogue or stats or batchnorm or diffaug or mask'
timeout -k 10 1500 compute-sanitizer --tool memcheck --error-exitcode 3 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_diffaug.py -m gpu -q -x -k "$SEL" --timeout 1400 > gpurun_out/sanitizer_memcheck.log 2>&1
echo "== memcheck exit $?"; grep -E "ERROR SUMMARY|passed|failed" gpurun_out/sanitizer_memcheck.log | tail -3
timeout -k 10 1500 compute-sanitizer --tool racecheck --error-exitcode 3 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "tma_valid3x3 or spade_cls_fwd_bwd or heads_4 or epilogue_statistics or channel_stats" --timeout 1400 > gpurun_out/sanitizer_racecheck.log 2>&1
echo "== racecheck exit $?"; grep -E "RACECHECK SUMMARY|ERROR SUMMARY|passed|failed" gpurun_out/sanitizer_racecheck.log | tail -3

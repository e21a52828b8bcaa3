#!/bin/bash
# ncu --set full on the batched W4A8 kernel alone (tools/gemm_bench.py shapes); run after the plain command exited 0
mkdir -p gpurun_out
python tools/gemm_bench.py > gpurun_out/gemm_bench.log 2>&1 && cat gpurun_out/gemm_bench.log | grep -v Deprec &&
ncu --set full --clock-control none --import-source on -k regex:gemm_w4a8 -s 22 -c 2 -f -o gpurun_out/prof_gemm2 python tools/gemm_bench.py > gpurun_out/ncu_gemm2.log 2>&1
echo "ncu exit $?"

#!/bin/bash
# ncu evidence for one round: (1) per-launch device times of a short decode run, (2) one
# `--set full` capture of the fused GEMV (four consecutive launches = qkv, wo, gate|up, down).
# Each ncu command runs only after the identical command exited 0 without ncu.
mkdir -p gpurun_out
CMD="python bench.py --steps 8 --warmup 3 --ctx 16 --no-cpu"
$CMD > gpurun_out/plain1.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s ${SKIP_ALL:-3500} -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "launch list exit $?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemv_kernel -s ${SKIP_GEMV:-2500} -c 4 -f -o gpurun_out/prof_gemv $CMD > gpurun_out/ncu2.log 2>&1
echo "full capture exit $?"
tail -2 gpurun_out/ncu1.log gpurun_out/ncu2.log

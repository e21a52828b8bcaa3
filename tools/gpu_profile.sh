#!/bin/bash
# ncu evidence for one round: (1) per-launch device times of a short run (chunked prefill of the
# prompt through the batched kernels, then decode steps = step_begin, persistent layer kernel, head
# GEMV, step_end), (2) one `--set full` capture of the persistent kernel, (3) one of the batched
# W4A8 kernel.  Each ncu command runs only after the identical command exited 0 without ncu.
mkdir -p gpurun_out
export B2_MODEL_DIR=/tmp/b2gguf-models
CMD="python bench.py --steps 8 --warmup 3 --ctx ${CTX:-64} --no-cpu --no-batch"
$CMD > gpurun_out/plain1.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c ${COUNT:-1200} --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "launch list exit $?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:mega_kernel -s 3 -c 1 -f -o gpurun_out/prof_mega $CMD > gpurun_out/ncu2.log 2>&1
echo "full capture (persistent kernel) exit $?"
$CMD > gpurun_out/plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_w4a8 -s 8 -c 4 -f -o gpurun_out/prof_gemm $CMD > gpurun_out/ncu3.log 2>&1
echo "full capture (batched kernel) exit $?"
tail -n 2 gpurun_out/ncu1.log; tail -n 2 gpurun_out/ncu2.log; tail -n 2 gpurun_out/ncu3.log

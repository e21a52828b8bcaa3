#!/bin/bash
# Round-2 evidence in one gpurun call: the default bench line, then (each only after the identical command exited 0
# without ncu) the ncu launch list of a short run and one full capture of the persistent kernel.
mkdir -p gpurun_out
export B2_MODEL_DIR=/tmp/b2gguf-models
python bench.py > gpurun_out/r02_bench.json 2> gpurun_out/r02_bench.err; echo "bench exit $?"
CMD="python bench.py --steps 8 --warmup 3 --ctx 512 --no-cpu --no-batch --no-tp"
$CMD > gpurun_out/plain1.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "launch list exit $?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:mega_kernel -s 3 -c 1 -f -o gpurun_out/r02_mega $CMD > gpurun_out/ncu2.log 2>&1
echo "full capture (persistent kernel) exit $?"
tail -c 1200 gpurun_out/r02_bench.json; tail -n 3 gpurun_out/r02_bench.err

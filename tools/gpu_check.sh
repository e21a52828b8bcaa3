#!/bin/bash
# One gpurun call: GPU test-suite (one pytest process per file, so a faulting kernel cannot take
# the other files down), smoke, a GEMV sweep and a short bench.  Logs land in gpurun_out/.
mkdir -p gpurun_out
export B2_MODEL_DIR=/tmp/b2gguf-models
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/gpu.txt 2>&1
for f in test_gpu_kernels test_gpu_model test_gpu_batch test_gpu_tp test_zz_no_alloc; do
  timeout -s KILL 900 python -m pytest tests/$f.py -m gpu -q --timeout=180 --timeout-method=thread -p no:cacheprovider \
      > gpurun_out/$f.log 2>&1
  echo "$f exit $?" | tee -a gpurun_out/summary.txt
done
timeout -s KILL 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/summary.txt
timeout -s KILL 600 python tools/gemv_sweep.py > gpurun_out/gemv_sweep.log 2>&1; echo "sweep exit $?" | tee -a gpurun_out/summary.txt
timeout -s KILL 1200 python bench.py --steps ${STEPS:-128} --warmup 8 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?" | tee -a gpurun_out/summary.txt
for f in test_gpu_kernels test_gpu_model test_gpu_batch test_gpu_tp test_zz_no_alloc smoke; do tail -n 2 gpurun_out/$f.log; done
tail -c 1500 gpurun_out/bench.json

#!/bin/bash
# Builds ollama-operator_b200/libb200gguf.so for sm_100a, in-tree (the .so travels to the GPU box).
set -euo pipefail
cd "$(dirname "$0")/ollama-operator_b200/csrc"
OUT=../libb200gguf.so
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS=(-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-Wall,-Wno-unused-function
       --fmad=true -Xptxas -v -ccbin /usr/bin/g++ ${B2_CXXFLAGS:-})
mkdir -p ../build
objs=()
for f in dequant.cu gemv.cu attn.cu mega.cu batch.cu gemm.cu gemm_tc.cu step.cu engine.cu engine_batch.cu tp.cu api.cu noalloc.cu gguf.cpp tokenizer.cpp tokenizer_api.cpp store.cpp; do
  o=../build/${f%.*}.o
  if [ ! -f "$o" ] || [ "$f" -nt "$o" ] || [ -n "$(find . ../../include -newer "$o" \( -name '*.h' -o -name '*.cuh' \) -print -quit)" ]; then
    echo "nvcc $f" >&2
    "$NVCC" "${FLAGS[@]}" -x cu -c "$f" -o "$o" 2> "../build/${f%.*}.log" || { cat "../build/${f%.*}.log" >&2; exit 1; }
  fi
  objs+=("$o")
done
"$NVCC" -shared -o "$OUT" "${objs[@]}" -gencode arch=compute_100a,code=sm_100a -ccbin /usr/bin/g++ -lcudart_static -lrt -lpthread -ldl
echo "built $(readlink -f $OUT)" >&2
# the HTTP front end (`serve`): uses the C ABI only (cwd is csrc/)
g++ -O2 -std=c++17 -Wall -pthread -o ../../server/b2-server ../../server/b2_server.cpp -L.. -lb200gguf -Wl,-rpath,'$ORIGIN/../ollama-operator_b200'
echo "built $(readlink -f ../../server/b2-server)" >&2

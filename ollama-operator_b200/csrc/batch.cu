// batch.cu -- several activation columns per pass over the weights: batched decode (one column per
// sequence) and chunked prefill (consecutive tokens of one sequence).
//
// The arithmetic is the GEMV's, column by column (ggml quantises every activation row to Q8_0 and
// takes exact integer block dots, also in batches): each 32-weight block contributes
// ((float)sumi * d_w) * d_a, added in f64.  Only the integer part changes engine: one
// mma.sync.m16n8k32.s8 computes the block sums of 16 rows x 8 columns at once.  The tensor-core
// shape is chosen for exactness, not peak: K = 32 is exactly one quantisation block, so every MMA
// starts from a zero accumulator and its s32 results are scaled per (row, column, block) -- that
// per-block scaling (one I2F, two FMUL, one F2F, one DADD per result) is the real cost, and it is
// inherent to block-scaled int4 x int8 with exact ggml semantics.
#include <math.h>

#include <algorithm>

#include "common.cuh"
#include "formats.cuh"
#include "gemv_dev.cuh"
#include "kernels.h"
#include "qmat_dev.cuh"

namespace b2 {

bool& batch_pdl_enabled() {
    static bool on = true;
    return on;
}


namespace {
using namespace gemvdev;

// ---- batch_begin: embedding rows ------------------------------------------------------------------
__global__ void batch_begin_kernel(QMat embd, float* xb, int64_t x_stride, const int* tok, unsigned long long* keys) {
    const int c = blockIdx.y;
    const int t = tok[c];
    const int E = (int)embd.k;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < E; i += gridDim.x * blockDim.x) xb[(size_t)c * x_stride + i] = qmat_elem(embd, t, i);
    if (blockIdx.x == 0 && threadIdx.x == 0) keys[c] = 0ull;
}

__global__ void batch_end_kernel(const unsigned long long* keys, int* next, int B) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < B) next[c] = (int)(0xFFFFFFFFu - (uint32_t)(keys[c] & 0xFFFFFFFFull));
}

// ---- quant_cols: one CTA per column; the GEMV prologue (same statistics, same quantiser as
// prologue_q8_0 in gemv_dev.cuh) writing the chunk-major global layout of Q8Cols -------------------------
constexpr int QC_SLICE = 1024;           // elements one CTA quantises: one 4-element chunk per thread, one global round trip
__global__ void __launch_bounds__(256) quant_cols_kernel(GemvArgs a, int64_t x_stride, Q8Cols out) {
    __shared__ double dscratch[32];
    const int c = blockIdx.x, tid = threadIdx.x;
    a.x += (size_t)c * x_stride;
    pdl_trigger();       // the batched kernel that consumes these columns may start streaming its weights
    pdl_wait();          // the activations come from the preceding kernel, which may also still be reading the buffer we overwrite
    // every CTA of a column computes the column's norm statistics (a streaming pass over K, identical in all of them),
    // then quantises its own slice
    const ProStats st = prologue_stats(a, dscratch, tid, 256);
    const int k0 = blockIdx.y * QC_SLICE, k1 = min((int)a.W.k, k0 + QC_SLICE);
    prologue_q8_0_each(a, st, tid, 256, [=](int blk, int o, uint32_t w, float d, int s8) {
        const int kc = blk / GEMM_KC, bl = blk % GEMM_KC;
        *reinterpret_cast<uint32_t*>(out.q + ((size_t)kc * BATCH_MAX + c) * GEMM_XCOL + (o < 16 ? 0 : GEMM_KC * 16) + bl * 16 + (o & 15)) = w;
        if (o == 0) {
            out.d[((size_t)kc * GEMM_KC + bl) * BATCH_MAX + c] = d;
            out.s8[((size_t)kc * GEMM_KC + bl) * BATCH_MAX + c] = -s8;        // the MMA accumulator starts from it
        }
    }, k0, k1);
}
}  // namespace

int batch_begin_launch(const QMat& embd, float* xb, int64_t x_stride, const int* d_tok, unsigned long long* keys, int B, cudaStream_t st) {
    const int E = (int)embd.k;
    batch_begin_kernel<<<dim3((E + 255) / 256, B), 256, 0, st>>>(embd, xb, x_stride, d_tok, keys);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}
int batch_end_launch(const unsigned long long* keys, int* d_next, int B, cudaStream_t st) {
    batch_end_kernel<<<(B + 127) / 128, 128, 0, st>>>(keys, d_next, B);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}
int quant_cols_launch(const float* x, int64_t x_stride, int64_t K, int B, int pro, const float* gain, const float* nbias, float eps, const Q8Cols& out,
                      cudaStream_t st) {
    if (K % 32 || K > out.k_cap) B2_FAIL(B2_ERR_ARG, "quant_cols: K=%lld", (long long)K);
    GemvArgs a;
    a.W.k = K; a.pro = pro; a.x = x; a.gain = gain; a.nbias = nbias; a.eps = eps;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(B, (unsigned)((K + QC_SLICE - 1) / QC_SLICE));
    cfg.blockDim = dim3(256);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = batch_pdl_enabled() ? 1 : 0;
    B2_CUDA(cudaLaunchKernelEx(&cfg, quant_cols_kernel, a, x_stride, out));
    return B2_OK;
}

}  // namespace b2

// noalloc.cu -- b2_gemv_noalloc: the b2_gemv hook for callers that must not allocate.
//
// Who cannot allocate.  b2_gemv stages its operands with five cudaMalloc/cudaFree pairs per call.
// That is fine for tests, but a serving process that sized its KV cache to fill the GPU at model
// load has nothing left to hand out (the allocation fails with out-of-memory), a process that runs
// inside another framework's memory pool is not allowed to go around that pool, and cudaMalloc is
// a device-wide synchronisation point that a latency-sensitive caller does not want on its path.
// This variant computes the same thing with no allocation at all: every byte of device staging --
// the file-format blocks, their split-plane repack, the activation vector, the norm gain and the
// result -- lives in ONE file-scope device array that is part of the library image.
//
// What that costs and what the caller has to know:
//   * the array is B2_NOALLOC_SCRATCH_BYTES (8 MiB) of HBM on every GPU that loads the library,
//     held until the process exits whether or not this entry point is ever called;
//   * requests whose staging does not fit are rejected with B2_ERR_NOMEM (no partial work);
//   * the first call on a device looks the array's device address up (cudaGetSymbolAddress, with
//     that device current) and remembers it; make that first call before starting any stream
//     capture;
//   * all calls share the array and it keeps whatever the previous call left behind: the regions
//     that must start from zero (the repack planes, whose row padding is never written) are
//     cleared on every call, everything else is fully overwritten before it is read;
//   * for the same reason two calls that overlap on the GPU -- from two host threads, or two
//     processes' worth of streams in one context -- race on the array.  Serialise them.
#include <string.h>

#include "common.cuh"
#include "formats.cuh"
#include "kernels.h"

namespace {

constexpr size_t kScratchBytes = (size_t)B2_NOALLOC_SCRATCH_BYTES;

// In host code this name is only a handle for cudaGetSymbolAddress; it is not a device pointer.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
// Device address of the array, per device.  Written without a lock: two threads that race on a device's first call
// both store the same address (cudaGetSymbolAddress is deterministic per device), so the race is benign.
unsigned char* g_scratch_ptr[kMaxDevices] = {};

// Set-up, once per device, with that device current.
int scratch_base(unsigned char** out) {
    int dev = -1;
    B2_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) B2_FAIL(B2_ERR_UNSUPPORTED, "b2_gemv_noalloc: device ordinal %d not supported", dev);
    if (!g_scratch_ptr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            B2_FAIL(B2_ERR_CUDA, "b2_gemv_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
        }
        g_scratch_ptr[dev] = (unsigned char*)p;
    }
    *out = g_scratch_ptr[dev];
    return B2_OK;
}

inline size_t up256(size_t v) { return (v + 255) / 256 * 256; }

}  // namespace

extern "C" int b2_gemv_noalloc(int ggml_type, const void* W, int64_t rows, int64_t k, const float* x, const float* gain, float eps,
                               float* y) {
    using namespace b2;
    if (!W || !x || !y) B2_FAIL(B2_ERR_ARG, "b2_gemv_noalloc: null buffer");
    if (!type_is_quant(ggml_type)) B2_FAIL(B2_ERR_UNSUPPORTED, "ggml type %d is not supported (Q4_0=2, Q8_0=8, Q4_K=12, Q6_K=14)", ggml_type);
    if (rows <= 0 || k <= 0 || k % type_block_elems(ggml_type) || (rows & 1))
        B2_FAIL(B2_ERR_ARG, "b2_gemv_noalloc: rows=%lld must be positive and even, k=%lld a positive multiple of %d", (long long)rows, (long long)k,
                type_block_elems(ggml_type));
    // a row count or length beyond the scratch itself can never fit; refusing it here also keeps the byte
    // arithmetic below far from wrapping (each product is then below 2^23 * 2^25)
    if ((uint64_t)rows > kScratchBytes || (uint64_t)k > kScratchBytes * 4)
        B2_FAIL(B2_ERR_NOMEM, "b2_gemv_noalloc: a %lld x %lld matrix cannot fit the %zu-byte static scratch; use b2_gemv", (long long)rows, (long long)k,
                kScratchBytes);
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) {
        cudaGetLastError();
        B2_FAIL(B2_ERR_CUDA, "no CUDA device available; this library has no CPU fallback");
    }

    // carve the staging regions out of the static array; refuse what does not fit
    QMat m;
    qmat_shape(m, ggml_type, rows, k);
    const size_t raw_bytes = (size_t)rows * type_row_bytes(ggml_type, k);
    size_t need = up256(raw_bytes);
    size_t plane_off[4] = {0, 0, 0, 0};
    for (int p = 0; p < m.n_planes; p++) { plane_off[p] = need; need += up256((size_t)rows * m.stride[p]); }
    const size_t planes_begin = plane_off[0], planes_end = need;
    const size_t x_off = need; need += up256((size_t)k * 4);
    const size_t g_off = need; need += up256((size_t)k * 4);
    const size_t y_off = need; need += up256((size_t)rows * 4);
    if (need > kScratchBytes)
        B2_FAIL(B2_ERR_NOMEM, "b2_gemv_noalloc: a %lld x %lld matrix of type %d needs %zu bytes of device staging, the static scratch holds %zu; use b2_gemv",
                (long long)rows, (long long)k, ggml_type, need, kScratchBytes);

    unsigned char* base = nullptr;
    B2_TRY(scratch_base(&base));
    for (int p = 0; p < m.n_planes; p++) m.plane[p] = base + plane_off[p];

    // the planes keep the previous call's bytes: clear them (row padding is never written by the repack)
    B2_CUDA(cudaMemsetAsync(base + planes_begin, 0, planes_end - planes_begin, 0));
    B2_CUDA(cudaMemcpyAsync(base, W, raw_bytes, cudaMemcpyHostToDevice, 0));
    B2_CUDA(cudaMemcpyAsync(base + x_off, x, (size_t)k * 4, cudaMemcpyHostToDevice, 0));
    if (gain) B2_CUDA(cudaMemcpyAsync(base + g_off, gain, (size_t)k * 4, cudaMemcpyHostToDevice, 0));
    B2_TRY(repack(ggml_type, base, rows, k, m, 0, 1, 0));
    GemvArgs a;
    a.W = m;
    a.pro = gain ? PRO_RMSNORM : PRO_PLAIN;
    a.epi = EPI_STORE;
    a.x = reinterpret_cast<const float*>(base + x_off);
    a.gain = reinterpret_cast<const float*>(base + g_off);
    a.eps = eps;
    a.y = reinterpret_cast<float*>(base + y_off);
    B2_TRY(gemv_launch(a, 0, false));
    B2_CUDA(cudaMemcpyAsync(y, base + y_off, (size_t)rows * 4, cudaMemcpyDeviceToHost, 0));
    B2_CUDA(cudaStreamSynchronize(0));
    return B2_OK;
}

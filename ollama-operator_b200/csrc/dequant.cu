// dequant.cu -- block dequantisation (bit-exact with ggml's dequantize_row_* as restated by
// gguf/quants.py:241-251,396-401,504-522,554-572), the load-time repack of file blocks into
// split planes, and row extraction from the planes (ggml get_rows for the embedding).
//
// Every float operation below is written with explicit single-rounding intrinsics so that
// nvcc cannot contract a multiply and an add into an FMA: the golden vectors were produced
// by numpy, which rounds after every operation.
#include "common.cuh"
#include "formats.cuh"
#include "kernels.h"
#include "qmat_dev.cuh"

namespace b2 {

// ---- raw file blocks -> f32, one thread per element ------------------------------------------
__global__ void dequant_raw_kernel(int type, const uint8_t* __restrict__ src, int64_t n, float* __restrict__ out) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    float v = 0.f;
    switch (type) {
    case T_F32: v = reinterpret_cast<const float*>(src)[e]; break;
    case T_F16: v = h2f_bits(reinterpret_cast<const uint16_t*>(src)[e]); break;
    case T_Q4_0: {
        const BlockQ4_0* b = reinterpret_cast<const BlockQ4_0*>(src) + e / 32;
        const int j = (int)(e % 32);
        const int q = j < 16 ? (b->qs[j] & 0x0F) : (b->qs[j - 16] >> 4);
        v = __fmul_rn((float)(q - 8), h2f_bits(b->d));
    } break;
    case T_Q8_0: {
        const BlockQ8_0* b = reinterpret_cast<const BlockQ8_0*>(src) + e / 32;
        v = __fmul_rn((float)b->qs[e % 32], h2f_bits(b->d));
    } break;
    case T_Q4_K: {
        const BlockQ4_K* b = reinterpret_cast<const BlockQ4_K*>(src) + e / 256;
        const int j = (int)(e % 256), sub = j / 32, l = j % 32;
        int sc, mn;
        q4k_scale_min(sub, b->scales, sc, mn);
        const uint8_t byte = b->qs[(sub / 2) * 32 + l];
        const int q = (sub & 1) ? (byte >> 4) : (byte & 0x0F);
        const float d = __fmul_rn(h2f_bits(b->d), (float)sc), m = __fmul_rn(h2f_bits(b->dmin), (float)mn);
        v = __fsub_rn(__fmul_rn(d, (float)q), m);
    } break;
    case T_Q6_K: {
        const BlockQ6_K* b = reinterpret_cast<const BlockQ6_K*>(src) + e / 256;
        const int j = (int)(e % 256), half = j / 128, r = j % 128, quad = r / 32, l = r % 32;
        const uint8_t lo = b->ql[half * 64 + (quad & 1) * 32 + l];
        const int q4 = (quad & 2) ? (lo >> 4) : (lo & 0x0F);
        const int q2 = (b->qh[half * 32 + l] >> (2 * quad)) & 3;
        const int q = (int)(int8_t)(q4 | (q2 << 4)) - 32;
        const int sc = b->scales[half * 8 + quad * 2 + l / 16];
        v = __fmul_rn(__fmul_rn(h2f_bits(b->d), (float)sc), (float)q);
    } break;
    default: break;
    }
    out[e] = v;
}

int dequant_raw(int type, const void* d_blocks, int64_t n_elem, float* d_out, cudaStream_t st) {
    if (n_elem == 0) return B2_OK;
    const int threads = 256;
    const int64_t blocks = (n_elem + threads - 1) / threads;
    dequant_raw_kernel<<<(unsigned)blocks, threads, 0, st>>>(type, (const uint8_t*)d_blocks, n_elem, d_out);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}

// ---- repack: file blocks -> planes, one thread per 16-bit word of the file ----------------------
// All four block sizes are even and every field starts on an even offset, so a block is a
// whole number of 16-bit words and each word belongs to exactly one plane.
__global__ void repack_kernel(int type, const uint16_t* __restrict__ src, int64_t rows, int64_t nb, QMat dst,
                              int64_t dst_row0, int64_t dst_row_step) {
    const int wpb = type == T_Q4_0 ? 9 : type == T_Q8_0 ? 17 : type == T_Q4_K ? 72 : 105;  // words per block
    const int64_t total = rows * nb * wpb;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int w = (int)(i % wpb);
    const int64_t blk = (i / wpb) % nb, row = i / ((int64_t)wpb * nb);
    const uint16_t v = src[i];
    const int64_t drow = dst_row0 + row * dst_row_step;
    int p, off;   // plane and byte offset inside the block's slice of that plane
    switch (type) {
    case T_Q4_0: if (w == 0) { p = 1; off = 0; } else { p = 0; off = (w - 1) * 2; } break;
    case T_Q8_0: if (w == 0) { p = 1; off = 0; } else { p = 0; off = (w - 1) * 2; } break;
    case T_Q4_K:
        if (w < 2) { p = 2; off = w * 2; }
        else if (w < 8) { p = 1; off = (w - 2) * 2; }
        else { p = 0; off = (w - 8) * 2; }
        break;
    default:  // Q6_K
        if (w < 64) { p = 0; off = w * 2; }
        else if (w < 96) { p = 1; off = (w - 64) * 2; }
        else if (w < 104) { p = 2; off = (w - 96) * 2; }
        else { p = 3; off = 0; }
        break;
    }
    // byte offset inside the destination row (see qmat_dev.cuh for the chunk-interleaved planes)
    size_t ro;
    if (type == T_Q4_0) ro = p == 0 ? (size_t)blk * 16 + off : (size_t)blk * 2;
    else if (type == T_Q8_0) ro = p == 0 ? il_off(nb, blk, off) : (size_t)blk * 2;
    else if (type == T_Q4_K) ro = p == 0 ? il_off(2 * nb, blk * 2 + off / 64, off % 64) : p == 1 ? (size_t)blk * 12 + off : (size_t)blk * 4 + off;
    else ro = p == 0 ? il_off(2 * nb, blk * 2 + off / 64, off % 64)
            : p == 1 ? il_off(2 * nb, blk * 2 + off / 32, off % 32)
            : p == 2 ? (size_t)blk * 16 + off : (size_t)blk * 2;
    uint8_t* d = dst.plane[p] + (size_t)drow * dst.stride[p] + ro;
    *reinterpret_cast<uint16_t*>(d) = v;
}

int repack(int type, const void* d_raw, int64_t rows, int64_t k, const QMat& dst, int64_t dst_row0, int64_t dst_row_step,
           cudaStream_t st) {
    if (!type_is_quant(type) || dst.type != type || dst.k != k) B2_FAIL(B2_ERR_ARG, "repack: type/shape mismatch");
    if (k % type_block_elems(type)) B2_FAIL(B2_ERR_ARG, "repack: k=%lld not a multiple of the block size", (long long)k);
    if (dst_row0 + (rows - 1) * dst_row_step >= dst.rows) B2_FAIL(B2_ERR_ARG, "repack: destination rows out of range");
    if (rows == 0) return B2_OK;
    const int64_t nb = k / type_block_elems(type);
    const int64_t total = rows * nb * (type_block_bytes(type) / 2);
    const int threads = 256;
    repack_kernel<<<(unsigned)((total + threads - 1) / threads), threads, 0, st>>>(type, (const uint16_t*)d_raw, rows, nb, dst,
                                                                                    dst_row0, dst_row_step);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}

// ---- planes -> f32 (qmat_elem lives in qmat_dev.cuh) ----------------------------------------------
__global__ void dequant_rows_kernel(QMat m, int64_t row0, int64_t nrows, float* __restrict__ out) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nrows * m.k) return;
    out[e] = qmat_elem(m, row0 + e / m.k, e % m.k);
}

int dequant_rows(const QMat& m, int64_t row0, int64_t nrows, float* d_out, cudaStream_t st) {
    if (row0 < 0 || row0 + nrows > m.rows) B2_FAIL(B2_ERR_ARG, "dequant_rows: rows [%lld,%lld) out of range", (long long)row0,
                                                    (long long)(row0 + nrows));
    const int64_t n = nrows * m.k;
    if (n == 0) return B2_OK;
    dequant_rows_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(m, row0, nrows, d_out);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}

}  // namespace b2

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

namespace {
using namespace gemvdev;

__host__ __device__ inline size_t q8_col_bytes(int64_t k_cap) { return (size_t)k_cap; }     // two planes of k/2 bytes

// ---- batch_begin: embedding rows ------------------------------------------------------------------
__global__ void batch_begin_kernel(QMat embd, float* xb, int64_t x_stride, const int* tok, unsigned long long* keys) {
    const int c = blockIdx.y;
    const int t = tok[c];
    const int E = (int)embd.k;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < E; i += gridDim.x * blockDim.x) xb[(size_t)c * x_stride + i] = qmat_elem(embd, t, i);
    if (blockIdx.x == 0 && threadIdx.x == 0) keys[c] = 0ull;
}

__global__ void batch_end_kernel(const unsigned long long* keys, int* next, int B) {
    const int c = threadIdx.x;
    if (c < B) next[c] = (int)(0xFFFFFFFFu - (uint32_t)(keys[c] & 0xFFFFFFFFull));
}

// ---- quant_cols: one CTA per column, the GEMV prologue writing to global memory -----------------------
__global__ void __launch_bounds__(256) quant_cols_kernel(GemvArgs a, int64_t x_stride, Q8Cols out) {
    __shared__ double dscratch[32];
    const int c = blockIdx.x, tid = threadIdx.x;
    a.x += (size_t)c * x_stride;
    const int nbk_cap = (int)(out.k_cap / 32);
    uint8_t* xq = out.q + (size_t)c * out.k_cap;
    float* xd = out.d + (size_t)c * nbk_cap;
    int* xs = out.s8 + (size_t)c * nbk_cap;
    const ProStats st = prologue_stats(a, dscratch, tid, 256);
    prologue_q8_0(a, st, xq, xd, xs, tid, 256);        // planes at xq and xq + (K/32)*16
}
}  // namespace

int batch_begin_launch(const QMat& embd, float* xb, int64_t x_stride, const int* d_tok, unsigned long long* keys, int B, cudaStream_t st) {
    const int E = (int)embd.k;
    batch_begin_kernel<<<dim3((E + 255) / 256, B), 256, 0, st>>>(embd, xb, x_stride, d_tok, keys);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}
int batch_end_launch(const unsigned long long* keys, int* d_next, int B, cudaStream_t st) {
    batch_end_kernel<<<1, 32, 0, st>>>(keys, d_next, B);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}
int quant_cols_launch(const float* x, int64_t x_stride, int64_t K, int B, int pro, const float* gain, const float* nbias, float eps, const Q8Cols& out,
                      cudaStream_t st) {
    if (K % 32 || K > out.k_cap) B2_FAIL(B2_ERR_ARG, "quant_cols: K=%lld", (long long)K);
    GemvArgs a;
    a.W.k = K; a.pro = pro; a.x = x; a.gain = gain; a.nbias = nbias; a.eps = eps;
    quant_cols_kernel<<<B, 256, 0, st>>>(a, x_stride, out);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}

}  // namespace b2

// =====================================================================================================
// gemm_w4a8: Q4_0 weights x up to 32 Q8_0 activation columns
// =====================================================================================================
namespace b2 {
namespace {

constexpr int GB_WARPS = 8;                    // consumer warps; each owns 16 rows of a tile
constexpr int GB_ROWS = 16 * GB_WARPS;         // rows per tile
constexpr int GB_KC = 16;                      // 32-weight blocks per K chunk
constexpr int GB_ROWQ = GB_KC * 16 + 16;       // shared-memory row stride of the nibbles (padded: conflict-free LDS.32)
constexpr int GB_ROWD = GB_KC * 2;             // scales per row and chunk
constexpr int GB_STAGE = GB_ROWS * (GB_ROWQ + GB_ROWD);
constexpr int GB_NS = 3;
constexpr int GB_XCOL = 2 * GB_KC * 16 + 16;   // shared-memory stride of one activation column's chunk (padded)

struct GemmCfg {
    int rows_cta_max;                          // capacity of the per-CTA row-sum buffer
    uint32_t off_x, off_xd, off_xs, off_ysum, off_ring;
};

__device__ __forceinline__ void mma_s8_16x8x32(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%12,%13};"
        : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(0), "r"(0), "r"(0), "r"(0));
}

template <int NT>     // NT = 8-column MMA tiles (columns padded to 8 * NT)
__global__ void __launch_bounds__((GB_WARPS + 1) * 32, 1) gemm_w4a8_kernel(const QMat W, const Q8Cols xin, int B, const GemvArgs* __restrict__ colargs,
                                                                             const int* __restrict__ col_pos, const int* __restrict__ col_pt, int pt_stride,
                                                                             const GemmCfg cfg) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int BP = 8 * NT;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    uint64_t* empty = full + GB_NS;
    uint8_t* xsm = smem + cfg.off_x;                                  // [BP][GB_XCOL]
    float* xd = reinterpret_cast<float*>(smem + cfg.off_xd);          // [BP][GB_KC]
    int* xs = reinterpret_cast<int*>(smem + cfg.off_xs);              // [BP][GB_KC]
    double* ysum = reinterpret_cast<double*>(smem + cfg.off_ysum);    // [rows_cta][BP]
    uint8_t* ring = smem + cfg.off_ring;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    constexpr int ntc = GB_WARPS * 32;
    const int nbk = (int)(W.k / 32), n_kc = (nbk + GB_KC - 1) / GB_KC;
    const int64_t pairs = W.rows >> 1;
    const int64_t r_begin = 2 * ((int64_t)blockIdx.x * pairs / gridDim.x), r_end = 2 * ((int64_t)(blockIdx.x + 1) * pairs / gridDim.x);
    const int rows_cta = (int)(r_end - r_begin), n_tiles = (rows_cta + GB_ROWS - 1) / GB_ROWS;

    if (tid == 0) {
        for (int s = 0; s < GB_NS; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], GB_WARPS); }
        mbar_fence_init();
    }
    __syncthreads();

    if (warp == GB_WARPS) {
        // ===== producer warp: one bulk copy per row and plane (a K chunk of a row is 256 + 32 bytes) =====
        const uint64_t pol = l2_policy_evict_first();
        uint32_t it = 0;
        for (int kc = 0; kc < n_kc; kc++) {
            const int kb = min(GB_KC, nbk - kc * GB_KC);
            for (int tl = 0; tl < n_tiles; tl++, it++) {
                const uint32_t s = it % GB_NS;
                const int64_t row0 = r_begin + (int64_t)tl * GB_ROWS;
                const int rows_here = (int)min((int64_t)GB_ROWS, r_end - row0);
                if (lane == 0) {
                    if (it >= (uint32_t)GB_NS) mbar_wait(&empty[s], ((it / GB_NS) - 1) & 1);
                    mbar_expect_tx(&full[s], (uint32_t)rows_here * (uint32_t)(kb * 18));
                }
                __syncwarp();
                uint8_t* sq = ring + (size_t)s * GB_STAGE;
                uint8_t* sd = sq + GB_ROWS * GB_ROWQ;
                for (int r = lane; r < rows_here; r += 32) {
                    bulk_g2s(sq + r * GB_ROWQ, W.plane[0] + (size_t)(row0 + r) * W.stride[0] + (size_t)kc * GB_KC * 16, (uint32_t)kb * 16, &full[s], pol);
                    bulk_g2s(sd + r * GB_ROWD, W.plane[1] + (size_t)(row0 + r) * W.stride[1] + (size_t)kc * GB_KC * 2, (uint32_t)kb * 2, &full[s], pol);
                }
            }
        }
        return;
    }

    // ===== consumers =====
    for (int i = tid; i < rows_cta * BP; i += ntc) ysum[i] = 0.0;
    const int g = lane >> 2, t = lane & 3;
    uint32_t it = 0;
    for (int kc = 0; kc < n_kc; kc++) {
        const int kb = min(GB_KC, nbk - kc * GB_KC);
        named_bar_sync(1, ntc);                     // everybody is done with the previous chunk of x (and ysum is zeroed)
        // stage this K chunk of every column: 16-byte pieces of the two planes, scales and sums
        for (int i = tid; i < BP * 2 * GB_KC; i += ntc) {
            const int c = i / (2 * GB_KC), pl = (i / GB_KC) & 1, b = i % GB_KC;
            uint4 v = make_uint4(0, 0, 0, 0);
            if (c < B && b < kb) v = *reinterpret_cast<const uint4*>(xin.q + (size_t)c * xin.k_cap + ((size_t)pl * nbk + kc * GB_KC + b) * 16);
            *reinterpret_cast<uint4*>(xsm + c * GB_XCOL + (pl * GB_KC + b) * 16) = v;
        }
        for (int i = tid; i < BP * GB_KC; i += ntc) {
            const int c = i / GB_KC, b = i % GB_KC;
            const bool ok = c < B && b < kb;
            xd[i] = ok ? xin.d[(size_t)c * (xin.k_cap / 32) + kc * GB_KC + b] : 0.f;
            xs[i] = ok ? xin.s8[(size_t)c * (xin.k_cap / 32) + kc * GB_KC + b] : 0;
        }
        named_bar_sync(1, ntc);
        for (int tl = 0; tl < n_tiles; tl++, it++) {
            const uint32_t s = it % GB_NS;
            mbar_wait(&full[s], (it / GB_NS) & 1);
            const int rows_here = min(GB_ROWS, rows_cta - tl * GB_ROWS);
            if (16 * warp < rows_here) {
                const uint8_t* sq = ring + (size_t)s * GB_STAGE + (16 * warp) * GB_ROWQ;
                const __half* sd = reinterpret_cast<const __half*>(ring + (size_t)s * GB_STAGE + GB_ROWS * GB_ROWQ) + (16 * warp) * GB_KC;
                double acc[NT][4];
#pragma unroll
                for (int j = 0; j < NT; j++)
#pragma unroll
                    for (int q = 0; q < 4; q++) acc[j][q] = 0.0;
                for (int b = 0; b < kb; b++) {
                    const uint32_t w0 = *reinterpret_cast<const uint32_t*>(sq + g * GB_ROWQ + b * 16 + 4 * t);
                    const uint32_t w1 = *reinterpret_cast<const uint32_t*>(sq + (g + 8) * GB_ROWQ + b * 16 + 4 * t);
                    const float dw0 = __half2float(sd[g * GB_KC + b]), dw1 = __half2float(sd[(g + 8) * GB_KC + b]);
                    const uint32_t a0 = w0 & 0x0F0F0F0Fu, a2 = (w0 >> 4) & 0x0F0F0F0Fu, a1 = w1 & 0x0F0F0F0Fu, a3 = (w1 >> 4) & 0x0F0F0F0Fu;
#pragma unroll
                    for (int j = 0; j < NT; j++) {
                        const uint8_t* xc = xsm + (8 * j + g) * GB_XCOL + b * 16 + 4 * t;
                        const uint32_t b0 = *reinterpret_cast<const uint32_t*>(xc), b1 = *reinterpret_cast<const uint32_t*>(xc + GB_KC * 16);
                        int c[4];
                        mma_s8_16x8x32(c, a0, a1, a2, a3, b0, b1);
                        const int c0 = 8 * j + 2 * t;
                        const float da0 = xd[c0 * GB_KC + b], da1 = xd[(c0 + 1) * GB_KC + b];
                        const int s0 = xs[c0 * GB_KC + b], s1 = xs[(c0 + 1) * GB_KC + b];
                        acc[j][0] += (double)__fmul_rn(__fmul_rn((float)(c[0] - s0), dw0), da0);
                        acc[j][1] += (double)__fmul_rn(__fmul_rn((float)(c[1] - s1), dw0), da1);
                        acc[j][2] += (double)__fmul_rn(__fmul_rn((float)(c[2] - s0), dw1), da0);
                        acc[j][3] += (double)__fmul_rn(__fmul_rn((float)(c[3] - s1), dw1), da1);
                    }
                }
                // every (row, column) of the CTA belongs to exactly one thread: plain read-modify-write
                const int rl0 = tl * GB_ROWS + 16 * warp + g, rl1 = rl0 + 8;
#pragma unroll
                for (int j = 0; j < NT; j++) {
                    const int c0 = 8 * j + 2 * t;
                    if (rl0 < rows_cta) { ysum[rl0 * BP + c0] += acc[j][0]; ysum[rl0 * BP + c0 + 1] += acc[j][1]; }
                    if (rl1 < rows_cta) { ysum[rl1 * BP + c0] += acc[j][2]; ysum[rl1 * BP + c0 + 1] += acc[j][3]; }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
        }
    }
    named_bar_sync(1, ntc);
    // ===== epilogue: one thread per (row pair, column), the GEMV's epilogues with that column's operands =====
    const int npairs = rows_cta >> 1;
    for (int i = tid; i < npairs * B; i += ntc) {
        const int p = i / B, c = i % B;
        unsigned long long best = 0ull;
        const GemvArgs& a = colargs[c];
        epilogue_pair(a, a.epi_row0 + r_begin + 2 * p, (float)ysum[(2 * p) * BP + c], (float)ysum[(2 * p + 1) * BP + c], best, col_pos[c],
                      col_pt + (size_t)c * pt_stride);
        if (a.epi == EPI_LOGITS && best != 0ull) atomicMax(a.argmax_key, best);
    }
}

template <int NT>
int gemm_launch_t(const QMat& W, const Q8Cols& x, int B, const GemvArgs* colargs, const int* pos, const int* pt, int pt_stride, const GemmCfg& cfg,
                  size_t smem, int grid, cudaStream_t st) {
    auto kern = gemm_w4a8_kernel<NT>;
    static size_t configured = 0;
    if (smem > configured) {
        B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    kern<<<grid, (GB_WARPS + 1) * 32, smem, st>>>(W, x, B, colargs, pos, pt, pt_stride, cfg);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}
}  // namespace

int gemm_w4a8_launch(const QMat& W, const Q8Cols& x, int B, const GemvArgs* d_colargs, const int* d_pos, const int* d_pt, int pt_stride, cudaStream_t st) {
    if (W.type != T_Q4_0) B2_FAIL(B2_ERR_UNSUPPORTED, "gemm: weight type %d (only Q4_0 has a batched kernel)", W.type);
    if (B < 1 || B > BATCH_MAX) B2_FAIL(B2_ERR_ARG, "gemm: %d columns (1..%d)", B, BATCH_MAX);
    if ((W.rows & 1) || (W.k % 256) || W.k > x.k_cap) B2_FAIL(B2_ERR_UNSUPPORTED, "gemm: shape %lld x %lld", (long long)W.rows, (long long)W.k);
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = (int)std::min<int64_t>(W.rows / 2, sms > 0 ? sms : 148);
    const int nt = (B + 7) / 8, bp = 8 * nt;
    GemmCfg cfg;
    cfg.rows_cta_max = (int)((W.rows / 2 + grid - 1) / grid) * 2 + 2;
    size_t off = 2 * GB_NS * 8;
    off = (off + 15) / 16 * 16;
    cfg.off_x = (uint32_t)off; off += (size_t)bp * GB_XCOL;
    cfg.off_xd = (uint32_t)off; off += (size_t)bp * GB_KC * 4;
    cfg.off_xs = (uint32_t)off; off += (size_t)bp * GB_KC * 4;
    off = (off + 15) / 16 * 16;
    cfg.off_ysum = (uint32_t)off; off += (size_t)cfg.rows_cta_max * bp * 8;
    off = (off + 127) / 128 * 128;
    cfg.off_ring = (uint32_t)off; off += (size_t)GB_NS * GB_STAGE;
    if (off > 227 * 1024) B2_FAIL(B2_ERR_UNSUPPORTED, "gemm: %zu bytes of shared memory needed", off);
    switch (nt) {
    case 1: return gemm_launch_t<1>(W, x, B, d_colargs, d_pos, d_pt, pt_stride, cfg, off, grid, st);
    case 2: return gemm_launch_t<2>(W, x, B, d_colargs, d_pos, d_pt, pt_stride, cfg, off, grid, st);
    case 3: return gemm_launch_t<3>(W, x, B, d_colargs, d_pos, d_pt, pt_stride, cfg, off, grid, st);
    default: return gemm_launch_t<4>(W, x, B, d_colargs, d_pos, d_pt, pt_stride, cfg, off, grid, st);
    }
}

}  // namespace b2

// gemm.cu -- Q4_0 weights x up to 32 Q8_0 activation columns per pass over the weights
// (batched decode, chunked prefill).
//
// The arithmetic is the GEMV's, column by column (ggml quantises every activation row to Q8_0 and
// takes exact integer block dots, also in batches): each 32-weight block contributes
// ((float)sumi * d_w) * d_a, added in f64.  Only the integer part changes engine: one
// mma.sync.m16n8k32.s8 computes the block sums of 16 rows x 8 columns at once.  The tensor-core
// shape is chosen for exactness, not peak: K = 32 is exactly one quantisation block, so every MMA
// starts from a zero accumulator and its s32 results are scaled per (row, column, block) -- that
// per-block scaling (I2F, two FMUL, F2F to f64, DADD per result) is the real cost, and it is
// inherent to block-scaled int4 x int8 with exact ggml semantics.
//
// Data movement: a producer warp fills a 3-stage shared-memory ring with TMA.  A stage is 16
// blocks (512 weights) of K for up to 10 row tiles of 16 rows: per row tile two 2-D tensor-map
// boxes of [16 rows x 128 B] nibbles (hardware 128-byte swizzle: the A-fragment loads of eight
// rows at the same K offset hit eight different 16-byte chunks, no bank conflicts), one box of
// [16 rows x 32 B] scales, and this K chunk of the quantised activation columns (one bulk copy:
// quant_cols_kernel writes them chunk-major, already in the padded shared-memory layout).
// Work units are (row tile, 8-column tile); warp w owns units w, w+8, ... and keeps
// their f64 accumulators in registers over the whole K loop, so no two warps ever touch the same
// output.  The -8 * sum(q) correction of the unsigned nibbles enters as the MMA's C operand.
#include <cuda.h>
#include <cudaTypedefs.h>
#include <math.h>

#include <algorithm>

#include "common.cuh"
#include "formats.cuh"
#include "gemv_dev.cuh"
#include "kernels.h"

namespace b2 {

namespace {
using namespace gemvdev;

constexpr int GB_WARPS = 16;                      // consumer warps
constexpr int GB_NTC = GB_WARPS * 32;
constexpr int GB_RTMAX = 10;                      // row tiles (of 16 rows) per row group
constexpr int GB_MAXI = 3;                       // work units per warp: RTMAX * 4 column tiles / 8 warps
constexpr int GB_NS = 3;
constexpr int GB_QBOX = 16 * 128;                 // one swizzled box: 16 rows x 8 blocks of nibbles
constexpr int GB_DBOX = 16 * GEMM_KC * 2;         // scales of 16 rows x 16 blocks
constexpr int GB_OFF_D = GB_RTMAX * 2 * GB_QBOX;
constexpr int GB_OFF_XQ = GB_OFF_D + GB_RTMAX * GB_DBOX;
constexpr int GB_OFF_XD = GB_OFF_XQ + BATCH_MAX * GEMM_XCOL;
constexpr int GB_OFF_XS = GB_OFF_XD + GEMM_KC * BATCH_MAX * 4;
constexpr int GB_STAGE = (GB_OFF_XS + GEMM_KC * BATCH_MAX * 4 + 1023) / 1024 * 1024;
constexpr int GB_OUT_STRIDE = 33;
constexpr int GB_OUT_BYTES = GB_RTMAX * 16 * GB_OUT_STRIDE * 4;
constexpr size_t GB_SMEM = 1024 + (size_t)GB_NS * GB_STAGE + GB_OUT_BYTES + 2 * GB_NS * 8;
static_assert(GB_OFF_D % 128 == 0 && GB_OFF_XQ % 128 == 0 && GB_OFF_XD % 16 == 0, "stage layout alignment");
static_assert(GB_SMEM <= 227 * 1024, "shared memory budget");

// c = A . B + {cx, cy, cx, cy}: the accumulator starts from the per-column constant
__device__ __forceinline__ void mma_s8_16x8x32(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1, int cx, int cy) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%10,%11};"
        : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(cx), "r"(cy));
}

// 2-D tiled TMA load: box of the tensor map at (c0 = byte offset inside the row, c1 = row)
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2, %3}], [%4], %5;" ::"r"(
            smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}
__device__ __forceinline__ float half_of(uint32_t packed, int hi) {
    return __half2float(__ushort_as_half((unsigned short)(hi ? packed >> 16 : packed & 0xFFFFu)));
}

// NT = 8-column MMA tiles (columns padded to 8 * NT); SPLIT: every CTA has so few units (row tiles x NT <= 8)
// that two warps share one, each taking one half of every K chunk
template <int NT, bool SPLIT>
__global__ void __launch_bounds__((GB_WARPS + 1) * 32, 1)
gemm_w4a8_kernel(const __grid_constant__ GemmMaps maps, const Q8Cols xin, const int rows, const int nbk, const int B, const GemvArgs* __restrict__ colargs,
                 const int* __restrict__ col_pos, const int* __restrict__ col_pt, const int pt_stride) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);      // swizzled boxes need 1024-byte alignment
    uint8_t* ring = smem;
    float* outb = reinterpret_cast<float*>(smem + (size_t)GB_NS * GB_STAGE);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)GB_NS * GB_STAGE + GB_OUT_BYTES);
    uint64_t* empty = full + GB_NS;
    constexpr int BP = 8 * NT;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int T = (rows + 15) >> 4;
    const int t0 = (int)((int64_t)blockIdx.x * T / gridDim.x), t1 = (int)((int64_t)(blockIdx.x + 1) * T / gridDim.x);
    const int nt_cta = t1 - t0;
    const int n_grp = (nt_cta + GB_RTMAX - 1) / GB_RTMAX, per = n_grp ? (nt_cta + n_grp - 1) / n_grp : 0;
    const int n_kc = (nbk + GEMM_KC - 1) / GEMM_KC;

    if (tid == 0) {
        for (int s = 0; s < GB_NS; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], GB_WARPS); }
        mbar_fence_init();
    }
    __syncthreads();

    pdl_trigger();
    if (warp == GB_WARPS) {
        // ===== producer warp: every lane issues (at most) one TMA operation per stage =====
        // The weights do not depend on the preceding kernel: their boxes for the first stages are issued
        // before griddepcontrol.wait, the activation chunks (written by the preceding kernel) after it.
        const uint64_t pol_w = l2_policy_evict_first(), pol_x = l2_policy_evict_last();
        const uint32_t n_it = (uint32_t)n_grp * (uint32_t)n_kc;
        auto stage_of = [&](uint32_t it, int& gt0, int& n_rt, int& kc) {
            const int grp = (int)(it / (uint32_t)n_kc);
            kc = (int)(it % (uint32_t)n_kc);
            gt0 = t0 + grp * per;
            n_rt = min(per, t1 - gt0);
        };
        auto issue_w = [&](uint32_t it) {
            int gt0, n_rt, kc;
            stage_of(it, gt0, n_rt, kc);
            const uint32_t s = it % GB_NS;
            if (lane == 0) mbar_expect_tx(&full[s], (uint32_t)n_rt * (2 * GB_QBOX + GB_DBOX) + BP * GEMM_XCOL + 2 * GEMM_KC * BATCH_MAX * 4);
            __syncwarp();
            uint8_t* sb = ring + (size_t)s * GB_STAGE;
            if (lane < 2 * n_rt) {
                const int rt = lane >> 1, sub = lane & 1;
                tma_load_2d(sb + (rt * 2 + sub) * GB_QBOX, &maps.q, (kc * GEMM_KC + sub * 8) * 16, (gt0 + rt) * 16, &full[s], pol_w);
            } else if (lane >= 2 * GB_RTMAX && lane - 2 * GB_RTMAX < n_rt) {
                const int rt = lane - 2 * GB_RTMAX;
                tma_load_2d(sb + GB_OFF_D + rt * GB_DBOX, &maps.d, kc * GEMM_KC * 2, (gt0 + rt) * 16, &full[s], pol_w);
            }
        };
        auto issue_x = [&](uint32_t it) {
            const int kc = (int)(it % (uint32_t)n_kc);
            const uint32_t s = it % GB_NS;
            uint8_t* sb = ring + (size_t)s * GB_STAGE;
            if (lane == 30) {
                bulk_g2s(sb + GB_OFF_XQ, xin.q + (size_t)kc * BATCH_MAX * GEMM_XCOL, BP * GEMM_XCOL, &full[s], pol_x);
            } else if (lane == 31) {
                bulk_g2s(sb + GB_OFF_XD, xin.d + (size_t)kc * GEMM_KC * BATCH_MAX, GEMM_KC * BATCH_MAX * 4, &full[s], pol_x);
                bulk_g2s(sb + GB_OFF_XS, xin.s8 + (size_t)kc * GEMM_KC * BATCH_MAX, GEMM_KC * BATCH_MAX * 4, &full[s], pol_x);
            }
        };
        const uint32_t n_pre = min(n_it, (uint32_t)GB_NS);
        for (uint32_t it = 0; it < n_pre; it++) issue_w(it);
        pdl_wait();
        for (uint32_t it = 0; it < n_pre; it++) issue_x(it);
        for (uint32_t it = n_pre; it < n_it; it++) {
            const uint32_t s = it % GB_NS;
            if (lane == 0) mbar_wait(&empty[s], ((it / GB_NS) - 1) & 1);
            __syncwarp();
            issue_w(it);
            issue_x(it);
        }
        return;
    }

    // ===== consumers =====
    // Work units are (row tile, 8-column tile) pairs; warp w owns units w, w+16, ... (warps w and w+4 share an
    // SM sub-partition, so the four schedulers carry the same load for the usual 20 or 24 units).  With few
    // units (small batches, narrow matrices) two warps share a unit, each taking one half of every K chunk;
    // their f64 partial sums meet in shared memory, in a fixed order.
    const int g = lane >> 2, t = lane & 3;
    const bool partial = colargs[0].epi == EPI_PARTIAL;      // tensor-parallel K slice: the f64 sums leave unrounded (same for every column)
    uint32_t it = 0;
    for (int grp = 0; grp < n_grp; grp++) {
        const int gt0 = t0 + grp * per, n_rt = min(per, t1 - gt0), tot = n_rt * NT;
        constexpr int KS = SPLIT ? 2 : 1;
        double acc[GB_MAXI][4];
#pragma unroll
        for (int m = 0; m < GB_MAXI; m++)
#pragma unroll
            for (int q = 0; q < 4; q++) acc[m][q] = 0.0;
        for (int kc = 0; kc < n_kc; kc++, it++) {
            const uint32_t s = it % GB_NS;
            mbar_wait(&full[s], (it / GB_NS) & 1);
            const uint8_t* sb = ring + (size_t)s * GB_STAGE;
            const int nsub = min(2, (nbk - kc * GEMM_KC) >> 3);
#pragma unroll
            for (int m = 0; m < GB_MAXI; m++) {
                const int u = warp + GB_WARPS * m;
                if (u < tot * KS) {
                    const int unit = u % tot, ks = u / tot;
                    const int rt = unit / NT, j = unit - rt * NT;
                    const uint8_t* qb = sb + rt * 2 * GB_QBOX + g * 128 + 4 * t;
                    const uint8_t* db = sb + GB_OFF_D + rt * GB_DBOX + g * (GEMM_KC * 2);
                    const uint8_t* xc = sb + GB_OFF_XQ + (8 * j + g) * GEMM_XCOL + 4 * t;
                    const float2* xdp = reinterpret_cast<const float2*>(sb + GB_OFF_XD) + 4 * j + t;     // [b][32] floats: pair (c0, c0+1), c0 = 8j + 2t
                    const int2* xsp = reinterpret_cast<const int2*>(sb + GB_OFF_XS) + 4 * j + t;
                    const int sub_begin = KS == 2 ? ks : 0, sub_end = KS == 2 ? min(nsub, ks + 1) : nsub;
                    for (int sub = sub_begin; sub < sub_end; sub++) {
                        const uint4 dv0 = *reinterpret_cast<const uint4*>(db + sub * 16);
                        const uint4 dv1 = *reinterpret_cast<const uint4*>(db + 8 * (GEMM_KC * 2) + sub * 16);
                        const uint32_t dw0p[4] = {dv0.x, dv0.y, dv0.z, dv0.w}, dw1p[4] = {dv1.x, dv1.y, dv1.z, dv1.w};
#pragma unroll
                        for (int b8 = 0; b8 < 8; b8++) {
                            const int b = sub * 8 + b8;
                            const uint32_t w0 = *reinterpret_cast<const uint32_t*>(qb + sub * GB_QBOX + ((b8 ^ g) << 4));
                            const uint32_t w1 = *reinterpret_cast<const uint32_t*>(qb + sub * GB_QBOX + 1024 + ((b8 ^ g) << 4));
                            const uint32_t a0 = w0 & 0x0F0F0F0Fu, a2 = (w0 >> 4) & 0x0F0F0F0Fu, a1 = w1 & 0x0F0F0F0Fu, a3 = (w1 >> 4) & 0x0F0F0F0Fu;
                            const uint32_t b0 = *reinterpret_cast<const uint32_t*>(xc + b * 16), b1 = *reinterpret_cast<const uint32_t*>(xc + 256 + b * 16);
                            const float2 da = xdp[b * (BATCH_MAX / 2)];
                            const int2 ss = xsp[b * (BATCH_MAX / 2)];          // -8 * sum(q): the MMA starts from it
                            int c[4];
                            mma_s8_16x8x32(c, a0, a1, a2, a3, b0, b1, ss.x, ss.y);
                            const float dw0 = half_of(dw0p[b8 >> 1], b8 & 1), dw1 = half_of(dw1p[b8 >> 1], b8 & 1);
                            acc[m][0] += (double)__fmul_rn(__fmul_rn((float)c[0], dw0), da.x);
                            acc[m][1] += (double)__fmul_rn(__fmul_rn((float)c[1], dw0), da.y);
                            acc[m][2] += (double)__fmul_rn(__fmul_rn((float)c[2], dw1), da.x);
                            acc[m][3] += (double)__fmul_rn(__fmul_rn((float)c[3], dw1), da.y);
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
        }
        // results of the row group -> shared memory -> the GEMV's epilogues with each column's operands
        double* outd = reinterpret_cast<double*>(outb);        // KS == 2: [ks][unit][4][32] partial sums
        if (KS == 2) {
            if (warp < tot * 2) {
#pragma unroll
                for (int q = 0; q < 4; q++) outd[(warp * 4 + q) * 32 + lane] = acc[0][q];      // u = warp: ks = warp / tot, unit = warp % tot
            }
        } else {
#pragma unroll
            for (int m = 0; m < GB_MAXI; m++) {
                const int unit = warp + GB_WARPS * m;
                if (unit < tot) {
                    const int rt = unit / NT, j = unit - rt * NT;
                    if (partial) {          // [row][32] doubles (the host keeps such matrices to <= 5 row tiles per CTA)
                        double* o = outd + (rt * 16 + g) * BATCH_MAX + 8 * j + 2 * t;
                        o[0] = acc[m][0]; o[1] = acc[m][1];
                        o[8 * BATCH_MAX] = acc[m][2]; o[8 * BATCH_MAX + 1] = acc[m][3];
                    } else {
                        float* o = outb + (rt * 16 + g) * GB_OUT_STRIDE + 8 * j + 2 * t;
                        o[0] = (float)acc[m][0]; o[1] = (float)acc[m][1];
                        o[8 * GB_OUT_STRIDE] = (float)acc[m][2]; o[8 * GB_OUT_STRIDE + 1] = (float)acc[m][3];
                    }
                }
            }
        }
        named_bar_sync(1, GB_NTC);
        const int npairs = n_rt * 8;
        for (int i = tid; i < npairs * B; i += GB_NTC) {
            const int p = i / B, c = i - p * B;
            const int64_t r0 = (int64_t)gt0 * 16 + 2 * p;
            if (r0 < rows) {
                float v0 = 0.f, v1 = 0.f;
                double s0 = 0.0, s1 = 0.0;
                if (KS == 2) {
                    // rows 2p, 2p+1 of the group, column c: unit, lane and register of the MMA accumulator layout
                    const int rl = 2 * p, unit = (rl >> 4) * NT + (c >> 3), ln = (rl & 7) * 4 + ((c & 7) >> 1), q = ((rl & 8) ? 2 : 0) + (c & 1);
                    const double* p0 = outd + (unit * 4 + q) * 32 + ln;
                    const double* p1 = outd + ((tot + unit) * 4 + q) * 32 + ln;
                    s0 = p0[0] + p1[0];
                    s1 = p0[4] + p1[4];                     // row + 1: the next lane group (g + 1), same column
                    v0 = (float)s0; v1 = (float)s1;
                } else if (partial) {
                    s0 = outd[(2 * p) * BATCH_MAX + c];
                    s1 = outd[(2 * p + 1) * BATCH_MAX + c];
                } else {
                    v0 = outb[(2 * p) * GB_OUT_STRIDE + c];
                    v1 = outb[(2 * p + 1) * GB_OUT_STRIDE + c];
                }
                const GemvArgs& a = colargs[c];
                if (partial) {
                    a.y64[a.epi_row0 + r0] = s0;
                    a.y64[a.epi_row0 + r0 + 1] = s1;
                } else {
                    unsigned long long best = 0ull;
                    epilogue_pair(a, a.epi_row0 + r0, v0, v1, best, col_pos[c], col_pt + (size_t)c * pt_stride);
                    if (a.epi == EPI_LOGITS && best != 0ull) atomicMax(a.argmax_key, best);
                }
            }
        }
        named_bar_sync(1, GB_NTC);
    }
}

template <int NT, bool SPLIT>
int gemm_launch_t(const GemmMaps& maps, const Q8Cols& x, int rows, int nbk, int B, const GemvArgs* colargs, const int* pos, const int* pt, int pt_stride, int grid,
                  cudaStream_t st) {
    auto kern = gemm_w4a8_kernel<NT, SPLIT>;
    static PerDeviceSmem configured;
    if (configured.raise(GB_SMEM)) B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GB_SMEM));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3((GB_WARPS + 1) * 32);
    cfg.dynamicSmemBytes = GB_SMEM;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = batch_pdl_enabled() ? 1 : 0;
    B2_CUDA(cudaLaunchKernelEx(&cfg, kern, maps, x, rows, nbk, B, colargs, pos, pt, pt_stride));
    return B2_OK;
}

PFN_cuTensorMapEncodeTiled encode_fn() {
    static PFN_cuTensorMapEncodeTiled fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled>(p);
    }
    return fn;
}
}  // namespace

bool gemm_supported(const QMat& W) { return W.type == T_Q4_0 && W.k % 256 == 0 && (W.rows & 1) == 0 && W.stride[0] % 16 == 0 && W.stride[1] % 16 == 0; }

int gemm_make_maps(const QMat& W, GemmMaps* out) {
    if (!gemm_supported(W)) B2_FAIL(B2_ERR_UNSUPPORTED, "gemm: type %d shape %lld x %lld has no batched kernel", W.type, (long long)W.rows, (long long)W.k);
    PFN_cuTensorMapEncodeTiled enc = encode_fn();
    if (!enc) B2_FAIL(B2_ERR_CUDA, "gemm: cuTensorMapEncodeTiled is not available from this driver");
    const cuuint32_t estr[2] = {1, 1};
    {
        const cuuint64_t dim[2] = {(cuuint64_t)(W.k / 2), (cuuint64_t)W.rows}, str[1] = {(cuuint64_t)W.stride[0]};
        const cuuint32_t box[2] = {128, 16};
        const CUresult r = enc(&out->q, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, W.plane[0], dim, str, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) B2_FAIL(B2_ERR_CUDA, "gemm: tensor map (nibbles) rejected, CUresult %d", (int)r);
    }
    {
        const cuuint64_t dim[2] = {(cuuint64_t)(W.k / 16), (cuuint64_t)W.rows}, str[1] = {(cuuint64_t)W.stride[1]};
        const cuuint32_t box[2] = {GEMM_KC * 2, 16};
        const CUresult r = enc(&out->d, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, W.plane[1], dim, str, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) B2_FAIL(B2_ERR_CUDA, "gemm: tensor map (scales) rejected, CUresult %d", (int)r);
    }
    return B2_OK;
}

int gemm_w4a8_launch(const QMat& W, const GemmMaps& maps, const Q8Cols& x, int B, const GemvArgs* d_colargs, const int* d_pos, const int* d_pt, int pt_stride,
                     cudaStream_t st, bool partial_out) {
    if (!gemm_supported(W)) B2_FAIL(B2_ERR_UNSUPPORTED, "gemm: type %d shape %lld x %lld has no batched kernel", W.type, (long long)W.rows, (long long)W.k);
    if (B < 1 || B > BATCH_MAX) B2_FAIL(B2_ERR_ARG, "gemm: %d columns (1..%d)", B, BATCH_MAX);
    if (W.k > x.k_cap) B2_FAIL(B2_ERR_ARG, "gemm: K=%lld beyond the activation buffer (%lld)", (long long)W.k, (long long)x.k_cap);
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int T = (int)((W.rows + 15) / 16);
    const int grid = std::min(T, sms > 0 ? sms : 148);
    const int nt = (B + 7) / 8, rows = (int)W.rows, nbk = (int)(W.k / 32);
    const int max_rt = (T + grid - 1) / grid;                 // row tiles of the busiest CTA
    static_assert(5 * 16 * BATCH_MAX * 8 <= GB_OUT_BYTES, "f64 output staging of a K-split matrix");
    if (partial_out && max_rt > 5) B2_FAIL(B2_ERR_UNSUPPORTED, "gemm: a K-split matrix with %d rows leaves %d row tiles per CTA (at most 5)", rows, max_rt);
    const bool split = max_rt <= GB_RTMAX && 2 * max_rt * nt <= GB_WARPS;
#define B2_GEMM_CASE(N) \
    case N: return split ? gemm_launch_t<N, true>(maps, x, rows, nbk, B, d_colargs, d_pos, d_pt, pt_stride, grid, st) \
                         : gemm_launch_t<N, false>(maps, x, rows, nbk, B, d_colargs, d_pos, d_pt, pt_stride, grid, st)
    switch (nt) {
        B2_GEMM_CASE(1);
        B2_GEMM_CASE(2);
        B2_GEMM_CASE(3);
    default: break;
    }
    return split ? gemm_launch_t<4, true>(maps, x, rows, nbk, B, d_colargs, d_pos, d_pt, pt_stride, grid, st)
                 : gemm_launch_t<4, false>(maps, x, rows, nbk, B, d_colargs, d_pos, d_pt, pt_stride, grid, st);
#undef B2_GEMM_CASE
}

}  // namespace b2

// gemv.cu -- the decode hot op: y = epilogue( W . q8(prologue(x)) ) for one activation vector.
//
// Arithmetic follows ggml's CPU path (the llama.cpp code inside the ollama/ollama image the
// reference deploys, reference pkg/model/pod.go:14-20): the activation vector is quantised to
// Q8_0 (Q4_0/Q8_0 weights) or Q8_K (Q4_K/Q6_K weights), every block dot is an exact integer
// (dp4a), and only the per-block scale-accumulate is floating point.  Restated in
// oracle/ggml_ref.c (ref_vec_dot_*), which the -m gpu tests compare against.
//
// Shape of the kernel (HBM-bound; no tensor cores on purpose):
//   * one CTA streams a contiguous range of rows; a producer warp issues 1-D bulk async
//     copies (cp.async.bulk + mbarrier complete_tx) of whole row tiles, one copy per weight
//     plane, into an NS-deep shared-memory ring with an L2 evict-first policy (every weight
//     byte is read exactly once per step);
//   * the ring is primed BEFORE griddepcontrol.wait, so under programmatic dependent launch
//     the next layer's weights are already in flight while the previous kernel drains;
//   * consumer warps first run the fused prologue (RMSNorm / LayerNorm + Q8 quantisation of
//     x into shared memory), cache their slice of the quantised x in registers for the whole
//     kernel, then take rows from the ring: LDS.128 of packed nibbles, dp4a, one shuffle
//     reduction per row;
//   * the epilogue (residual add, SiLU-gate, RoPE + f16 KV append, logits + arg-max) runs on
//     the finished row sums, so no intermediate vector makes a round trip through HBM.
#include <math.h>

#include "common.cuh"
#include "formats.cuh"
#include "kernels.h"
#include "gemv_dev.cuh"

namespace b2 {

namespace {

using namespace gemvdev;

constexpr int MAX_NS = 8;

struct GemvCfg {
    int S;          // warps along K per row group
    int G;          // row groups per CTA
    int RPG;        // rows per group per tile (even)
    int RT;         // rows per tile = G * RPG
    int NS;         // ring depth
    int n_tiles;
    int nunits;     // units per row (Q4_0/Q8_0: 32-weight blocks; K-quants: 128-weight half super-blocks)
    uint32_t stage_bytes;
    uint32_t off_red, off_xq, off_xd, off_xs, off_stage;   // shared-memory carve-up (bytes)
};

// ---------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------
template <int TYPE, int NB>
__global__ void __launch_bounds__((TYPE == T_Q4_K || TYPE == T_Q6_K) ? 288 : 544) gemv_kernel(const GemvArgs a, const GemvCfg c) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    uint64_t* empty = full + MAX_NS;
    double* dscratch = reinterpret_cast<double*>(smem + 2 * MAX_NS * 8);     // 32 doubles
    double* red = reinterpret_cast<double*>(smem + c.off_red);
    uint8_t* xq = smem + c.off_xq;
    float* xd = reinterpret_cast<float*>(smem + c.off_xd);
    int* xs = reinterpret_cast<int*>(smem + c.off_xs);
    uint8_t* stages = smem + c.off_stage;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_cons_warps = c.G * c.S, ntc = n_cons_warps * 32;
    constexpr int NP = TYPE == T_Q4_0 || TYPE == T_Q8_0 ? 2 : TYPE == T_Q4_K ? 3 : 4;

    // Rows owned by this CTA: an even split of the row PAIRS over the grid (pairs, because the
    // RoPE / SiLU-gate epilogues consume two adjacent rows), walked in tiles of RT rows.
    const int64_t pairs = a.W.rows >> 1;
    const int64_t r_begin = 2 * ((int64_t)blockIdx.x * pairs / gridDim.x), r_end = 2 * ((int64_t)(blockIdx.x + 1) * pairs / gridDim.x);
    const int my_tiles = (int)((r_end - r_begin + c.RT - 1) / c.RT);

    if (tid == 0) {
        for (int s = 0; s < c.NS; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], n_cons_warps); }
        mbar_fence_init();
    }
    __syncthreads();
    pdl_trigger();   // the next kernel in the stream may start priming its own ring

    if (warp == n_cons_warps) {
        // ===== producer: stream this CTA's row tiles into the ring =====
        if (lane == 0) {
            const uint64_t pol = l2_policy_evict_first();
            for (int it = 0; it < my_tiles; it++) {
                const int s = it % c.NS;
                if (it >= c.NS) mbar_wait(&empty[s], ((it / c.NS) - 1) & 1);
                const int64_t row0 = r_begin + (int64_t)it * c.RT;
                const int rows_here = (int)min((int64_t)c.RT, r_end - row0);
                uint32_t total = 0;
#pragma unroll
                for (int p = 0; p < NP; p++) total += (uint32_t)rows_here * a.W.stride[p];
                mbar_expect_tx(&full[s], total);
                uint8_t* dst = stages + (size_t)s * c.stage_bytes;
#pragma unroll
                for (int p = 0; p < NP; p++) {
                    bulk_g2s(dst, a.W.plane[p] + (size_t)row0 * a.W.stride[p], (uint32_t)rows_here * a.W.stride[p], &full[s], pol);
                    dst += (size_t)c.RT * a.W.stride[p];
                }
            }
        }
        return;
    }

    // ===== consumers =====
    pdl_wait();      // everything below reads activations written by the previous kernel
    {
        const ProStats st = prologue_stats(a, dscratch, tid, ntc);
        if (TYPE == T_Q4_0 || TYPE == T_Q8_0) prologue_q8_0(a, st, xq, xd, xs, tid, ntc);
        else prologue_q8_K(a, st, xq, xd, xs, tid, ntc);
        named_bar_sync(1, ntc);
    }

    const int g = warp / c.S, ch = warp % c.S;
    XRegs<TYPE> xr[NB];
    int unit[NB];
#pragma unroll
    for (int j = 0; j < NB; j++) {
        const int u = (ch * NB + j) * 32 + lane;
        const bool valid = u < c.nunits;
        unit[j] = valid ? u : c.nunits - 1;      // clamped: reads stay inside the row, x is zero
        xr[j].load(xq, xd, xs, unit[j], c.nunits, valid);
    }

    unsigned long long best = 0ull;
    for (int it = 0; it < my_tiles; it++) {
        const int s = it % c.NS;
        mbar_wait(&full[s], (it / c.NS) & 1);
        const int64_t row0 = r_begin + (int64_t)it * c.RT;
        const int rows_here = (int)min((int64_t)c.RT, r_end - row0);
        const uint8_t* sp = stages + (size_t)s * c.stage_bytes;
        double* redt = red + (size_t)(it & 1) * c.RT * c.S;
        for (int rr = 0; rr < c.RPG; rr++) {
            const int r = g * c.RPG + rr;
            if (r >= rows_here) break;
            RowPtrs rp;
            const uint8_t* pb = sp;
#pragma unroll
            for (int p = 0; p < NP; p++) { rp.p[p] = pb + (size_t)r * a.W.stride[p]; pb += (size_t)c.RT * a.W.stride[p]; }
            // per-unit float terms exactly as in ggml's generic path, added in f64: the sum no longer
            // depends on the order, so any split over lanes / warps gives the same f32 result
            double acc = 0.0;
#pragma unroll
            for (int j = 0; j < NB; j++) acc += (double)unit_dot(xr[j], rp, unit[j], c.nunits);
            acc = warp_sum_d(acc);
            if (lane == 0) redt[r * c.S + ch] = acc;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);           // this warp no longer reads stage s
        if (c.S > 1) named_bar_sync(2 + g, c.S * 32); else __syncwarp();
        if (ch == 0) {
            const int r = g * c.RPG + 2 * lane;
            if (2 * lane < c.RPG && r < rows_here) {
                double d0 = 0.0, d1 = 0.0;
                for (int k = 0; k < c.S; k++) { d0 += redt[r * c.S + k]; d1 += redt[(r + 1) * c.S + k]; }
                if (a.epi == EPI_PARTIAL) {          // tensor-parallel K slice: the f64 partial sums go to the all-reduce unrounded
                    a.y64[a.epi_row0 + row0 + r] = d0;
                    a.y64[a.epi_row0 + row0 + r + 1] = d1;
                } else {
                    epilogue_pair(a, a.epi_row0 + row0 + r, (float)d0, (float)d1, best);
                }
            }
        }
    }
    if (a.epi == EPI_LOGITS && ch == 0) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
        if (lane == 0 && best != 0ull) atomicMax(a.argmax_key, best);
    }
}

// ---------------------------------------------------------------------------------------------
// host side: configuration and launch
// ---------------------------------------------------------------------------------------------
int g_sm_count = 0;
int sm_count() {
    if (!g_sm_count) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (g_sm_count <= 0) g_sm_count = 148;
    }
    return g_sm_count;
}

template <int TYPE, int NB>
int launch_t(const GemvArgs& a, const GemvCfg& c, int grid, int threads, size_t smem, cudaStream_t st, bool pdl) {
    auto kern = gemv_kernel<TYPE, NB>;
    static PerDeviceSmem configured;     // per instantiation
    if (configured.raise(smem)) B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    B2_CUDA(cudaLaunchKernelEx(&cfg, kern, a, c));
    return B2_OK;
}

}  // namespace

size_t gemv_weight_bytes(const QMat& W) { return (size_t)W.rows * type_row_bytes(W.type, W.k); }

int gemv_launch(const GemvArgs& a, cudaStream_t st, bool pdl, const GemvTune* tune) {
    const QMat& W = a.W;
    if (!type_is_quant(W.type)) B2_FAIL(B2_ERR_UNSUPPORTED, "gemv: weight type %d not supported", W.type);
    if (W.rows <= 0 || W.k <= 0 || (W.rows & 1)) B2_FAIL(B2_ERR_ARG, "gemv: rows=%lld must be positive and even", (long long)W.rows);
    const bool kq = W.type == T_Q4_K || W.type == T_Q6_K;
    if (W.k % (kq ? 256 : 32)) B2_FAIL(B2_ERR_ARG, "gemv: k=%lld not a multiple of the block size", (long long)W.k);
    if (a.epi == EPI_QKV_ROPE && ((a.head_dim & 1) || (a.n_rot & 1) || (a.epi_row0 & 1) || a.epi_row0 + W.rows > a.n_q + 2 * a.n_kv))
        B2_FAIL(B2_ERR_ARG, "gemv: bad qkv geometry");

    GemvCfg c = {};
    c.nunits = (int)(kq ? W.k / 128 : W.k / 32);
    int nb = tune && tune->nb ? tune->nb : (kq ? 1 : (c.nunits >= 128 ? 4 : c.nunits >= 64 ? 2 : 1));
    if (kq && nb > 2) nb = 2;
    if (nb == 3) nb = 2;
    if (nb > 4) nb = 4;
    c.S = (c.nunits + 32 * nb - 1) / (32 * nb);
    if (c.S > 16) B2_FAIL(B2_ERR_UNSUPPORTED, "gemv: k=%lld too large", (long long)W.k);
    // defaults from the round-1 sweep on B200 (profiles/r01a_gemv_sweep.txt): 8 consumer warps per
    // CTA when a warp covers the whole row, two CTAs per SM; long rows take one bigger CTA per SM
    const int ctas_per_sm = tune && tune->ctas_per_sm ? tune->ctas_per_sm : (c.S == 1 ? 2 : 1);
    c.G = tune && tune->g ? tune->g : (c.S == 1 ? 8 : c.S <= 4 ? 4 : 2);
    const int max_cons = kq ? 8 : 16;      // bounded by the kernels' __launch_bounds__
    if (c.S > max_cons) B2_FAIL(B2_ERR_UNSUPPORTED, "gemv: k=%lld too large for type %d", (long long)W.k, W.type);
    if (c.G * c.S > max_cons) c.G = max(1, max_cons / c.S);
    c.RPG = tune && tune->rpg ? (tune->rpg + 1) / 2 * 2 : (c.S == 1 ? 2 : 4);
    size_t row_bytes = 0;
    for (int p = 0; p < W.n_planes; p++) row_bytes += W.stride[p];
    // x planes + per-block scales/sums
    const size_t xq_bytes = (size_t)W.k;
    const size_t xd_bytes = (kq ? W.k / 256 : W.k / 32) * 4;
    const size_t xs_bytes = (kq ? W.k / 16 : W.k / 32) * 4;
    int ns = tune && tune->ns ? tune->ns : (c.S == 1 ? 3 : 4);
    if (ns > MAX_NS) ns = MAX_NS;
    size_t off = 0;
    for (;;) {
        c.RT = c.G * c.RPG;
        c.stage_bytes = (uint32_t)((c.RT * row_bytes + 127) / 128 * 128);
        off = 2 * MAX_NS * 8 + 32 * 8;
        c.off_red = (uint32_t)off; off += (size_t)2 * c.RT * c.S * 8; off = (off + 15) / 16 * 16;
        c.off_xq = (uint32_t)off; off += xq_bytes; off = (off + 15) / 16 * 16;
        c.off_xd = (uint32_t)off; off += xd_bytes; off = (off + 15) / 16 * 16;
        c.off_xs = (uint32_t)off; off += xs_bytes; off = (off + 127) / 128 * 128;
        c.off_stage = (uint32_t)off;
        // very long rows: shrink the tile until at least two stages fit
        if (off + 2 * (size_t)c.stage_bytes <= 227 * 1024) break;
        if (c.RPG > 2) c.RPG = 2;
        else if (c.G > 1) c.G /= 2;
        else break;
    }
    const size_t budget = (size_t)(227 * 1024) / ctas_per_sm - 1024;   // 1 KB per CTA is reserved by the driver
    while (ns > 1 && off + (size_t)ns * c.stage_bytes > budget) ns--;
    if (ns < 2 && ctas_per_sm > 1) { ns = 2; while (ns > 1 && off + (size_t)ns * c.stage_bytes > 226 * 1024) ns--; }
    if (off + (size_t)ns * c.stage_bytes > 227 * 1024) B2_FAIL(B2_ERR_UNSUPPORTED, "gemv: k=%lld does not fit shared memory", (long long)W.k);
    c.NS = ns;
    const size_t smem = off + (size_t)ns * c.stage_bytes;
    c.n_tiles = (int)((W.rows + c.RT - 1) / c.RT);
    const int grid = (int)min((int64_t)(W.rows / 2), (int64_t)sm_count() * ctas_per_sm);
    const int threads = (c.G * c.S + 1) * 32;

#define B2_GEMV_CASE(T, N) if (W.type == T && nb == N) return launch_t<T, N>(a, c, grid, threads, smem, st, pdl)
    B2_GEMV_CASE(T_Q4_0, 1); B2_GEMV_CASE(T_Q4_0, 2); B2_GEMV_CASE(T_Q4_0, 4);
    B2_GEMV_CASE(T_Q8_0, 1); B2_GEMV_CASE(T_Q8_0, 2); B2_GEMV_CASE(T_Q8_0, 4);
    B2_GEMV_CASE(T_Q4_K, 1); B2_GEMV_CASE(T_Q4_K, 2);
    B2_GEMV_CASE(T_Q6_K, 1); B2_GEMV_CASE(T_Q6_K, 2);
#undef B2_GEMV_CASE
    B2_FAIL(B2_ERR_UNSUPPORTED, "gemv: no kernel for type %d nb %d", W.type, nb);
}

}  // namespace b2

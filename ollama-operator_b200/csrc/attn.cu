// attn.cu -- decode attention for one new token over the paged f16 KV cache.
//
// Numerics follow llama.cpp's non-flash CPU path step by step (restated in the oracle's
// `attend`, canonical mode): K and V are f16; the query is rounded to f16; both dot products
// (q.K over the head dimension, P.V over the cached positions) are evaluated in the lane order of
// ggml_vec_dot_f16 on an 8-wide SIMD host -- 32 f32 FMA chains and a fixed reduction tree -- which
// a 32-lane warp reproduces bit for bit; scores and soft-max are f32 with ONE global max and ONE
// global sum (f64), exp rounded from f64; the normalised probabilities are rounded to f16 before
// the V product.  Staying this close matters: the next op quantises the attention output to
// Q8_0, and a 1e-4 relative difference here already flips enough int8 roundings to move the
// logits by 1e-1 (measured with an online-soft-max variant that kept f32 probabilities).
//
// Parallelisation (B200: 148 SMs, thread-block clusters with distributed shared memory):
// one CLUSTER per (sequence slot, KV head), serving all query heads that share the KV head
// (GQA).  Scores: positions split over the cluster's CTAs, one warp per position, all-gathered
// through DSMEM (one exchange).  Output: head dimension split over the CTAs, no merge.  No global
// scratch, no second kernel.  HBM traffic is K and V once: (pos+1) * kv_dim * 2 * 2 bytes.
#include <math.h>
#include <stdlib.h>

#include "attn_dev.cuh"
#include "common.cuh"
#include "kernels.h"

namespace b2 {

namespace {

using namespace attndev;

constexpr int ATT_THREADS = 256;

template <int GRP, int HD>
__global__ void __launch_bounds__(ATT_THREADS) attn_decode_kernel(const AttnArgs a, const AttnCfg c) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t xbar;
    if (threadIdx.x == 0) {
        mbar_init(&xbar, c.cluster);
        mbar_fence_init();
    }
    cluster_sync_all();          // every CTA's exchange barrier exists before a peer can arrive on it
    pdl_trigger();
    pdl_wait();                  // q and this step's K/V rows come from the preceding kernel
    uint32_t xphase = 0;
    attn_kv_head<GRP, HD, ATT_THREADS>(a, c, blockIdx.y, smem_raw, threadIdx.x, &xbar, xphase, 1);
}

// Batched variant: blockIdx.z selects the column (a sequence, or a prompt token of a chunk); each
// column brings its own query, output, page table and position through a device-resident array.
template <int GRP, int HD>
__global__ void __launch_bounds__(ATT_THREADS) attn_decode_batch_kernel(const AttnArgs* __restrict__ args, const AttnCfg c) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t xbar;
    if (threadIdx.x == 0) {
        mbar_init(&xbar, c.cluster);
        mbar_fence_init();
    }
    cluster_sync_all();
    pdl_trigger();
    pdl_wait();                  // q and this step's K/V rows come from the preceding kernel
    const AttnArgs a = args[blockIdx.z];
    uint32_t xphase = 0;
    attn_kv_head<GRP, HD, ATT_THREADS>(a, c, blockIdx.y, smem_raw, threadIdx.x, &xbar, xphase, 1);
}

template <int GRP, int HD>
int launch_grp_batch(const AttnArgs* d_args, int n_head_kv, int B, const AttnCfg& c, size_t smem, cudaStream_t st) {
    auto kern = attn_decode_batch_kernel<GRP, HD>;
    static PerDeviceSmem configured;
    if (configured.raise(smem)) B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(c.cluster, n_head_kv, B);
    cfg.blockDim = dim3(ATT_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = c.cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 2;
    B2_CUDA(cudaLaunchKernelEx(&cfg, kern, d_args, c));
    return B2_OK;
}

// ---- batched steps with many (column, head) pairs: one CTA per pair, K and V staged through shared memory -----
// The cluster kernel above is organised around one sequence's latency; a batched step is bandwidth work (every
// column brings its own KV rows: 268 MB per 7B layer at 32 x 512 tokens).  Here the rows of a 64-position chunk
// are copied with cp.async (no registers held while in flight, the next chunk streams while this one is used),
// and the arithmetic is attn_kv_head's, primitive for primitive and in the same order: q.K in the 32-lane SIMD
// order through reduce8_lanes32 (tail dims in f64), one max, exp from f64, f64 sum, f16 probabilities, P.V as one
// f32 FMA chain per (position mod 32, dim) in position order.  Same bits.
constexpr int ABT = 64;                       // positions per chunk

template <int HD>
__global__ void __launch_bounds__(256) attn_batch_staged_kernel(const AttnArgs* __restrict__ args, const int maxc) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    constexpr int NT = 256, NWARP = 8, HD32 = HD & ~31, NI = HD32 / 32;
    constexpr int ROWB = HD * 2 + 16;          // padded row: 16-byte reads of 32 different rows spread over the banks
    constexpr int PIECES = HD * 2 / 16;        // 16-byte pieces per row
    pdl_trigger();
    pdl_wait();                                // q and this step's K/V rows come from the preceding kernel
    const AttnArgs a = args[blockIdx.y];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = *a.pos + 1, h = blockIdx.x, kvh = h / (a.n_head / a.n_head_kv);
    int* rowoff = reinterpret_cast<int*>(smem_raw);                      // [maxc]
    float* sc = reinterpret_cast<float*>(smem_raw) + maxc;               // [maxc]
    float* qs = sc + maxc;                                               // [HD]
    float* red = qs + HD;                                                // [NWARP] floats, then [NWARP] doubles (16-byte aligned below)
    double* redd = reinterpret_cast<double*>(red + NWARP);
    uint8_t* tile = reinterpret_cast<uint8_t*>(redd + NWARP);            // [2][ABT][ROWB]
    const int pshift = __ffs(a.page_tokens) - 1, pmask = a.page_tokens - 1;
    for (int t = tid; t < n; t += NT) rowoff[t] = ((a.page_table[t >> pshift] << pshift) + (t & pmask)) * a.kv_dim;
    for (int i = tid; i < HD; i += NT) qs[i] = __half2float(__float2half_rn(__ldcg(a.q + (size_t)h * HD + i)));
    __syncthreads();
    const int n_chunks = (n + ABT - 1) / ABT;
    auto stage = [&](const __half* base, int c) {          // rows [c*ABT, c*ABT + ABT) of this head -> tile[c & 1]
        uint8_t* dst = tile + (size_t)(c & 1) * ABT * ROWB;
        for (int i = tid; i < ABT * PIECES; i += NT) {
            const int r = i / PIECES, pc = i - r * PIECES, t = min(c * ABT + r, n - 1);
            cp_async16(dst + r * ROWB + pc * 16, base + rowoff[t] + pc * 8);
        }
        cp_async_commit();
    };

    // ---- 1. scores ----
    {
        const __half* const k_base = a.k_base + (size_t)kvh * HD;
        float qreg[NI];
#pragma unroll
        for (int i = 0; i < NI; i++) qreg[i] = qs[i * 32 + lane];
        stage(k_base, 0);
        for (int c = 0; c < n_chunks; c++) {
            if (c + 1 < n_chunks) { stage(k_base, c + 1); cp_async_wait<1>(); } else cp_async_wait<0>();
            __syncthreads();
            const uint8_t* tl = tile + (size_t)(c & 1) * ABT * ROWB;
            const int r0 = warp * 8, tb = c * ABT + r0;          // eight positions per warp: one reduce8 batch
            if (tb < n) {
                float prod[8];
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    const __half* kr = reinterpret_cast<const __half*>(tl + (r0 + u) * ROWB);
                    float acc = 0.f;
#pragma unroll
                    for (int i = 0; i < NI; i++) acc = fmaf(__half2float(kr[i * 32 + lane]), qreg[i], acc);
                    prod[u] = acc;
                }
                float r = reduce8_lanes32(prod, lane);
                const int u = reduce8_index(lane), t = tb + u;
                if (t < n) {
                    if (HD32 != HD) {               // ggml adds the tail beyond a multiple of 32 in f64
                        const __half* kr = reinterpret_cast<const __half*>(tl + (r0 + u) * ROWB);
                        double sumf = (double)r;
                        for (int d = HD32; d < HD; d++) sumf += (double)__fmul_rn(__half2float(kr[d]), qs[d]);
                        r = (float)sumf;
                    }
                    if ((lane & 3) == 0) sc[t] = __fmul_rn(r, a.scale);
                }
            }
            __syncthreads();          // the chunk buffer is free again; after the last chunk: all scores are in place
        }
    }
    // ---- 2. soft-max (ggml soft_max: one max, exp, f64 sum, scale by 1/sum; probabilities to f16) ----
    {
        float lmax = -INFINITY;
        for (int t = tid; t < n; t += NT) lmax = fmaxf(lmax, sc[t]);
        lmax = warp_max(lmax);
        if (lane == 0) red[warp] = lmax;
        __syncthreads();
        float M = red[0];
#pragma unroll
        for (int w = 1; w < NWARP; w++) M = fmaxf(M, red[w]);
        double lsum = 0.0;
        for (int t = tid; t < n; t += NT) {
            const float e = (float)exp((double)__fsub_rn(sc[t], M));
            sc[t] = e;
            lsum += (double)e;
        }
        lsum = warp_sum_d(lsum);
        if (lane == 0) redd[warp] = lsum;
        __syncthreads();
        double S = 0;
#pragma unroll
        for (int w = 0; w < NWARP; w++) S += redd[w];
        const float inv = (float)(1.0 / S);
        for (int t = tid; t < n; t += NT) sc[t] = __half2float(__float2half_rn(__fmul_rn(sc[t], inv)));
        __syncthreads();
    }
    // ---- 3. P.V: warp w owns dims [16w, 16w+16); lane L owns the positions L mod 32, one FMA chain per dim ----
    {
        const __half* const v_base = a.v_base + (size_t)kvh * HD;
        const int d0 = 16 * warp;
        const bool active = d0 < HD;
        float acc[2][8];
#pragma unroll
        for (int j = 0; j < 8; j++) acc[0][j] = acc[1][j] = 0.f;
        stage(v_base, 0);
        for (int c = 0; c < n_chunks; c++) {
            if (c + 1 < n_chunks) { stage(v_base, c + 1); cp_async_wait<1>(); } else cp_async_wait<0>();
            __syncthreads();
            if (active) {
                const uint8_t* tl = tile + (size_t)(c & 1) * ABT * ROWB + d0 * 2;
#pragma unroll
                for (int u = 0; u < ABT / 32; u++) {
                    const int r = lane + 32 * u, t = c * ABT + r;
                    if (t < n) {
                        const float p = sc[t];
                        const uint4 v0 = *reinterpret_cast<const uint4*>(tl + r * ROWB), v1 = *reinterpret_cast<const uint4*>(tl + r * ROWB + 16);
                        const __half2* h0 = reinterpret_cast<const __half2*>(&v0);
                        const __half2* h1 = reinterpret_cast<const __half2*>(&v1);
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            const float2 f0 = __half22float2(h0[j]), f1 = __half22float2(h1[j]);
                            acc[0][2 * j] = fmaf(f0.x, p, acc[0][2 * j]); acc[0][2 * j + 1] = fmaf(f0.y, p, acc[0][2 * j + 1]);
                            acc[1][2 * j] = fmaf(f1.x, p, acc[1][2 * j]); acc[1][2 * j + 1] = fmaf(f1.y, p, acc[1][2 * j + 1]);
                        }
                    }
                }
            }
            __syncthreads();
        }
        if (active) {
            float* outh = a.out + (size_t)h * HD + d0;
#pragma unroll
            for (int half = 0; half < 2; half++) {
                const float r = reduce8_lanes32(acc[half], lane);
                if ((lane & 3) == 0) outh[8 * half + reduce8_index(lane)] = r;
            }
        }
    }
}

template <int HD>
int launch_batch_staged(const AttnArgs* d_args, const AttnArgs& a, int B, cudaStream_t st) {
    const int maxc = (a.max_ctx + 3) / 4 * 4;
    const size_t smem = (size_t)2 * maxc * 4 + (size_t)HD * 4 + 8 * 4 + 8 * 8 + (size_t)2 * ABT * (HD * 2 + 16);
    auto kern = attn_batch_staged_kernel<HD>;
    static PerDeviceSmem configured;
    if (smem > 48 * 1024 && configured.raise(smem)) B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(a.n_head, B, 1);
    cfg.blockDim = dim3(256);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = batch_pdl_enabled() ? 1 : 0;
    B2_CUDA(cudaLaunchKernelEx(&cfg, kern, d_args, maxc));
    return B2_OK;
}

template <int GRP, int HD>
int launch_grp(const AttnArgs& a, const AttnCfg& c, size_t smem, cudaStream_t st, bool pdl) {
    auto kern = attn_decode_kernel<GRP, HD>;
    static PerDeviceSmem configured;
    if (configured.raise(smem)) B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(c.cluster, a.n_head_kv, 1);
    cfg.blockDim = dim3(ATT_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = c.cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 2 : 1;
    B2_CUDA(cudaLaunchKernelEx(&cfg, kern, a, c));
    return B2_OK;
}

}  // namespace

int attn_decode_launch(const AttnArgs& a, cudaStream_t st, bool pdl, int cluster) {
    const int max_ctx = a.max_ctx;
    if (max_ctx <= 0) B2_FAIL(B2_ERR_ARG, "attn: max_ctx must be positive");
    if (a.n_head_kv <= 0 || a.n_head % a.n_head_kv) B2_FAIL(B2_ERR_ARG, "attn: n_head %d not a multiple of n_head_kv %d", a.n_head, a.n_head_kv);
    const int grp = a.n_head / a.n_head_kv;
    AttnCfg c;
    if (cluster <= 0) {
        // enough CTAs to cover the machine at batch 1, at most 8 per cluster (portable limit)
        cluster = 1;
        while (cluster < MAX_CLUSTER && cluster * a.n_head_kv * 2 <= 160) cluster *= 2;
    }
    if (cluster > MAX_CLUSTER) cluster = MAX_CLUSTER;
    while (cluster > 1 && a.head_dim % (8 * cluster)) cluster /= 2;      // the output dims are split eight per warp over the CTAs
    c.cluster = cluster;
    c.max_local = (max_ctx + 3) / 4 * 4;
    const size_t smem = attn_smem_bytes(grp, a.head_dim, c.max_local, cluster, ATT_THREADS);
    if (smem > 200 * 1024) B2_FAIL(B2_ERR_UNSUPPORTED, "attn: context %d needs %zu B of shared memory", max_ctx, smem);
#define B2_ATT_CASE(G, H) if (grp == G && a.head_dim == H) return launch_grp<G, H>(a, c, smem, st, pdl)
    B2_ATT_CASE(1, 64); B2_ATT_CASE(1, 80); B2_ATT_CASE(1, 128);
    B2_ATT_CASE(2, 64); B2_ATT_CASE(2, 80); B2_ATT_CASE(2, 128);
    B2_ATT_CASE(4, 64); B2_ATT_CASE(4, 80); B2_ATT_CASE(4, 128);
    B2_ATT_CASE(8, 64); B2_ATT_CASE(8, 80); B2_ATT_CASE(8, 128);
#undef B2_ATT_CASE
    if (a.head_dim != 64 && a.head_dim != 80 && a.head_dim != 128) B2_FAIL(B2_ERR_UNSUPPORTED, "attn: head_dim %d not supported (64, 80, 128)", a.head_dim);
    B2_FAIL(B2_ERR_UNSUPPORTED, "attn: %d query heads per KV head not supported (1, 2, 4, 8)", grp);
}

int attn_decode_batch_launch(const AttnArgs* d_args, const AttnArgs& a, int B, cudaStream_t st) {
    if (a.max_ctx <= 0 || B < 1) B2_FAIL(B2_ERR_ARG, "attn: bad batch arguments");
    if (a.n_head_kv <= 0 || a.n_head % a.n_head_kv) B2_FAIL(B2_ERR_ARG, "attn: n_head %d not a multiple of n_head_kv %d", a.n_head, a.n_head_kv);
    // many (column, head) pairs: the bandwidth-organised kernel; few: the latency-organised cluster kernel
    const size_t staged_smem = (size_t)2 * ((a.max_ctx + 3) / 4 * 4) * 4 + (size_t)2 * ABT * (a.head_dim * 2 + 16) + 1024;
    if ((size_t)B * a.n_head >= 128 && a.head_dim % 16 == 0 && a.head_dim <= 128 && staged_smem <= 100 * 1024 && !getenv("B2_ATTN_CLUSTER_ONLY")) {
        if (a.head_dim == 64) return launch_batch_staged<64>(d_args, a, B, st);
        if (a.head_dim == 80) return launch_batch_staged<80>(d_args, a, B, st);
        if (a.head_dim == 128) return launch_batch_staged<128>(d_args, a, B, st);
    }
    const int grp = a.n_head / a.n_head_kv;
    AttnCfg c;
    int cluster = 1;      // the columns already spread the work: only small batches split a head over CTAs
    while (cluster < MAX_CLUSTER && cluster * a.n_head_kv * B * 2 <= 160) cluster *= 2;
    while (cluster > 1 && a.head_dim % (8 * cluster)) cluster /= 2;
    c.cluster = cluster;
    c.max_local = (a.max_ctx + 3) / 4 * 4;
    const size_t smem = attn_smem_bytes(grp, a.head_dim, c.max_local, cluster, ATT_THREADS);
    if (smem > 200 * 1024) B2_FAIL(B2_ERR_UNSUPPORTED, "attn: context %d needs %zu B of shared memory", a.max_ctx, smem);
#define B2_ATT_CASE(G, H) if (grp == G && a.head_dim == H) return launch_grp_batch<G, H>(d_args, a.n_head_kv, B, c, smem, st)
    B2_ATT_CASE(1, 64); B2_ATT_CASE(1, 80); B2_ATT_CASE(1, 128);
    B2_ATT_CASE(2, 64); B2_ATT_CASE(2, 80); B2_ATT_CASE(2, 128);
    B2_ATT_CASE(4, 64); B2_ATT_CASE(4, 80); B2_ATT_CASE(4, 128);
    B2_ATT_CASE(8, 64); B2_ATT_CASE(8, 80); B2_ATT_CASE(8, 128);
#undef B2_ATT_CASE
    B2_FAIL(B2_ERR_UNSUPPORTED, "attn: head_dim %d with %d query heads per KV head not supported", a.head_dim, grp);
}

}  // namespace b2

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

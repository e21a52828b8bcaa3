// attn.cu -- decode attention for one new token over the paged f16 KV cache.
//
// Numerics follow llama.cpp's non-flash CPU path (oracle/ggml_ref.c `attend`): K and V are f16,
// the query is rounded to f16 before the dot, products are accumulated in f32, soft-max in
// f32.  One deliberate difference: ggml rounds the normalised probabilities to f16 before the
// V product; a split-context kernel only knows the normaliser at the end, so probabilities
// stay f32 here.  The parity tests bound the effect (<= 1e-3 relative on the output).
//
// Parallelisation (B200: 148 SMs, thread-block clusters with distributed shared memory):
// one CLUSTER per (sequence slot, KV head); the cluster's CTAs split the context, each runs
// an exact local soft-max over its slice for all query heads that share the KV head (GQA), and
// the partial (max, sum, output) triples are merged through DSMEM in rank 0 -- no global
// scratch, no second kernel.  HBM traffic is K and V once: (pos+1) * kv_dim * 2 * 2 bytes.
#include <math.h>

#include "common.cuh"
#include "kernels.h"

namespace b2 {

namespace {

constexpr int ATT_THREADS = 128;
constexpr int MAX_CLUSTER = 8;

struct AttnCfg {
    int cluster;        // CTAs per (slot, kv head)
    int max_local;      // capacity of the per-CTA score buffer (positions)
};

template <int GRP>
__global__ void __launch_bounds__(ATT_THREADS) attn_decode_kernel(const AttnArgs a, const AttnCfg c) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const int HD = a.head_dim;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t rk = cluster_ctarank();
    const int kvh = blockIdx.y;

    // shared-memory carve-up
    float* qs = reinterpret_cast<float*>(smem_raw);                 // [GRP][HD]   f16-rounded query
    float* sc = qs + GRP * HD;                                      // [GRP][max_local] scores -> probabilities
    float* wred = sc + (size_t)GRP * c.max_local;                   // [4][GRP] cross-warp scratch
    float* ml = wred + 4 * GRP;                                     // [2][GRP] local max, local sum
    float* ored = ml + 2 * GRP;                                     // [npl][GRP][HD] partial outputs
    const int ndg = HD / 8, npl = ATT_THREADS / ndg;
    float* merge = ored + (size_t)npl * GRP * HD;                   // rank 0 only: [cluster][GRP][HD + 2]

    pdl_trigger();
    pdl_wait();      // q and this step's K/V rows come from the preceding kernel

    const int n = *a.pos + 1;
    const int per = (n + c.cluster - 1) / c.cluster;
    const int t_begin = min(n, (int)rk * per), t_end = min(n, t_begin + per);
    const int n_loc = t_end - t_begin;

    for (int i = tid; i < GRP * HD; i += ATT_THREADS)
        qs[i] = __half2float(__float2half_rn(a.q[(size_t)kvh * GRP * HD + i]));
    __syncthreads();

    // ---- phase 1: scores, one thread per cached position ------------------------------------
    float lmax[GRP];
#pragma unroll
    for (int g = 0; g < GRP; g++) lmax[g] = -INFINITY;
    for (int t = t_begin + tid; t < t_end; t += ATT_THREADS) {
        const int page = a.page_table[t / a.page_tokens];
        const uint4* kr = reinterpret_cast<const uint4*>(a.k_base + ((size_t)page * a.page_tokens + t % a.page_tokens) * a.kv_dim +
                                                         (size_t)kvh * HD);
        float s[GRP];
#pragma unroll
        for (int g = 0; g < GRP; g++) s[g] = 0.f;
        for (int ch = 0; ch < ndg; ch++) {
            const uint4 kv = kr[ch];
            const __half2* h2 = reinterpret_cast<const __half2*>(&kv);
            float kf[8];
#pragma unroll
            for (int i = 0; i < 4; i++) { const float2 f = __half22float2(h2[i]); kf[2 * i] = f.x; kf[2 * i + 1] = f.y; }
#pragma unroll
            for (int g = 0; g < GRP; g++) {
                const float4 q0 = *reinterpret_cast<const float4*>(qs + g * HD + ch * 8);
                const float4 q1 = *reinterpret_cast<const float4*>(qs + g * HD + ch * 8 + 4);
                s[g] = fmaf(kf[0], q0.x, s[g]); s[g] = fmaf(kf[1], q0.y, s[g]); s[g] = fmaf(kf[2], q0.z, s[g]); s[g] = fmaf(kf[3], q0.w, s[g]);
                s[g] = fmaf(kf[4], q1.x, s[g]); s[g] = fmaf(kf[5], q1.y, s[g]); s[g] = fmaf(kf[6], q1.z, s[g]); s[g] = fmaf(kf[7], q1.w, s[g]);
            }
        }
#pragma unroll
        for (int g = 0; g < GRP; g++) {
            const float v = s[g] * a.scale;
            sc[(size_t)g * c.max_local + (t - t_begin)] = v;
            lmax[g] = fmaxf(lmax[g], v);
        }
    }
#pragma unroll
    for (int g = 0; g < GRP; g++) {
        const float m = warp_max(lmax[g]);
        if (lane == 0) wred[warp * GRP + g] = m;
    }
    __syncthreads();
    if (tid < GRP) ml[tid] = fmaxf(fmaxf(wred[tid], wred[GRP + tid]), fmaxf(wred[2 * GRP + tid], wred[3 * GRP + tid]));
    __syncthreads();

    // ---- phase 2: p = exp(s - local max), local sum ---------------------------------------------
    float lsum[GRP];
#pragma unroll
    for (int g = 0; g < GRP; g++) lsum[g] = 0.f;
    for (int i = tid; i < n_loc; i += ATT_THREADS) {
#pragma unroll
        for (int g = 0; g < GRP; g++) {
            const float p = expf(sc[(size_t)g * c.max_local + i] - ml[g]);
            sc[(size_t)g * c.max_local + i] = p;
            lsum[g] += p;
        }
    }
#pragma unroll
    for (int g = 0; g < GRP; g++) {
        const float s = warp_sum(lsum[g]);
        if (lane == 0) wred[warp * GRP + g] = s;
    }
    __syncthreads();
    if (tid < GRP) ml[GRP + tid] = (wred[tid] + wred[GRP + tid]) + (wred[2 * GRP + tid] + wred[3 * GRP + tid]);
    __syncthreads();

    // ---- phase 3: o = sum_t p_t * V_t; thread = (8-wide dim group, position lane) ---------------
    {
        const int dg = tid % ndg, pl = tid / ndg;
        float acc[GRP][8];
#pragma unroll
        for (int g = 0; g < GRP; g++)
#pragma unroll
            for (int i = 0; i < 8; i++) acc[g][i] = 0.f;
        if (pl < npl) {
            for (int t = t_begin + pl; t < t_end; t += npl) {
                const int page = a.page_table[t / a.page_tokens];
                const uint4 vv = *reinterpret_cast<const uint4*>(a.v_base + ((size_t)page * a.page_tokens + t % a.page_tokens) * a.kv_dim +
                                                                 (size_t)kvh * HD + dg * 8);
                const __half2* h2 = reinterpret_cast<const __half2*>(&vv);
                float vf[8];
#pragma unroll
                for (int i = 0; i < 4; i++) { const float2 f = __half22float2(h2[i]); vf[2 * i] = f.x; vf[2 * i + 1] = f.y; }
#pragma unroll
                for (int g = 0; g < GRP; g++) {
                    const float p = sc[(size_t)g * c.max_local + (t - t_begin)];
#pragma unroll
                    for (int i = 0; i < 8; i++) acc[g][i] = fmaf(p, vf[i], acc[g][i]);
                }
            }
#pragma unroll
            for (int g = 0; g < GRP; g++)
#pragma unroll
                for (int i = 0; i < 8; i++) ored[((size_t)pl * GRP + g) * HD + dg * 8 + i] = acc[g][i];
        }
    }
    __syncthreads();

    // ---- ship (max, sum, output) of this slice to rank 0 through DSMEM ---------------------------
    const int rec = HD + 2;   // floats per (rank, head) record
    for (int i = tid; i < GRP * HD; i += ATT_THREADS) {
        const int g = i / HD, d = i % HD;
        float o = 0.f;
        for (int p = 0; p < npl; p++) o += ored[((size_t)p * GRP + g) * HD + d];
        dsmem_st_f32(dsmem_addr(merge + ((size_t)rk * GRP + g) * rec + d, 0), o);
    }
    if (tid < GRP) {
        dsmem_st_f32(dsmem_addr(merge + ((size_t)rk * GRP + tid) * rec + HD, 0), ml[tid]);
        dsmem_st_f32(dsmem_addr(merge + ((size_t)rk * GRP + tid) * rec + HD + 1, 0), ml[GRP + tid]);
    }
    cluster_sync_all();
    if (rk != 0) return;

    // ---- rank 0: log-sum-exp merge of the slices --------------------------------------------------
    for (int i = tid; i < GRP * HD; i += ATT_THREADS) {
        const int g = i / HD, d = i % HD;
        float M = -INFINITY;
        for (int r = 0; r < c.cluster; r++) M = fmaxf(M, merge[((size_t)r * GRP + g) * rec + HD]);
        float L = 0.f, O = 0.f;
        for (int r = 0; r < c.cluster; r++) {
            const float* m = merge + ((size_t)r * GRP + g) * rec;
            const float w = expf(m[HD] - M);        // empty slice: exp(-inf) = 0
            L = fmaf(m[HD + 1], w, L);
            O = fmaf(m[d], w, O);
        }
        a.out[((size_t)kvh * GRP + g) * HD + d] = O / L;
    }
}

size_t attn_smem_bytes(int grp, int hd, int max_local, int cluster) {
    const int ndg = hd / 8, npl = ATT_THREADS / ndg;
    size_t f = (size_t)grp * hd + (size_t)grp * max_local + 4 * grp + 2 * grp + (size_t)npl * grp * hd + (size_t)cluster * grp * (hd + 2);
    return f * sizeof(float);
}

template <int GRP>
int launch_grp(const AttnArgs& a, const AttnCfg& c, size_t smem, cudaStream_t st, bool pdl) {
    auto kern = attn_decode_kernel<GRP>;
    static size_t configured = 0;
    if (smem > configured) {
        B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
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
    if (a.head_dim % 8 || a.head_dim > 256 || a.head_dim < 8) B2_FAIL(B2_ERR_UNSUPPORTED, "attn: head_dim %d", a.head_dim);
    const int grp = a.n_head / a.n_head_kv;
    AttnCfg c;
    if (cluster <= 0) {
        // enough CTAs to cover the machine at batch 1, at most 8 per cluster (portable limit)
        cluster = 1;
        while (cluster < MAX_CLUSTER && cluster * a.n_head_kv * 2 <= 160) cluster *= 2;
    }
    if (cluster > MAX_CLUSTER) cluster = MAX_CLUSTER;
    c.cluster = cluster;
    c.max_local = ((max_ctx + cluster - 1) / cluster + 3) / 4 * 4;
    const size_t smem = attn_smem_bytes(grp, a.head_dim, c.max_local, cluster);
    if (smem > 200 * 1024) B2_FAIL(B2_ERR_UNSUPPORTED, "attn: context %d needs %zu B of shared memory", max_ctx, smem);
    switch (grp) {
    case 1: return launch_grp<1>(a, c, smem, st, pdl);
    case 2: return launch_grp<2>(a, c, smem, st, pdl);
    case 4: return launch_grp<4>(a, c, smem, st, pdl);
    case 8: return launch_grp<8>(a, c, smem, st, pdl);
    }
    B2_FAIL(B2_ERR_UNSUPPORTED, "attn: %d query heads per KV head not supported (1, 2, 4, 8)", grp);
}

}  // namespace b2

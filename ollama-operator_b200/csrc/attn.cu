// attn.cu -- decode attention for one new token over the paged f16 KV cache.
//
// Numerics follow llama.cpp's non-flash CPU path step by step (restated in the oracle's
// `attend`): K and V are f16; the query is rounded to f16 before the dot; scores and the
// soft-max are f32 with ONE global max and ONE global sum; the normalised probabilities are
// rounded to f16 before the V product; accumulation is f32.  Staying this close matters: the
// next op quantises the attention output to Q8_0, and a 1e-4 relative difference here already
// flips enough int8 roundings to move the logits by 1e-1 (measured; an online-soft-max variant
// that kept f32 probabilities failed the 2e-2 logit tolerance).
//
// Parallelisation (B200: 148 SMs, thread-block clusters with distributed shared memory):
// one CLUSTER per (sequence slot, KV head); the cluster's CTAs split the context and serve all
// query heads that share the KV head (GQA).  The global max and sum are formed by all-gathers
// of the per-CTA values through DSMEM, the partial outputs are summed in rank 0 -- no global
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
    const int ndg = HD / 4, npl = ATT_THREADS / ndg;     // phase 3: 4-wide dim groups x position lanes

    // shared-memory carve-up (doubles first: 8-byte alignment)
    double* ored = reinterpret_cast<double*>(smem_raw);             // [npl][GRP][HD] partial outputs
    double* merge = ored + (size_t)npl * GRP * HD;                  // rank 0 only: [cluster][GRP][HD]
    double* csum = merge + (size_t)c.cluster * GRP * HD;            // [MAX_CLUSTER][GRP] every CTA's sum (all-gathered)
    double* wsum = csum + MAX_CLUSTER * GRP;                        // [4][GRP] cross-warp scratch
    float* qs = reinterpret_cast<float*>(wsum + 4 * GRP);           // [GRP][HD]   f16-rounded query
    float* sc = qs + GRP * HD;                                      // [GRP][max_local] scores -> probabilities
    float* wred = sc + (size_t)GRP * c.max_local;                   // [4][GRP] cross-warp scratch
    float* ml = wred + 4 * GRP;                                     // [2][GRP] global max, 1/sum
    float* cmax = ml + 2 * GRP;                                     // [MAX_CLUSTER][GRP] every CTA's max (all-gathered)

    pdl_trigger();
    pdl_wait();      // q and this step's K/V rows come from the preceding kernel

    const int n = *a.pos + 1;
    const int per = (n + c.cluster - 1) / c.cluster;
    const int t_begin = min(n, (int)rk * per), t_end = min(n, t_begin + per);
    const int n_loc = t_end - t_begin;

    for (int i = tid; i < GRP * HD; i += ATT_THREADS)
        qs[i] = __half2float(__float2half_rn(a.q[(size_t)kvh * GRP * HD + i]));
    __syncthreads();

    // ---- phase 1: scores, one thread per cached position.  f16 x f16 products are exact in f32;
    //      they are added in f64 (as ggml's scalar f16 dot does), so the order is immaterial. ----------
    float lmax[GRP];
#pragma unroll
    for (int g = 0; g < GRP; g++) lmax[g] = -INFINITY;
    for (int t = t_begin + tid; t < t_end; t += ATT_THREADS) {
        const int page = a.page_table[t / a.page_tokens];
        const uint4* kr = reinterpret_cast<const uint4*>(a.k_base + ((size_t)page * a.page_tokens + t % a.page_tokens) * a.kv_dim +
                                                         (size_t)kvh * HD);
        double s[GRP];
#pragma unroll
        for (int g = 0; g < GRP; g++) s[g] = 0.0;
        for (int ch = 0; ch < HD / 8; ch++) {
            const uint4 kv = kr[ch];
            const __half2* h2 = reinterpret_cast<const __half2*>(&kv);
            float kf[8];
#pragma unroll
            for (int i = 0; i < 4; i++) { const float2 f = __half22float2(h2[i]); kf[2 * i] = f.x; kf[2 * i + 1] = f.y; }
#pragma unroll
            for (int g = 0; g < GRP; g++) {
                const float4 q0 = *reinterpret_cast<const float4*>(qs + g * HD + ch * 8);
                const float4 q1 = *reinterpret_cast<const float4*>(qs + g * HD + ch * 8 + 4);
                s[g] += (double)__fmul_rn(kf[0], q0.x); s[g] += (double)__fmul_rn(kf[1], q0.y);
                s[g] += (double)__fmul_rn(kf[2], q0.z); s[g] += (double)__fmul_rn(kf[3], q0.w);
                s[g] += (double)__fmul_rn(kf[4], q1.x); s[g] += (double)__fmul_rn(kf[5], q1.y);
                s[g] += (double)__fmul_rn(kf[6], q1.z); s[g] += (double)__fmul_rn(kf[7], q1.w);
            }
        }
#pragma unroll
        for (int g = 0; g < GRP; g++) {
            const float v = __fmul_rn((float)s[g], a.scale);
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
    // all-gather the per-CTA maxima -> global max (ggml soft_max: one max over the whole row)
    for (int i = tid; i < GRP * c.cluster; i += ATT_THREADS) {
        const int r = i / GRP, g = i % GRP;
        const float m = fmaxf(fmaxf(wred[g], wred[GRP + g]), fmaxf(wred[2 * GRP + g], wred[3 * GRP + g]));
        dsmem_st_f32(dsmem_addr(cmax + rk * GRP + g, r), m);
    }
    cluster_sync_all();
    if (tid < GRP) {
        float M = -INFINITY;
        for (int r = 0; r < c.cluster; r++) M = fmaxf(M, cmax[r * GRP + tid]);
        ml[tid] = M;
    }
    __syncthreads();

    // ---- phase 2: e = exp(s - max) rounded from f64, f64 sum, p = f16(e * (1/sum)) -------------------
    double lsum[GRP];
#pragma unroll
    for (int g = 0; g < GRP; g++) lsum[g] = 0.0;
    for (int i = tid; i < n_loc; i += ATT_THREADS) {
#pragma unroll
        for (int g = 0; g < GRP; g++) {
            const float e = (float)exp((double)__fsub_rn(sc[(size_t)g * c.max_local + i], ml[g]));
            sc[(size_t)g * c.max_local + i] = e;
            lsum[g] += (double)e;
        }
    }
#pragma unroll
    for (int g = 0; g < GRP; g++) {
        const double s = warp_sum_d(lsum[g]);
        if (lane == 0) wsum[warp * GRP + g] = s;
    }
    __syncthreads();
    for (int i = tid; i < GRP * c.cluster; i += ATT_THREADS) {
        const int r = i / GRP, g = i % GRP;
        dsmem_st_f64(dsmem_addr(csum + rk * GRP + g, r), (wsum[g] + wsum[GRP + g]) + (wsum[2 * GRP + g] + wsum[3 * GRP + g]));
    }
    cluster_sync_all();
    if (tid < GRP) {
        double S = 0;
        for (int r = 0; r < c.cluster; r++) S += csum[r * GRP + tid];
        ml[GRP + tid] = (float)(1.0 / S);          // ggml: sum = 1.0/sum in f64, then an f32 scale
    }
    __syncthreads();
    for (int i = tid; i < n_loc; i += ATT_THREADS) {
#pragma unroll
        for (int g = 0; g < GRP; g++)
            sc[(size_t)g * c.max_local + i] = __half2float(__float2half_rn(__fmul_rn(sc[(size_t)g * c.max_local + i], ml[GRP + g])));
    }
    __syncthreads();

    // ---- phase 3: o = sum_t V_t * p_t (exact f32 products, f64 sum); thread = (4-wide dim group,
    //      position lane) -----------------------------------------------------------------------------
    {
        const int dg = tid % ndg, pl = tid / ndg;
        double acc[GRP][4];
#pragma unroll
        for (int g = 0; g < GRP; g++)
#pragma unroll
            for (int i = 0; i < 4; i++) acc[g][i] = 0.0;
        if (pl < npl) {
            for (int t = t_begin + pl; t < t_end; t += npl) {
                const int page = a.page_table[t / a.page_tokens];
                const uint2 vv = *reinterpret_cast<const uint2*>(a.v_base + ((size_t)page * a.page_tokens + t % a.page_tokens) * a.kv_dim +
                                                                 (size_t)kvh * HD + dg * 4);
                const __half2* h2 = reinterpret_cast<const __half2*>(&vv);
                const float2 f0 = __half22float2(h2[0]), f1 = __half22float2(h2[1]);
#pragma unroll
                for (int g = 0; g < GRP; g++) {
                    const float p = sc[(size_t)g * c.max_local + (t - t_begin)];
                    acc[g][0] += (double)__fmul_rn(f0.x, p); acc[g][1] += (double)__fmul_rn(f0.y, p);
                    acc[g][2] += (double)__fmul_rn(f1.x, p); acc[g][3] += (double)__fmul_rn(f1.y, p);
                }
            }
#pragma unroll
            for (int g = 0; g < GRP; g++)
#pragma unroll
                for (int i = 0; i < 4; i++) ored[((size_t)pl * GRP + g) * HD + dg * 4 + i] = acc[g][i];
        }
    }
    __syncthreads();

    // ---- partial outputs of this slice -> rank 0 through DSMEM, summed there (f64) -------------------
    for (int i = tid; i < GRP * HD; i += ATT_THREADS) {
        const int g = i / HD, d = i % HD;
        double o = 0.0;
        for (int p = 0; p < npl; p++) o += ored[((size_t)p * GRP + g) * HD + d];
        dsmem_st_f64(dsmem_addr(merge + ((size_t)rk * GRP + g) * HD + d, 0), o);
    }
    cluster_sync_all();
    if (rk != 0) return;
    for (int i = tid; i < GRP * HD; i += ATT_THREADS) {
        const int g = i / HD, d = i % HD;
        double O = 0.0;
        for (int r = 0; r < c.cluster; r++) O += merge[((size_t)r * GRP + g) * HD + d];
        a.out[((size_t)kvh * GRP + g) * HD + d] = (float)O;
    }
}

size_t attn_smem_bytes(int grp, int hd, int max_local, int cluster) {
    const int ndg = hd / 4, npl = ATT_THREADS / ndg;
    const size_t d = (size_t)npl * grp * hd + (size_t)cluster * grp * hd + MAX_CLUSTER * grp + 4 * grp;            // doubles
    const size_t f = (size_t)grp * hd + (size_t)grp * max_local + 4 * grp + 2 * grp + MAX_CLUSTER * grp;          // floats
    return d * sizeof(double) + f * sizeof(float);
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
    if (a.head_dim % 8 || a.head_dim > 256 || a.head_dim < 32) B2_FAIL(B2_ERR_UNSUPPORTED, "attn: head_dim %d", a.head_dim);
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

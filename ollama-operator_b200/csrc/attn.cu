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

constexpr int ATT_THREADS = 256;
constexpr int ATT_WARPS = ATT_THREADS / 32;
constexpr int MAX_CLUSTER = 8;
constexpr int VCH = 128;          // V positions per shared-memory chunk (two chunks in flight)

struct AttnCfg {
    int cluster;        // CTAs per (slot, kv head)
    int max_local;      // capacity of the per-CTA score buffer (positions)
};

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int GRP, int HD>
__global__ void __launch_bounds__(ATT_THREADS) attn_decode_kernel(const AttnArgs a, const AttnCfg c) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t rk = cluster_ctarank();
    const int kvh = blockIdx.y;
    // phase 3 mapping: DW-wide dim groups x position lanes; DW shrinks with GRP so that the f64
    // accumulators stay at 32 registers
    constexpr int DW = GRP <= 2 ? 8 : GRP == 4 ? 4 : 2;
    constexpr int ndg = HD / DW, npl = ATT_THREADS / ndg;
    constexpr int ROW16 = HD / 8;                                   // 16-byte pieces per K/V row

    // shared-memory carve-up (doubles first: 8-byte alignment)
    double* ored = reinterpret_cast<double*>(smem_raw);             // [npl][GRP][HD] partial outputs
    double* merge = ored + (size_t)npl * GRP * HD;                  // rank 0 only: [cluster][GRP][HD]
    double* csum = merge + (size_t)c.cluster * GRP * HD;            // [MAX_CLUSTER][GRP] every CTA's sum (all-gathered)
    double* wsum = csum + MAX_CLUSTER * GRP;                        // [ATT_WARPS][GRP] cross-warp scratch
    uint4* vbuf = reinterpret_cast<uint4*>(wsum + ATT_WARPS * GRP); // [2][VCH][ROW16] V chunks (cp.async)
    float* qs = reinterpret_cast<float*>(vbuf + 2 * VCH * ROW16);   // [GRP][HD]   f16-rounded query
    float* sc = qs + GRP * HD;                                      // [GRP][max_local] scores -> probabilities
    float* wred = sc + (size_t)GRP * c.max_local;                   // [ATT_WARPS][GRP] cross-warp scratch
    float* ml = wred + ATT_WARPS * GRP;                             // [2][GRP] global max, 1/sum
    float* cmax = ml + 2 * GRP;                                     // [MAX_CLUSTER][GRP] every CTA's max (all-gathered)

    pdl_trigger();
    pdl_wait();      // q and this step's K/V rows come from the preceding kernel

    const int n = *a.pos + 1;
    const int per = (n + c.cluster - 1) / c.cluster;
    const int t_begin = min(n, (int)rk * per), t_end = min(n, t_begin + per);
    const int n_loc = t_end - t_begin;
    const int n_chunks = (n_loc + VCH - 1) / VCH;

    // V does not depend on the scores: start streaming the first two chunks into shared memory now
    auto issue_v_chunk = [&](int ch) {
        if (ch < n_chunks) {
            uint4* dst = vbuf + (size_t)(ch & 1) * VCH * ROW16;
            const int t0 = t_begin + ch * VCH, cnt = min(VCH, t_end - t0);
            for (int i = tid; i < cnt * ROW16; i += ATT_THREADS) {
                const int t = t0 + i / ROW16;
                const int page = a.page_table[t / a.page_tokens];
                const __half* src = a.v_base + ((size_t)page * a.page_tokens + t % a.page_tokens) * a.kv_dim + (size_t)kvh * HD + (i % ROW16) * 8;
                cp_async16(dst + i, src);
            }
        }
        cp_async_commit();
    };
    issue_v_chunk(0);
    issue_v_chunk(1);

    for (int i = tid; i < GRP * HD; i += ATT_THREADS)
        qs[i] = __half2float(__float2half_rn(a.q[(size_t)kvh * GRP * HD + i]));
    __syncthreads();

    // ---- phase 1: scores.  Two threads share a cached position (half of the dims each).  f16 x f16
    //      products are exact in f32 and are added in f64 (as ggml's scalar f16 dot does), so neither
    //      the split nor the four partial sums per thread change the result. ----------------------------
    constexpr int HALF16 = ROW16 / 2;          // 16-byte pieces per thread
    float lmax[GRP];
#pragma unroll
    for (int g = 0; g < GRP; g++) lmax[g] = -INFINITY;
    for (int base = 0; base < n_loc; base += ATT_THREADS / 2) {
        const int i = base + (tid >> 1), hsel = tid & 1;
        const bool active = i < n_loc;
        double s[GRP];
#pragma unroll
        for (int g = 0; g < GRP; g++) s[g] = 0.0;
        if (active) {
            const int t = t_begin + i;
            const int page = a.page_table[t / a.page_tokens];
            const uint4* kr = reinterpret_cast<const uint4*>(a.k_base + ((size_t)page * a.page_tokens + t % a.page_tokens) * a.kv_dim +
                                                             (size_t)kvh * HD) + hsel * HALF16;
            uint4 kreg[HALF16];
#pragma unroll
            for (int ch = 0; ch < HALF16; ch++) kreg[ch] = kr[ch];
            double part[GRP][4];
#pragma unroll
            for (int g = 0; g < GRP; g++)
#pragma unroll
                for (int j = 0; j < 4; j++) part[g][j] = 0.0;
#pragma unroll
            for (int ch = 0; ch < HALF16; ch++) {
                const __half2* h2 = reinterpret_cast<const __half2*>(&kreg[ch]);
                float kf[8];
#pragma unroll
                for (int j = 0; j < 4; j++) { const float2 f = __half22float2(h2[j]); kf[2 * j] = f.x; kf[2 * j + 1] = f.y; }
#pragma unroll
                for (int g = 0; g < GRP; g++) {
                    const float* qp = qs + g * HD + (hsel * HALF16 + ch) * 8;
                    const float4 q0 = *reinterpret_cast<const float4*>(qp), q1 = *reinterpret_cast<const float4*>(qp + 4);
                    part[g][0] += (double)__fmul_rn(kf[0], q0.x); part[g][1] += (double)__fmul_rn(kf[1], q0.y);
                    part[g][2] += (double)__fmul_rn(kf[2], q0.z); part[g][3] += (double)__fmul_rn(kf[3], q0.w);
                    part[g][0] += (double)__fmul_rn(kf[4], q1.x); part[g][1] += (double)__fmul_rn(kf[5], q1.y);
                    part[g][2] += (double)__fmul_rn(kf[6], q1.z); part[g][3] += (double)__fmul_rn(kf[7], q1.w);
                }
            }
#pragma unroll
            for (int g = 0; g < GRP; g++) s[g] = (part[g][0] + part[g][1]) + (part[g][2] + part[g][3]);
        }
#pragma unroll
        for (int g = 0; g < GRP; g++) {
            s[g] += __shfl_xor_sync(0xffffffffu, s[g], 1);          // the partner's half of the dims
            if (active && hsel == 0) {
                const float v = __fmul_rn((float)s[g], a.scale);
                sc[(size_t)g * c.max_local + i] = v;
                lmax[g] = fmaxf(lmax[g], v);
            }
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
        float m = wred[g];
#pragma unroll
        for (int w = 1; w < ATT_WARPS; w++) m = fmaxf(m, wred[w * GRP + g]);
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
    for (int i = tid; i < n_loc * GRP; i += ATT_THREADS) {
        const int g = i / n_loc, j = i % n_loc;
        const float e = (float)exp((double)__fsub_rn(sc[(size_t)g * c.max_local + j], ml[g]));
        sc[(size_t)g * c.max_local + j] = e;
#pragma unroll
        for (int gg = 0; gg < GRP; gg++) lsum[gg] += gg == g ? (double)e : 0.0;
    }
#pragma unroll
    for (int g = 0; g < GRP; g++) {
        const double s = warp_sum_d(lsum[g]);
        if (lane == 0) wsum[warp * GRP + g] = s;
    }
    __syncthreads();
    for (int i = tid; i < GRP * c.cluster; i += ATT_THREADS) {
        const int r = i / GRP, g = i % GRP;
        double sacc = 0.0;
#pragma unroll
        for (int w = 0; w < ATT_WARPS; w++) sacc += wsum[w * GRP + g];
        dsmem_st_f64(dsmem_addr(csum + rk * GRP + g, r), sacc);
    }
    cluster_sync_all();
    if (tid < GRP) {
        double S = 0;
        for (int r = 0; r < c.cluster; r++) S += csum[r * GRP + tid];
        ml[GRP + tid] = (float)(1.0 / S);          // ggml: sum = 1.0/sum in f64, then an f32 scale
    }
    __syncthreads();
    for (int i = tid; i < n_loc * GRP; i += ATT_THREADS) {
        const int g = i / n_loc, j = i % n_loc;
        sc[(size_t)g * c.max_local + j] = __half2float(__float2half_rn(__fmul_rn(sc[(size_t)g * c.max_local + j], ml[GRP + g])));
    }
    __syncthreads();

    // ---- phase 3: o = sum_t V_t * p_t (exact f32 products, f64 sum) from the shared-memory chunks;
    //      thread = (DW-wide dim group, position lane) -------------------------------------------------
    {
        const int dg = tid % ndg, pl = tid / ndg;
        double acc[GRP][DW];
#pragma unroll
        for (int g = 0; g < GRP; g++)
#pragma unroll
            for (int i = 0; i < DW; i++) acc[g][i] = 0.0;
        for (int ch = 0; ch < n_chunks; ch++) {
            cp_async_wait<1>();                  // chunk ch has landed (at most the next one is pending)
            __syncthreads();
            const __half* vch = reinterpret_cast<const __half*>(vbuf + (size_t)(ch & 1) * VCH * ROW16);
            const int cnt = min(VCH, n_loc - ch * VCH);
            if (pl < npl) {
                for (int j = pl; j < cnt; j += npl) {
                    const __half* vp = vch + (size_t)j * HD + dg * DW;
                    uint32_t vw[DW / 2];
                    if constexpr (DW == 8) { const uint4 v4 = *reinterpret_cast<const uint4*>(vp); vw[0] = v4.x; vw[1] = v4.y; vw[2] = v4.z; vw[3] = v4.w; }
                    else if constexpr (DW == 4) { const uint2 v2 = *reinterpret_cast<const uint2*>(vp); vw[0] = v2.x; vw[1] = v2.y; }
                    else { vw[0] = *reinterpret_cast<const uint32_t*>(vp); }
#pragma unroll
                    for (int i = 0; i < DW / 2; i++) {
                        const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&vw[i]));
#pragma unroll
                        for (int g = 0; g < GRP; g++) {
                            const float p = sc[(size_t)g * c.max_local + ch * VCH + j];
                            acc[g][2 * i] += (double)__fmul_rn(f.x, p);
                            acc[g][2 * i + 1] += (double)__fmul_rn(f.y, p);
                        }
                    }
                }
            }
            __syncthreads();                     // everyone is done with this buffer
            issue_v_chunk(ch + 2);
        }
        if (pl < npl) {
#pragma unroll
            for (int g = 0; g < GRP; g++)
#pragma unroll
                for (int i = 0; i < DW; i++) ored[((size_t)pl * GRP + g) * HD + dg * DW + i] = acc[g][i];
        }
    }
    cp_async_wait<0>();
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
    const int dw = grp <= 2 ? 8 : grp == 4 ? 4 : 2;
    const int ndg = hd / dw, npl = ATT_THREADS / ndg;
    const size_t d = (size_t)npl * grp * hd + (size_t)cluster * grp * hd + MAX_CLUSTER * grp + ATT_WARPS * grp;     // doubles
    const size_t v = (size_t)2 * VCH * (hd / 8) * 16;                                                               // V chunks
    const size_t f = (size_t)grp * hd + (size_t)grp * max_local + ATT_WARPS * grp + 2 * grp + MAX_CLUSTER * grp;    // floats
    return d * sizeof(double) + v + f * sizeof(float);
}

template <int GRP, int HD>
int launch_grp(const AttnArgs& a, const AttnCfg& c, size_t smem, cudaStream_t st, bool pdl) {
    auto kern = attn_decode_kernel<GRP, HD>;
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
#define B2_ATT_CASE(G, H) if (grp == G && a.head_dim == H) return launch_grp<G, H>(a, c, smem, st, pdl)
    B2_ATT_CASE(1, 64); B2_ATT_CASE(1, 80); B2_ATT_CASE(1, 128);
    B2_ATT_CASE(2, 64); B2_ATT_CASE(2, 80); B2_ATT_CASE(2, 128);
    B2_ATT_CASE(4, 64); B2_ATT_CASE(4, 80); B2_ATT_CASE(4, 128);
    B2_ATT_CASE(8, 64); B2_ATT_CASE(8, 80); B2_ATT_CASE(8, 128);
#undef B2_ATT_CASE
    if (a.head_dim != 64 && a.head_dim != 80 && a.head_dim != 128) B2_FAIL(B2_ERR_UNSUPPORTED, "attn: head_dim %d not supported (64, 80, 128)", a.head_dim);
    B2_FAIL(B2_ERR_UNSUPPORTED, "attn: %d query heads per KV head not supported (1, 2, 4, 8)", grp);
}

}  // namespace b2

// attn_dev.cuh -- decode attention of one KV head as a device function, shared by the stand-alone
// kernel (attn.cu) and the persistent decode kernel (mega.cu).  See attn.cu for the numerics.
#pragma once
#include <math.h>

#include "common.cuh"
#include "kernels.h"

namespace b2 {
namespace attndev {

constexpr int MAX_CLUSTER = 8;
constexpr int VCH = 64;           // V positions per shared-memory chunk (two chunks in flight)

struct AttnCfg {
    int cluster;        // CTAs per (slot, kv head)
    int max_local;      // capacity of the per-CTA score buffer (positions)
};

// One all-to-all step between the CTAs of a cluster without involving threads outside the NT
// callers: every CTA has written its contribution into its peers' shared memory (DSMEM stores
// above the call); one thread then arrives, with release.cluster semantics, on the exchange
// barrier of every peer, and everybody waits until all peers have arrived on its own.
static __device__ __forceinline__ void cluster_exchange(uint64_t* xb, uint32_t& phase, int cluster, int tid, int barid, int nt) {
    named_bar_sync(barid, nt);                       // this CTA's DSMEM stores are issued
    if (tid < cluster) {
        const uint32_t remote = dsmem_addr(xb, (uint32_t)tid);
        asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
    }
    uint32_t ok = 0;
    while (!ok) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(smem_u32(xb)), "r"(phase & 1u)
            : "memory");
    }
    phase++;
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Attention of one KV head (all GRP query heads that share it) by the NT calling threads of every
// CTA of the cluster.  `xb` is this CTA's exchange barrier (initialised to `cluster` arrivals,
// same shared-memory offset in every CTA), `xphase` its running phase; `barid` a CTA-local named
// barrier covering exactly the NT callers.
template <int GRP, int HD, int NT>
static __device__ __noinline__ void attn_kv_head(const AttnArgs& a, const AttnCfg& c, int kvh, uint8_t* smem_raw, int tid, uint64_t* xb, uint32_t& xphase,
                                    int barid, int pos_value = -1, const int* page_table_override = nullptr,
                                    unsigned long long* tr = nullptr) {
    auto stamp = [&](int k) { if (tr) { unsigned long long tt; tt = (unsigned long long)clock64(); tr[k] = tt; } };
    stamp(0);
    constexpr int ATT_THREADS = NT, ATT_WARPS = NT / 32;
    const int lane = tid & 31, warp = tid >> 5;
    const uint32_t rk = cluster_ctarank();
    // phase 3 mapping: DW-wide dim groups x position lanes; DW shrinks with GRP so that the f64
    // accumulators stay at 32 registers
    constexpr int DW = GRP <= 2 ? 8 : GRP == 4 ? 4 : 2;
    constexpr int NTV = ATT_THREADS < 256 ? ATT_THREADS : 256;      // threads that take part in phase 3
    constexpr int ndg = HD / DW, npl = NTV / ndg;
    constexpr int ROW16 = HD / 8;                                   // 16-byte pieces per K/V row

    // shared-memory carve-up (doubles first: 8-byte alignment)
    double* ored = reinterpret_cast<double*>(smem_raw);             // [npl][GRP][HD] partial outputs
    double* merge = ored + (size_t)npl * GRP * HD;                  // rank 0 only: [cluster][GRP][HD]
    double* csum = merge + (size_t)c.cluster * GRP * HD;            // [MAX_CLUSTER][GRP] every CTA's sum (all-gathered)
    double* wsum = csum + MAX_CLUSTER * GRP;                        // [ATT_WARPS][GRP] cross-warp scratch
    uint4* vbuf = reinterpret_cast<uint4*>(wsum + (ATT_WARPS * GRP + 1) / 2 * 2); // [2][VCH][ROW16] V chunks (cp.async), 16-byte aligned
    float* qs = reinterpret_cast<float*>(vbuf + 2 * VCH * ROW16);   // [GRP][HD]   f16-rounded query
    float* sc = qs + GRP * HD;                                      // [GRP][max_local] scores -> probabilities
    float* wred = sc + (size_t)GRP * c.max_local;                   // [ATT_WARPS][GRP] cross-warp scratch
    float* ml = wred + ATT_WARPS * GRP;                             // [2][GRP] global max, 1/sum
    float* cmax = ml + 2 * GRP;                                     // [MAX_CLUSTER][GRP] every CTA's max (all-gathered)

    // position and page table: the persistent kernel hands in copies it made once per step
    const int* page_table = page_table_override ? page_table_override : a.page_table;
    const int n = (pos_value >= 0 ? pos_value : *a.pos) + 1;
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
                const int page = page_table[t / a.page_tokens];
                const __half* src = a.v_base + ((size_t)page * a.page_tokens + t % a.page_tokens) * a.kv_dim + (size_t)kvh * HD + (i % ROW16) * 8;
                cp_async16(dst + i, src);
            }
        }
        cp_async_commit();
    };
    issue_v_chunk(0);
    issue_v_chunk(1);

    // ---- phase 1: scores.  Two threads share a cached position (half of the dims each).  f16 x f16
    //      products are exact in f32 and are added in f64 (as ggml's scalar f16 dot does), so neither
    //      the split nor the four partial sums per thread change the result.  The K rows of the first
    //      round are requested before the query is staged, so both latencies overlap. -------------------
    constexpr int HALF16 = ROW16 / 2;          // 16-byte pieces per thread
    const int hsel = tid & 1;
    uint4 kreg[HALF16];
    auto load_k = [&](int i) {
        if (i < n_loc) {
            const int t = t_begin + i;
            const int page = page_table[t / a.page_tokens];
            const uint4* kr = reinterpret_cast<const uint4*>(a.k_base + ((size_t)page * a.page_tokens + t % a.page_tokens) * a.kv_dim +
                                                             (size_t)kvh * HD) + hsel * HALF16;
#pragma unroll
            for (int ch = 0; ch < HALF16; ch++) kreg[ch] = kr[ch];
        }
    };
    load_k(tid >> 1);
    for (int i = tid; i < GRP * HD; i += ATT_THREADS)
        qs[i] = __half2float(__float2half_rn(__ldcg(a.q + (size_t)kvh * GRP * HD + i)));
    named_bar_sync(barid, NT);
    stamp(1);

    float lmax[GRP];
#pragma unroll
    for (int g = 0; g < GRP; g++) lmax[g] = -INFINITY;
    for (int base = 0; base < n_loc; base += ATT_THREADS / 2) {
        const int i = base + (tid >> 1);
        const bool active = i < n_loc;
        if (base > 0) load_k(i);
        double s[GRP];
#pragma unroll
        for (int g = 0; g < GRP; g++) s[g] = 0.0;
        if (active) {
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
    stamp(2);
#pragma unroll
    for (int g = 0; g < GRP; g++) {
        const float m = warp_max(lmax[g]);
        if (lane == 0) wred[warp * GRP + g] = m;
    }
    named_bar_sync(barid, NT);
    // all-gather the per-CTA maxima -> global max (ggml soft_max: one max over the whole row)
    for (int i = tid; i < GRP * c.cluster; i += ATT_THREADS) {
        const int r = i / GRP, g = i % GRP;
        float m = wred[g];
#pragma unroll
        for (int w = 1; w < ATT_WARPS; w++) m = fmaxf(m, wred[w * GRP + g]);
        dsmem_st_f32(dsmem_addr(cmax + rk * GRP + g, r), m);
    }
    cluster_exchange(xb, xphase, c.cluster, tid, barid, NT);
    if (tid < GRP) {
        float M = -INFINITY;
        for (int r = 0; r < c.cluster; r++) M = fmaxf(M, cmax[r * GRP + tid]);
        ml[tid] = M;
    }
    named_bar_sync(barid, NT);

    stamp(3);
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
    named_bar_sync(barid, NT);
    for (int i = tid; i < GRP * c.cluster; i += ATT_THREADS) {
        const int r = i / GRP, g = i % GRP;
        double sacc = 0.0;
#pragma unroll
        for (int w = 0; w < ATT_WARPS; w++) sacc += wsum[w * GRP + g];
        dsmem_st_f64(dsmem_addr(csum + rk * GRP + g, r), sacc);
    }
    cluster_exchange(xb, xphase, c.cluster, tid, barid, NT);
    stamp(4);
    if (tid < GRP) {
        double S = 0;
        for (int r = 0; r < c.cluster; r++) S += csum[r * GRP + tid];
        ml[GRP + tid] = (float)(1.0 / S);          // ggml: sum = 1.0/sum in f64, then an f32 scale
    }
    named_bar_sync(barid, NT);
    for (int i = tid; i < n_loc * GRP; i += ATT_THREADS) {
        const int g = i / n_loc, j = i % n_loc;
        sc[(size_t)g * c.max_local + j] = __half2float(__float2half_rn(__fmul_rn(sc[(size_t)g * c.max_local + j], ml[GRP + g])));
    }
    named_bar_sync(barid, NT);

    stamp(5);
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
            named_bar_sync(barid, NT);
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
            named_bar_sync(barid, NT);                     // everyone is done with this buffer
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
    named_bar_sync(barid, NT);

    stamp(6);
    // ---- partial outputs of this slice -> rank 0 through DSMEM, summed there (f64) -------------------
    for (int i = tid; i < GRP * HD; i += ATT_THREADS) {
        const int g = i / HD, d = i % HD;
        double o = 0.0;
        for (int p = 0; p < npl; p++) o += ored[((size_t)p * GRP + g) * HD + d];
        dsmem_st_f64(dsmem_addr(merge + ((size_t)rk * GRP + g) * HD + d, 0), o);
    }
    cluster_exchange(xb, xphase, c.cluster, tid, barid, NT);
    stamp(7);
    if (rk != 0) return;
    for (int i = tid; i < GRP * HD; i += ATT_THREADS) {
        const int g = i / HD, d = i % HD;
        double O = 0.0;
        for (int r = 0; r < c.cluster; r++) O += merge[((size_t)r * GRP + g) * HD + d];
        a.out[((size_t)kvh * GRP + g) * HD + d] = (float)O;
    }
}


static inline size_t attn_smem_bytes(int grp, int hd, int max_local, int cluster, int nthreads) {
    const int dw = grp <= 2 ? 8 : grp == 4 ? 4 : 2;
    const int ntv = nthreads < 256 ? nthreads : 256;
    const int ndg = hd / dw, npl = ntv / ndg, nwarps = nthreads / 32;
    const size_t d = (size_t)npl * grp * hd + (size_t)cluster * grp * hd + MAX_CLUSTER * grp + (size_t)(nwarps * grp + 1) / 2 * 2;   // doubles (even count: what follows needs 16-byte alignment)
    const size_t v = (size_t)2 * VCH * (hd / 8) * 16;                                                                   // V chunks
    const size_t f = (size_t)grp * hd + (size_t)grp * max_local + (size_t)nwarps * grp + 2 * grp + MAX_CLUSTER * grp;   // floats
    return (d * sizeof(double) + v + f * sizeof(float) + 15) / 16 * 16;
}

}  // namespace attndev
}  // namespace b2

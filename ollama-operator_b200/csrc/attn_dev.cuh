// attn_dev.cuh -- decode attention of one KV head as a device function, shared by the stand-alone
// kernel (attn.cu) and the persistent decode kernel (mega.cu).  See attn.cu for the numerics.
#pragma once
#include <math.h>

#include "common.cuh"
#include "kernels.h"

namespace b2 {
namespace attndev {

constexpr int MAX_CLUSTER = 8;

struct AttnCfg {
    int cluster;        // CTAs per (slot, kv head)
    int max_local;      // capacity of the score buffer: ALL positions of the sequence (every CTA holds them)
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

// Butterfly with the operand order of ggml's 8-wide SIMD reduction (GGML_F32x8_REDUCE over four
// accumulator vectors): lanes (L, L+16), then (L, L+8), (L, L+4), then the two horizontal adds.
__device__ __forceinline__ float reduce_lanes32(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    return v;
}

// L2 prefetch of the cache rows attn_kv_head will read for `kvh` in this CTA: its slice of K rows
// (whole rows) and its slice of the head dimension of every V row.  Issued well ahead of the
// attention (the persistent kernel does it while the QKV weights stream), so that the dependent
// loads of the attention itself hit L2 instead of queueing behind the weight stream in HBM.
template <int GRP, int HD, int NT>
static __device__ __forceinline__ void attn_prefetch(const AttnArgs& a, int cluster, int kvh, int tid, int pos_value, const int* page_table) {
    const uint32_t rk = cluster_ctarank();
    const int n = pos_value + 1;
    const int per = (n + cluster - 1) / cluster;
    const int t_begin = min(n, (int)rk * per), t_end = min(n, t_begin + per);
    constexpr int KLINES = (HD * 2 + 127) / 128;                       // 128-byte lines per K row
    for (int i = tid; i < (t_end - t_begin) * KLINES; i += NT) {
        const int t = t_begin + i / KLINES;
        const int page = page_table[t / a.page_tokens];
        const __half* p = a.k_base + ((size_t)page * a.page_tokens + t % a.page_tokens) * a.kv_dim + (size_t)kvh * HD + (i % KLINES) * 64;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    }
    const int dpc = HD / cluster;                                      // this CTA's output dims
    const int vlines = (dpc * 2 + 127) / 128;
    for (int i = tid; i < n * vlines; i += NT) {
        const int t = i / vlines;
        const int page = page_table[t / a.page_tokens];
        const __half* p = a.v_base + ((size_t)page * a.page_tokens + t % a.page_tokens) * a.kv_dim + (size_t)kvh * HD + (int)rk * dpc + (i % vlines) * 64;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    }
}

// Attention of one KV head (all GRP query heads that share it) by the NT calling threads of every
// CTA of the cluster.  `xb` is this CTA's exchange barrier (initialised to `cluster` arrivals,
// same shared-memory offset in every CTA), `xphase` its running phase; `barid` a CTA-local named
// barrier covering exactly the NT callers.
//   1. scores: the cached positions are split over the cluster's CTAs, one WARP per position; lane L
//      owns dims L, L+32, ... as an f32 FMA chain and the warp butterfly reproduces ggml's SIMD
//      reduction order, so a score is bit-identical to ggml_vec_dot_f16 on an AVX2 host.  Every score
//      is written into every CTA's score buffer through DSMEM (one exchange).
//   2. soft-max over all positions, redundantly in every CTA (no exchange): one max, exp rounded
//      from f64, f64 sum, probabilities rounded to f16 as ggml does before the V product.
//   3. output: the head dimension is split over the cluster's CTAs (no merge needed); lane L owns the
//      positions t = L mod 32 as an FMA chain -- again ggml's lane structure -- four dims per warp.
template <int GRP, int HD, int NT>
static __device__ __noinline__ void attn_kv_head(const AttnArgs& a, const AttnCfg& c, int kvh, uint8_t* smem_raw, int tid, uint64_t* xb, uint32_t& xphase,
                                                 int barid, int pos_value = -1, const int* page_table_override = nullptr,
                                                 unsigned long long* tr = nullptr) {
    auto stamp = [&](int k) { if (tr) tr[k] = (unsigned long long)clock64(); };
    stamp(0);
    constexpr int NWARP = NT / 32;
    constexpr int HD32 = HD & ~31;             // dims that go through the SIMD lanes
    const int lane = tid & 31, warp = tid >> 5;
    const uint32_t rk = cluster_ctarank();
    const int CL = c.cluster, maxc = c.max_local;
    const int* page_table = page_table_override ? page_table_override : a.page_table;
    const int n = (pos_value >= 0 ? pos_value : *a.pos) + 1;
    // operands into registers once (the descriptor lives in global memory)
    const __half* const k_base = a.k_base + (size_t)kvh * HD;
    const __half* const v_base = a.v_base + (size_t)kvh * HD;
    const float scale = a.scale;
    const int kv_dim = a.kv_dim, pshift = __ffs(a.page_tokens) - 1, pmask = a.page_tokens - 1;      // pages are a power of two

    double* wsum = reinterpret_cast<double*>(smem_raw);              // [NWARP][GRP]
    float* qs = reinterpret_cast<float*>(wsum + ((NWARP * GRP + 1) & ~1));   // [GRP][HD] f16-rounded query
    float* sc = qs + GRP * HD;                                       // [GRP][maxc] scores -> probabilities (all positions)
    float* wred = sc + (size_t)GRP * maxc;                           // [NWARP][GRP]
    float* ml = wred + NWARP * GRP;                                  // [2][GRP] max, 1/sum
    int* rowoff = reinterpret_cast<int*>(ml + 2 * GRP);              // [maxc] element offset of every cached row inside a layer's pool

    for (int i = tid; i < GRP * HD; i += NT) qs[i] = __half2float(__float2half_rn(__ldcg(a.q + (size_t)kvh * GRP * HD + i)));
    // page-table walk once per position, in parallel, instead of inside the latency-bound loops below
    for (int t = tid; t < n; t += NT) rowoff[t] = ((page_table[t >> pshift] << pshift) + (t & pmask)) * kv_dim;
    named_bar_sync(barid, NT);
    stamp(1);

    // ---- 1. scores of this CTA's slice of the context ---------------------------------------------
    const int per = (n + CL - 1) / CL;
    const int t_begin = min(n, (int)rk * per), t_end = min(n, t_begin + per);
    // U positions per warp are in flight at once (the loads of a batch are issued before any of them
    // is used): the phase is bound by memory latency, not by arithmetic
    constexpr int U = GRP >= 4 ? 4 : (HD <= 128 ? 12 : 8);
    constexpr int NI = HD32 / 32 > 0 ? HD32 / 32 : 1;
    for (int t0 = t_begin + warp; t0 < t_end; t0 += NWARP * U) {
        const __half* kr[U];
        float kv[U][NI];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int t = min(t0 + u * NWARP, t_end - 1);
            kr[u] = k_base + rowoff[t];
#pragma unroll
            for (int i = 0; i < HD32 / 32; i++) kv[u][i] = __half2float(kr[u][i * 32 + lane]);
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int t = t0 + u * NWARP;
            if (t < t_end) {
                float acc[GRP];
#pragma unroll
                for (int g = 0; g < GRP; g++) {
                    acc[g] = 0.f;
#pragma unroll
                    for (int i = 0; i < HD32 / 32; i++) acc[g] = fmaf(kv[u][i], qs[g * HD + i * 32 + lane], acc[g]);
                    acc[g] = reduce_lanes32(acc[g]);
                    if (HD32 != HD) {                   // ggml adds the tail beyond a multiple of 32 in f64
                        double sumf = (double)acc[g];
                        for (int d = HD32; d < HD; d++) sumf += (double)__fmul_rn(__half2float(kr[u][d]), qs[g * HD + d]);
                        acc[g] = (float)sumf;
                    }
                    acc[g] = __fmul_rn(acc[g], scale);
                }
                if (lane < CL) {
#pragma unroll
                    for (int g = 0; g < GRP; g++) dsmem_st_f32(dsmem_addr(sc + (size_t)g * maxc + t, (uint32_t)lane), acc[g]);
                }
            }
        }
    }
    stamp(2);
    cluster_exchange(xb, xphase, CL, tid, barid, NT);      // every CTA now holds all n scores
    stamp(3);

    // ---- 2. soft-max over all positions (ggml soft_max: one max, expf, f64 sum, scale by 1/sum) ------
    float lmax[GRP];
#pragma unroll
    for (int g = 0; g < GRP; g++) lmax[g] = -INFINITY;
    for (int t = tid; t < n; t += NT) {
#pragma unroll
        for (int g = 0; g < GRP; g++) lmax[g] = fmaxf(lmax[g], sc[(size_t)g * maxc + t]);
    }
#pragma unroll
    for (int g = 0; g < GRP; g++) {
        const float m = warp_max(lmax[g]);
        if (lane == 0) wred[warp * GRP + g] = m;
    }
    named_bar_sync(barid, NT);
    if (tid < GRP) {
        float M = wred[tid];
        for (int w = 1; w < NWARP; w++) M = fmaxf(M, wred[w * GRP + tid]);
        ml[tid] = M;
    }
    named_bar_sync(barid, NT);
    double lsum[GRP];
#pragma unroll
    for (int g = 0; g < GRP; g++) lsum[g] = 0.0;
    for (int t = tid; t < n; t += NT) {
#pragma unroll
        for (int g = 0; g < GRP; g++) {
            const float e = (float)exp((double)__fsub_rn(sc[(size_t)g * maxc + t], ml[g]));
            sc[(size_t)g * maxc + t] = e;
            lsum[g] += (double)e;
        }
    }
#pragma unroll
    for (int g = 0; g < GRP; g++) {
        const double s = warp_sum_d(lsum[g]);
        if (lane == 0) wsum[warp * GRP + g] = s;
    }
    named_bar_sync(barid, NT);
    if (tid < GRP) {
        double S = 0;
        for (int w = 0; w < NWARP; w++) S += wsum[w * GRP + tid];
        ml[GRP + tid] = (float)(1.0 / S);
    }
    named_bar_sync(barid, NT);
    for (int t = tid; t < n; t += NT) {
#pragma unroll
        for (int g = 0; g < GRP; g++)
            sc[(size_t)g * maxc + t] = __half2float(__float2half_rn(__fmul_rn(sc[(size_t)g * maxc + t], ml[GRP + g])));
    }
    named_bar_sync(barid, NT);
    stamp(4);

    // ---- 3. output dims of this CTA: V columns dotted with the probabilities, lane L = position mod 32 --
    const int dpc = HD / CL, dbase = (int)rk * dpc;      // HD % (4 * CL) == 0 (checked on the host)
    for (int d0 = 4 * warp; d0 < dpc; d0 += 4 * NWARP) {
        float acc[GRP][4];
#pragma unroll
        for (int g = 0; g < GRP; g++)
#pragma unroll
            for (int j = 0; j < 4; j++) acc[g][j] = 0.f;
        constexpr int UV = GRP >= 4 ? 8 : 20;       // V rows in flight per lane
        for (int tb = lane; tb < n; tb += 32 * UV) {
            uint2 vv[UV];
#pragma unroll
            for (int u = 0; u < UV; u++) {
                const int t = min(tb + 32 * u, n - 1);
                vv[u] = *reinterpret_cast<const uint2*>(v_base + rowoff[t] + dbase + d0);
            }
#pragma unroll
            for (int u = 0; u < UV; u++) {
                const int t = tb + 32 * u;
                if (t < n) {
                    const float2 f0 = __half22float2(*reinterpret_cast<const __half2*>(&vv[u].x)), f1 = __half22float2(*reinterpret_cast<const __half2*>(&vv[u].y));
#pragma unroll
                    for (int g = 0; g < GRP; g++) {
                        const float p = sc[(size_t)g * maxc + t];
                        acc[g][0] = fmaf(f0.x, p, acc[g][0]); acc[g][1] = fmaf(f0.y, p, acc[g][1]);
                        acc[g][2] = fmaf(f1.x, p, acc[g][2]); acc[g][3] = fmaf(f1.y, p, acc[g][3]);
                    }
                }
            }
        }
#pragma unroll
        for (int g = 0; g < GRP; g++) {
#pragma unroll
            for (int j = 0; j < 4; j++) acc[g][j] = reduce_lanes32(acc[g][j]);
            if (lane == 0)
                *reinterpret_cast<float4*>(a.out + ((size_t)kvh * GRP + g) * HD + dbase + d0) = make_float4(acc[g][0], acc[g][1], acc[g][2], acc[g][3]);
        }
    }
    stamp(5);
    named_bar_sync(barid, NT);      // the score buffer is free again (a following head reuses it)
}

static inline size_t attn_smem_bytes(int grp, int hd, int max_ctx_pad, int cluster, int nthreads) {
    (void)cluster;
    const int nwarps = nthreads / 32;
    const size_t d = (size_t)((nwarps * grp + 1) & ~1);                                        // doubles
    const size_t f = (size_t)grp * hd + (size_t)grp * max_ctx_pad + (size_t)nwarps * grp + 2 * grp + (size_t)max_ctx_pad;   // floats + row offsets
    return (d * sizeof(double) + f * sizeof(float) + 15) / 16 * 16;
}

}  // namespace attndev
}  // namespace b2

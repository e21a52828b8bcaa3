// attn_dev.cuh -- decode attention of one KV head as a device function, shared by the stand-alone
// kernel (attn.cu) and the persistent decode kernel (mega.cu).  See attn.cu for the numerics.
#pragma once
#include <math.h>

#include <type_traits>

#include "common.cuh"
#include "kernels.h"

namespace b2 {
namespace attndev {

constexpr int MAX_CLUSTER = 8;

struct AttnCfg {
    int cluster;        // CTAs per (slot, kv head)
    int max_local;      // capacity of the score buffer: ALL positions of the sequence (every CTA holds them)
    bool ll_out = false;    // also write the hand-over copy a.out_ll and re-poison a.out_ll_poison (persistent kernel, see GemvArgs::x_ll)
    bool ll_in = false;     // q and the current position's K/V rows arrive through a.qkv_ll (polled), not behind a grid barrier
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
    SpinGuard guard;
    while (!ok) {
        guard.tick();
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
    // cluster is 1, 2 or 4 and pages are a power of two: shifts, no integer divisions on the path of the stage that issues this
    const int cshift = __ffs(cluster) - 1, pshift = __ffs(a.page_tokens) - 1, pmask = a.page_tokens - 1;
    const int per = (n + cluster - 1) >> cshift;
    const int t_begin = min(n, (int)rk * per), t_end = min(n, t_begin + per);
    constexpr int KLINES = (HD * 2 + 127) / 128;                       // 128-byte lines per K row
    for (int i = tid; i < (t_end - t_begin) * KLINES; i += NT) {
        const int t = t_begin + i / KLINES;
        const int page = page_table[t >> pshift];
        const __half* p = a.k_base + (((size_t)page << pshift) + (t & pmask)) * a.kv_dim + (size_t)kvh * HD + (i % KLINES) * 64;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    }
    const int dpc = HD >> cshift;                                      // this CTA's output dims
    const int vlines = (dpc * 2 + 127) / 128;
    if (vlines == 1) {
        for (int t = tid; t < n; t += NT) {
            const int page = page_table[t >> pshift];
            const __half* p = a.v_base + (((size_t)page << pshift) + (t & pmask)) * a.kv_dim + (size_t)kvh * HD + (int)rk * dpc;
            asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
        }
    } else {
        for (int i = tid; i < n * vlines; i += NT) {
            const int t = i / vlines;
            const int page = page_table[t >> pshift];
            const __half* p = a.v_base + (((size_t)page << pshift) + (t & pmask)) * a.kv_dim + (size_t)kvh * HD + (int)rk * dpc + (i % vlines) * 64;
            asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
        }
    }
}

// Eight independent 32-lane sums at once, each reduced in exactly the order of reduce_lanes32: at
// the xor-16 / 8 / 4 steps a lane keeps half of its values and hands the other half to its partner,
// so 9 shuffles do the work of 40.  Lane L ends up with product number ((L>>4)&1)*4 + ((L>>3)&1)*2 +
// ((L>>2)&1); the four lanes that share bits 2..4 hold the same value.
__device__ __forceinline__ float reduce8_lanes32(const float (&v)[8], int lane) {
    const bool b4 = (lane & 16) != 0, b3 = (lane & 8) != 0, b2 = (lane & 4) != 0;
    float k[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const float keep = b4 ? v[4 + i] : v[i], send = b4 ? v[i] : v[4 + i];
        k[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
    float m[2];
#pragma unroll
    for (int i = 0; i < 2; i++) {
        const float keep = b3 ? k[2 + i] : k[i], send = b3 ? k[i] : k[2 + i];
        m[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    float r = (b2 ? m[1] : m[0]) + __shfl_xor_sync(0xffffffffu, b2 ? m[0] : m[1], 4);
    r += __shfl_xor_sync(0xffffffffu, r, 1);
    r += __shfl_xor_sync(0xffffffffu, r, 2);
    return r;
}
__device__ __forceinline__ int reduce8_index(int lane) { return ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1); }

// The wide-row layout of the P.V product applies when the CTA's slice of the head is 32 or 64 dims and the per-chain
// partial sums ([GRP][32][dpc] floats) stay small.
template <int GRP> __host__ __device__ __forceinline__ bool pv_wide(int dpc) { return (dpc == 32 || dpc == 64) && GRP * dpc <= 128; }

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
//      positions t = L mod 32 as an FMA chain -- again ggml's lane structure -- eight dims per warp.
// The whole thing is latency-bound (a few hundred KB per head), so the code is organised around
// having many independent loads in flight and few dependent shuffles.
template <int GRP, int HD, int NT>
static __device__ __noinline__ void attn_kv_head(const AttnArgs& a, const AttnCfg& c, int kvh, uint8_t* smem_raw, int tid, uint64_t* xb, uint32_t& xphase,
                                                 int barid, int pos_value = -1, const int* page_table_override = nullptr,
                                                 unsigned long long* tr = nullptr, int kv_index = -1) {
    // kvh indexes the query-head group (q and out); kv_index, when given, the KV head whose cache rows are read:
    // a GQA group can be processed one query head at a time (GRP = 1, kvh = query head, kv_index = its KV head)
    // by different clusters -- every (head, position) product is computed the same way, so the bits do not change.
    auto stamp = [&](int k) { if (tr) tr[k] = (unsigned long long)clock64(); };
    stamp(0);
    constexpr int NWARP = NT / 32;
    constexpr int HD32 = HD & ~31;             // dims that go through the SIMD lanes
    constexpr int NI = HD32 / 32 > 0 ? HD32 / 32 : 1;
    static_assert(8 % GRP == 0, "GRP must divide 8");
    constexpr int PB = 8 / GRP;                // positions per reduce8 batch (PB * GRP = 8 dot products)
    const int lane = tid & 31, warp = tid >> 5;
    const uint32_t rk = cluster_ctarank();
    const int CL = c.cluster, maxc = c.max_local;
    const int* page_table = page_table_override ? page_table_override : a.page_table;
    const int n = (pos_value >= 0 ? pos_value : *a.pos) + 1;
    // operands into registers once (the descriptor lives in global memory)
    const __half* const k_base = a.k_base + (size_t)(kv_index >= 0 ? kv_index : kvh) * HD;
    const __half* const v_base = a.v_base + (size_t)(kv_index >= 0 ? kv_index : kvh) * HD;
    const float scale = a.scale;
    const int kv_dim = a.kv_dim, pshift = __ffs(a.page_tokens) - 1, pmask = a.page_tokens - 1;      // pages are a power of two

    double* wsum = reinterpret_cast<double*>(smem_raw);              // [NWARP][GRP]
    float* qs = reinterpret_cast<float*>(wsum + ((NWARP * GRP + 1) & ~1));   // [GRP][HD] f16-rounded query
    float* sc = qs + GRP * HD;                                       // [GRP][maxc] scores -> probabilities (all positions)
    float* wred = sc + (size_t)GRP * maxc;                           // [NWARP][GRP]
    int* rowoff = reinterpret_cast<int*>(wred + NWARP * GRP);        // [maxc] element offset of every cached row inside a layer's pool

    if (c.ll_in) {
        // Hand-over from the QKV stage (no grid barrier in between): poll the query; and poll the new cache rows this CTA is
        // going to read -- the K row if it scores the last position, its slice of the V row -- and store them into the cache
        // itself.  The QKV stage stores the same bits there, but only a value seen in the hand-over buffer is known to have
        // arrived; after the block barrier below the loops read the cache as for every other position.
        const int kvi = kv_index >= 0 ? kv_index : kvh, nq = a.n_head * HD;
        const uint32_t* words = reinterpret_cast<const uint32_t*>(a.qkv_ll) + nq;
        const int last = n - 1, per0 = (n + CL - 1) / CL, dpc0 = HD / CL;
        const bool own_k = last >= (int)rk * per0 && last < (int)rk * per0 + per0;
        const size_t row = (size_t)((page_table[last >> pshift] << pshift) + (last & pmask)) * kv_dim;
        for (int i = tid; i < GRP * HD + HD / 2 + dpc0 / 2; i += NT) {
            SpinGuard guard;
            if (i < GRP * HD) {
                float v;
                do { v = __uint_as_float(ll_load_u32(reinterpret_cast<const uint32_t*>(a.qkv_ll) + (size_t)kvh * GRP * HD + i)); guard.tick(); } while (__float_as_uint(v) == LL_POISON);
                qs[i] = __half2float(__float2half_rn(v));
            } else if (i < GRP * HD + HD / 2) {
                if (own_k) {
                    const int w = i - GRP * HD;
                    uint32_t v;
                    do { v = ll_load_u32(words + (size_t)kvi * (HD / 2) + w); guard.tick(); } while (v == LL_POISON);
                    reinterpret_cast<uint32_t*>(const_cast<__half*>(k_base + row))[w] = v;
                }
            } else {
                const int w = i - GRP * HD - HD / 2;
                uint32_t v;
                do { v = ll_load_u32(words + (a.kv_dim >> 1) + (size_t)kvi * (HD / 2) + (int)rk * (dpc0 / 2) + w); guard.tick(); } while (v == LL_POISON);
                reinterpret_cast<uint32_t*>(const_cast<__half*>(v_base + row + (int)rk * dpc0))[w] = v;
            }
        }
    } else {
        for (int i = tid; i < GRP * HD; i += NT) qs[i] = __half2float(__float2half_rn(__ldcg(a.q + (size_t)kvh * GRP * HD + i)));
    }
    // page-table walk once per position, in parallel, instead of inside the latency-bound loops below
    for (int t = tid; t < n; t += NT) rowoff[t] = ((page_table[t >> pshift] << pshift) + (t & pmask)) * kv_dim;
    named_bar_sync(barid, NT);
    stamp(1);

    // ---- 1. scores of this CTA's slice of the context ---------------------------------------------
    const int per = (n + CL - 1) / CL;
    const int t_begin = min(n, (int)rk * per), t_end = min(n, t_begin + per);
    if (HD == 128 && GRP == 1) {
        // Wide K rows (one query head, head_dim 128): four lanes share a position, lane j0 of them loads the 16-byte chunks
        // j0, j0 + 4, j0 + 8, j0 + 12 of the 256-byte row and thereby owns the eight complete FMA chains L = 8 j0 + e
        // (chain L runs over dims L, L + 32, L + 64, L + 96, in that order -- ggml's lane structure).  Four 128-bit loads
        // per lane replace thirty-two 16-bit ones (the 2-byte loads, not their latency, bounded this phase); the chains
        // are then combined in exactly the order of reduce_lanes32: L^16 and L^8 live in the lanes j0^2 and j0^1, L^4, L^1
        // and L^2 partly in the thread itself.
        const int j0 = lane & 3, pg = lane >> 2;
        const bool b1 = (j0 & 2) != 0, b0 = (j0 & 1) != 0;
        float qr[4][8];
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
            for (int e = 0; e < 8; e++) qr[i][e] = qs[32 * i + 8 * j0 + e];
        for (int t0 = t_begin + warp * 8; t0 < t_end; t0 += NWARP * 8) {
            const int t = min(t0 + pg, t_end - 1);
            const __half* kr = k_base + rowoff[t] + 8 * j0;
            uint4 kq[4];
#pragma unroll
            for (int i = 0; i < 4; i++) kq[i] = *reinterpret_cast<const uint4*>(kr + 32 * i);
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; e++) v[e] = 0.f;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const __half2* h2 = reinterpret_cast<const __half2*>(&kq[i]);
#pragma unroll
                for (int e2 = 0; e2 < 4; e2++) {
                    const float2 f = __half22float2(h2[e2]);
                    v[2 * e2] = fmaf(f.x, qr[i][2 * e2], v[2 * e2]);
                    v[2 * e2 + 1] = fmaf(f.y, qr[i][2 * e2 + 1], v[2 * e2 + 1]);
                }
            }
            float k4[4], k2[2];
#pragma unroll
            for (int i = 0; i < 4; i++) {                 // L ^ 16: lane ^ 2; this lane keeps e in [4 b1, 4 b1 + 4)
                const float keep = b1 ? v[4 + i] : v[i], send = b1 ? v[i] : v[4 + i];
                k4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
            }
#pragma unroll
            for (int i = 0; i < 2; i++) {                 // L ^ 8: lane ^ 1; keeps e = 4 b1 + 2 b0 + i
                const float keep = b0 ? k4[2 + i] : k4[i], send = b0 ? k4[i] : k4[2 + i];
                k2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
            }
#pragma unroll
            for (int i = 0; i < 2; i++) k2[i] = k2[i] + __shfl_xor_sync(0xffffffffu, k2[i], 2);      // L ^ 4
            float r = k2[0] + k2[1];                                                                   // L ^ 1
            r = r + __shfl_xor_sync(0xffffffffu, r, 1);                                                // L ^ 2
            if (t0 + pg < t_end) {
                r = __fmul_rn(r, scale);
                for (int peer = j0; peer < CL; peer += 4) dsmem_st_f32(dsmem_addr(sc + t, (uint32_t)peer), r);
            }
        }
    } else
    {
        float qreg[GRP][NI];
#pragma unroll
        for (int g = 0; g < GRP; g++)
#pragma unroll
            for (int i = 0; i < HD32 / 32; i++) qreg[g][i] = qs[g * HD + i * 32 + lane];
        // a warp takes PB consecutive positions per batch and keeps NBATCH batches of loads in flight
        constexpr int NBATCH = 2;
        for (int t0 = t_begin + warp * PB; t0 < t_end; t0 += NWARP * PB * NBATCH) {
            float kv[NBATCH][PB][NI];
#pragma unroll
            for (int bq = 0; bq < NBATCH; bq++)
#pragma unroll
                for (int u = 0; u < PB; u++) {
                    const int t = min(t0 + bq * NWARP * PB + u, t_end - 1);
                    const __half* kr = k_base + rowoff[t];
#pragma unroll
                    for (int i = 0; i < HD32 / 32; i++) kv[bq][u][i] = __half2float(kr[i * 32 + lane]);
                }
#pragma unroll
            for (int bq = 0; bq < NBATCH; bq++) {
                const int tb = t0 + bq * NWARP * PB;
                if (tb < t_end) {
                    float prod[8];
#pragma unroll
                    for (int u = 0; u < PB; u++)
#pragma unroll
                        for (int g = 0; g < GRP; g++) {
                            float acc = 0.f;
#pragma unroll
                            for (int i = 0; i < HD32 / 32; i++) acc = fmaf(kv[bq][u][i], qreg[g][i], acc);
                            prod[u * GRP + g] = acc;
                        }
                    float r = reduce8_lanes32(prod, lane);
                    const int idx = reduce8_index(lane), u = idx / GRP, g = idx % GRP, t = tb + u;
                    if (t < t_end) {
                        if (HD32 != HD) {               // ggml adds the tail beyond a multiple of 32 in f64
                            const __half* kr = k_base + rowoff[t];
                            double sumf = (double)r;
                            for (int d = HD32; d < HD; d++) sumf += (double)__fmul_rn(__half2float(kr[d]), qs[g * HD + d]);
                            r = (float)sumf;
                        }
                        r = __fmul_rn(r, scale);
                        // the four lanes that hold this product write it to up to four peers
                        for (int peer = lane & 3; peer < CL; peer += 4) dsmem_st_f32(dsmem_addr(sc + (size_t)g * maxc + t, (uint32_t)peer), r);
                    }
                }
            }
        }
    }
    stamp(2);
    cluster_exchange(xb, xphase, CL, tid, barid, NT);      // every CTA now holds all n scores
    stamp(3);

    // ---- 2. soft-max over all positions (ggml soft_max: one max, expf, f64 sum, scale by 1/sum) ------
    float M[GRP], inv[GRP];
    {
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
#pragma unroll
        for (int g = 0; g < GRP; g++) {
            M[g] = wred[g];
#pragma unroll
            for (int w = 1; w < NWARP; w++) M[g] = fmaxf(M[g], wred[w * GRP + g]);
        }
        double lsum[GRP];
#pragma unroll
        for (int g = 0; g < GRP; g++) lsum[g] = 0.0;
        for (int t = tid; t < n; t += NT) {
#pragma unroll
            for (int g = 0; g < GRP; g++) {
                const float e = (float)exp((double)__fsub_rn(sc[(size_t)g * maxc + t], M[g]));
                sc[(size_t)g * maxc + t] = e;          // only this thread touches element t until the barrier below
                lsum[g] += (double)e;
            }
        }
#pragma unroll
        for (int g = 0; g < GRP; g++) {
            const double s = warp_sum_d(lsum[g]);
            if (lane == 0) wsum[warp * GRP + g] = s;
        }
        named_bar_sync(barid, NT);
#pragma unroll
        for (int g = 0; g < GRP; g++) {
            double S = 0;
#pragma unroll
            for (int w = 0; w < NWARP; w++) S += wsum[w * GRP + g];
            inv[g] = (float)(1.0 / S);                  // ggml: sum = 1.0/sum in f64, then an f32 scale
        }
    }
    // the scaling by 1/sum and the rounding to f16 happen where the P.V product reads a probability (one pass and one barrier less)
    auto prob = [&](int g, int t) { return __half2float(__float2half_rn(__fmul_rn(sc[(size_t)g * maxc + t], inv[g]))); };
    stamp(4);

    // ---- 3. output dims of this CTA: V columns dotted with the probabilities ------------------------------
    // ggml's chain structure: for one output dim, positions t = c, c + 32, ... form FMA chain c (32 chains),
    // combined by the fixed tree of reduce_lanes32.
    const int dpc = HD / CL, dbase = (int)rk * dpc;      // HD % (8 * CL) == 0 (checked on the host)
    if (pv_wide<GRP>(dpc)) {
        // Wide rows: `lpd` lanes cover the CTA's dims of one cached row (16 bytes each), so a warp load touches
        // 32 / lpd whole row slices instead of 32 different rows (the L1 tag stage handles one 128-byte line per
        // cycle: 16-byte pieces of 32 rows cost 32 cycles per load, and every row was visited by dpc / 8 warps).
        // Warp w owns chains c = w * rpw + lane / lpd; a chain stays in one thread, in position order, so the sums
        // are the ones the narrow layout below produces.  The 32 chain sums per dim meet in shared memory.
        float* pvs = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(rowoff + maxc) + 15) & ~(uintptr_t)15);        // [GRP][32 chains][dpc]
        // One query head (GRP = 1): 4 dims (8 bytes) per lane instead of 8 -- twice the warps share the work and a lane
        // holds half the registers per row, which pays for more rows in flight (a context just past a multiple of 512
        // would otherwise cost a second, exposed L2 round trip for one or two rows per chain).
        auto pv_part = [&](auto dpl_tag) {
        constexpr int DPL = decltype(dpl_tag)::value;           // dims per lane
        const int lpd = dpc / DPL, rpw = 32 / lpd;
        if (warp < lpd) {
            const int dl = lane % lpd, c = warp * rpw + lane / lpd;
            float acc[GRP][DPL];
#pragma unroll
            for (int g = 0; g < GRP; g++)
#pragma unroll
                for (int j = 0; j < DPL; j++) acc[g][j] = 0.f;
            constexpr int UV = GRP >= 4 ? 4 : GRP > 1 ? 8 : DPL == 4 ? 20 : 16;        // V rows in flight per lane
            for (int tb = c; tb < n; tb += 32 * UV) {
                uint32_t vv[UV][DPL / 2];
#pragma unroll
                for (int u = 0; u < UV; u++) {
                    const int t = min(tb + 32 * u, n - 1);
                    const __half* src = v_base + rowoff[t] + dbase + dl * DPL;
                    if constexpr (DPL == 8) {
                        const uint4 q = *reinterpret_cast<const uint4*>(src);
                        vv[u][0] = q.x; vv[u][1] = q.y; vv[u][DPL / 2 - 2] = q.z; vv[u][DPL / 2 - 1] = q.w;
                    } else {
                        const uint2 q = *reinterpret_cast<const uint2*>(src);
                        vv[u][0] = q.x; vv[u][1] = q.y;
                    }
                }
#pragma unroll
                for (int u = 0; u < UV; u++) {
                    const int t = tb + 32 * u;
                    if (t < n) {
                        float vf[DPL];
#pragma unroll
                        for (int j = 0; j < DPL / 2; j++) {
                            const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&vv[u][j]));
                            vf[2 * j] = f.x; vf[2 * j + 1] = f.y;
                        }
#pragma unroll
                        for (int g = 0; g < GRP; g++) {
                            const float p = prob(g, t);
#pragma unroll
                            for (int j = 0; j < DPL; j++) acc[g][j] = fmaf(vf[j], p, acc[g][j]);
                        }
                    }
                }
            }
#pragma unroll
            for (int g = 0; g < GRP; g++) {
                float4* dst = reinterpret_cast<float4*>(pvs + ((size_t)g * 32 + c) * dpc + dl * DPL);
                dst[0] = make_float4(acc[g][0], acc[g][1], acc[g][2], acc[g][3]);
                if constexpr (DPL == 8) dst[1] = make_float4(acc[g][DPL - 4], acc[g][DPL - 3], acc[g][DPL - 2], acc[g][DPL - 1]);
            }
        }
        };
        if (GRP == 1 && dpc / 4 <= NWARP) pv_part(std::integral_constant<int, 4>{});
        else pv_part(std::integral_constant<int, 8>{});
        named_bar_sync(barid, NT);
        for (int o = tid; o < GRP * dpc; o += NT) {
            const int g = o / dpc, d = o - g * dpc;
            const float* v = pvs + (size_t)g * 32 * dpc + d;
            float s4[4];
#pragma unroll
            for (int c = 0; c < 4; c++) {       // lane c of reduce_lanes32 after the xor-16, 8, 4 steps
                const float a0 = v[c * dpc] + v[(c + 16) * dpc], a1 = v[(c + 8) * dpc] + v[(c + 24) * dpc];
                const float a2 = v[(c + 4) * dpc] + v[(c + 20) * dpc], a3 = v[(c + 12) * dpc] + v[(c + 28) * dpc];
                s4[c] = (a0 + a1) + (a2 + a3);
            }
            const float r = (s4[0] + s4[1]) + (s4[2] + s4[3]);
            a.out[((size_t)kvh * GRP + g) * HD + dbase + d] = r;
            if (c.ll_out) {
                ll_store(a.out_ll + ((size_t)kvh * GRP + g) * HD + dbase + d, r);
                ll_store(a.out_ll_poison + ((size_t)kvh * GRP + g) * HD + dbase + d, __uint_as_float(LL_POISON));
            }
        }
    } else
    {
    for (int d0 = 8 * warp; d0 < dpc; d0 += 8 * NWARP) {
        float acc[GRP][8];
#pragma unroll
        for (int g = 0; g < GRP; g++)
#pragma unroll
            for (int j = 0; j < 8; j++) acc[g][j] = 0.f;
        constexpr int UV = GRP >= 4 ? 4 : 10;       // V rows in flight per lane
        for (int tb = lane; tb < n; tb += 32 * UV) {
            uint4 vv[UV];
#pragma unroll
            for (int u = 0; u < UV; u++) {
                const int t = min(tb + 32 * u, n - 1);
                vv[u] = *reinterpret_cast<const uint4*>(v_base + rowoff[t] + dbase + d0);
            }
#pragma unroll
            for (int u = 0; u < UV; u++) {
                const int t = tb + 32 * u;
                if (t < n) {
                    const __half2* h2 = reinterpret_cast<const __half2*>(&vv[u]);
                    float vf[8];
#pragma unroll
                    for (int j = 0; j < 4; j++) { const float2 f = __half22float2(h2[j]); vf[2 * j] = f.x; vf[2 * j + 1] = f.y; }
#pragma unroll
                    for (int g = 0; g < GRP; g++) {
                        const float p = prob(g, t);
#pragma unroll
                        for (int j = 0; j < 8; j++) acc[g][j] = fmaf(vf[j], p, acc[g][j]);
                    }
                }
            }
        }
#pragma unroll
        for (int g = 0; g < GRP; g++) {
            const float r = reduce8_lanes32(acc[g], lane);
            if ((lane & 3) == 0) {
                a.out[((size_t)kvh * GRP + g) * HD + dbase + d0 + reduce8_index(lane)] = r;
                if (c.ll_out) {
                    ll_store(a.out_ll + ((size_t)kvh * GRP + g) * HD + dbase + d0 + reduce8_index(lane), r);
                    ll_store(a.out_ll_poison + ((size_t)kvh * GRP + g) * HD + dbase + d0 + reduce8_index(lane), __uint_as_float(LL_POISON));
                }
            }
        }
    }
    }
    stamp(5);
    named_bar_sync(barid, NT);      // the score buffer is free again (a following head reuses it)
}

static inline size_t attn_smem_bytes(int grp, int hd, int max_ctx_pad, int cluster, int nthreads) {
    const int nwarps = nthreads / 32;
    const int dpc = cluster > 0 ? hd / cluster : hd;
    // chain sums of the wide P.V layout (+ alignment slack); a GQA group may also be processed one query head at a time (GRP = 1)
    const size_t pv = (dpc == 32 || dpc == 64) ? (size_t)(grp * dpc <= 128 ? grp : 1) * 32 * dpc + 4 : 0;
    const size_t d = (size_t)((nwarps * grp + 1) & ~1);                                        // doubles
    const size_t f = (size_t)grp * hd + (size_t)grp * max_ctx_pad + (size_t)nwarps * grp + (size_t)max_ctx_pad + pv;   // floats + row offsets
    return (d * sizeof(double) + f * sizeof(float) + 15) / 16 * 16;
}

}  // namespace attndev
}  // namespace b2

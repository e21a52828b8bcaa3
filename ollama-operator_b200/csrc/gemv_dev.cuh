// gemv_dev.cuh -- device building blocks of the fused W4A8 GEMV, shared by the per-matrix kernel
// (gemv.cu) and the persistent decode kernel (mega.cu): the activation prologue (norm + Q8
// quantisation into shared memory), the per-type register cache of the quantised activation,
// the per-unit integer dot with ggml's float scale terms, and the epilogues.
#pragma once
#include <math.h>

#include "common.cuh"
#include "formats.cuh"
#include "kernels.h"

namespace b2 {
namespace gemvdev {

__device__ __forceinline__ float h2f(const __half h) { return __half2float(h); }

// ---------------------------------------------------------------------------------------------
// prologue: f32 activation -> normalised -> quantised into shared memory
// ---------------------------------------------------------------------------------------------
struct ProStats { float mean, scale; };

// Sum over the consumer threads of a double, same value returned to every thread.
// `reuse_barrier` = false: the caller guarantees a consumer barrier between this call and the next use of `scratch`
static __device__ double consumers_sum(double v, double* scratch, int tid, int ntc, bool reuse_barrier = true) {
    v = warp_sum_d(v);
    if ((tid & 31) == 0) scratch[tid >> 5] = v;
    named_bar_sync(1, ntc);
    double s = 0;
    for (int w = 0; w < ntc / 32; w++) s += scratch[w];
    if (reuse_barrier) named_bar_sync(1, ntc);
    return s;
}

// Four values of the stage input.  In the persistent kernel a stage may read a hand-over buffer (GemvArgs::x_ll) that other
// SMs wrote during this launch with no L1 invalidation in between: those reads go to L2 (the caller has polled the buffer
// complete before).  Everywhere else x was written by an earlier kernel and the plain load stands.
__device__ __forceinline__ float4 load_x4(const GemvArgs& a, int i) {
    return a.x_ll ? ll_load4(a.x_ll + i) : *reinterpret_cast<const float4*>(a.x + i);
}

static __device__ ProStats prologue_stats(const GemvArgs& a, double* scratch, int tid, int ntc) {
    ProStats st{0.f, 1.f};
    const int K = (int)a.W.k;
    if (a.pro == PRO_RMSNORM) {
        // ggml rms_norm: f64 sum of f32 squares; scale = 1/sqrtf(mean + eps)
        double ss = 0;
        for (int i = tid * 4; i < K; i += ntc * 4) {
            const float4 v = load_x4(a, i);
            ss += (double)__fmul_rn(v.x, v.x); ss += (double)__fmul_rn(v.y, v.y);
            ss += (double)__fmul_rn(v.z, v.z); ss += (double)__fmul_rn(v.w, v.w);
        }
        ss = consumers_sum(ss, scratch, tid, ntc);
        const float mean = (float)(ss / (double)K);
        st.scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, a.eps)));
    } else if (a.pro == PRO_LAYERNORM) {
        // ggml norm: mean, then centred sum of squares, both in f64
        double s = 0;
        for (int i = tid * 4; i < K; i += ntc * 4) {
            const float4 v = load_x4(a, i);
            s += (double)v.x; s += (double)v.y; s += (double)v.z; s += (double)v.w;
        }
        s = consumers_sum(s, scratch, tid, ntc);
        st.mean = (float)(s / (double)K);
        double ss = 0;
        for (int i = tid * 4; i < K; i += ntc * 4) {
            const float4 v = load_x4(a, i);
            const float c0 = __fsub_rn(v.x, st.mean), c1 = __fsub_rn(v.y, st.mean), c2 = __fsub_rn(v.z, st.mean), c3 = __fsub_rn(v.w, st.mean);
            ss += (double)__fmul_rn(c0, c0); ss += (double)__fmul_rn(c1, c1);
            ss += (double)__fmul_rn(c2, c2); ss += (double)__fmul_rn(c3, c3);
        }
        ss = consumers_sum(ss, scratch, tid, ntc);
        const float var = (float)(ss / (double)K);
        st.scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(var, a.eps)));
    }
    return st;
}

__device__ __forceinline__ float4 prologue_vals(const GemvArgs& a, const ProStats& st, int i) {
    float4 v = load_x4(a, i);
    if (a.pro == PRO_RMSNORM) {
        const float4 g = *reinterpret_cast<const float4*>(a.gain + i);
        v.x = __fmul_rn(__fmul_rn(v.x, st.scale), g.x); v.y = __fmul_rn(__fmul_rn(v.y, st.scale), g.y);
        v.z = __fmul_rn(__fmul_rn(v.z, st.scale), g.z); v.w = __fmul_rn(__fmul_rn(v.w, st.scale), g.w);
    } else if (a.pro == PRO_LAYERNORM) {
        const float4 g = *reinterpret_cast<const float4*>(a.gain + i);
        float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
        if (a.nbias) b = *reinterpret_cast<const float4*>(a.nbias + i);
        v.x = __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(v.x, st.mean), st.scale), g.x), b.x);
        v.y = __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(v.y, st.mean), st.scale), g.y), b.y);
        v.z = __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(v.z, st.mean), st.scale), g.z), b.z);
        v.w = __fadd_rn(__fmul_rn(__fmul_rn(__fsub_rn(v.w, st.mean), st.scale), g.w), b.w);
    }
    return v;
}

// amax / 127 correctly rounded without the generic division sequence: one Newton step on a
// correctly rounded reciprocal is exact division for normal results; tiny inputs take the slow path.
__device__ __forceinline__ float div127_rn(float a) {
    if (a < 1e-30f) return __fdiv_rn(a, 127.0f);
    const float c = 0.007874015718698502f;           // fl(1/127)
    const float q = __fmul_rn(a, c);
    const float r = __fmaf_rn(-q, 127.0f, a);         // exact remainder
    return __fmaf_rn(r, c, q);
}

// (int)roundf(t) for |t| <= 256 (round half away from zero) in three instructions: add the largest float below
// one half with the sign of t, then truncate.  Checked against roundf() for every float in that range on the host.
__device__ __forceinline__ int round_away_i(float t) {
    return __float2int_rz(__fadd_rn(t, copysignf(0x1.fffffep-2f, t)));
}

__device__ __forceinline__ uint32_t pack4(int a, int b, int c, int d) {
    return (uint32_t)(a & 0xFF) | ((uint32_t)(b & 0xFF) << 8) | ((uint32_t)(c & 0xFF) << 16) | ((uint32_t)(d & 0xFF) << 24);
}

// Q8_0 (quantize_row_q8_0_ref): per 32: d = amax/127, q = roundf(x/d) (half away from zero),
// stored scale rounded to f16.  Eight lanes own one block (a float4 each).
// Shared layout: xq = two planes of 16 B per block (elements 0..15 | 16..31) so that the
// per-lane LDS.128 of the main loop is conflict-free; xd[b] = f16-rounded d; xs[b] = 8*sum(q)
// (folds Q4_0's -8 offset: sum (w-8) q = sum w q - 8 sum q, exactly, in integers).
// The quantiser proper: store(block, offset in block (multiple of 4), four packed int8, f16-rounded scale, 8*sum(q));
// scale and sum are valid for offset 0 only.
template <class Store>
static __device__ __forceinline__ void prologue_q8_0_each(const GemvArgs& a, const ProStats& st, int tid, int ntc, Store store, int k0 = 0, int k1 = -1) {
    const int K = k1 < 0 ? (int)a.W.k : k1;          // elements [k0, K): k0 and K multiples of 32
    const unsigned lane = tid & 31, gmask = 0xFFu << (lane & ~7u);
    for (int i = k0 + tid * 4; i < K; i += ntc * 4) {
        const float4 v = prologue_vals(a, st, i);
        float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
        amax = fmaxf(amax, __shfl_xor_sync(gmask, amax, 1));
        amax = fmaxf(amax, __shfl_xor_sync(gmask, amax, 2));
        amax = fmaxf(amax, __shfl_xor_sync(gmask, amax, 4));
        const float d = div127_rn(amax);
        const float id = d != 0.0f ? __frcp_rn(d) : 0.0f;
        const int q0 = round_away_i(__fmul_rn(v.x, id)), q1 = round_away_i(__fmul_rn(v.y, id));
        const int q2 = round_away_i(__fmul_rn(v.z, id)), q3 = round_away_i(__fmul_rn(v.w, id));
        int sum = q0 + q1 + q2 + q3;
        sum += __shfl_xor_sync(gmask, sum, 1);
        sum += __shfl_xor_sync(gmask, sum, 2);
        sum += __shfl_xor_sync(gmask, sum, 4);
        store(i >> 5, i & 31, pack4(q0, q1, q2, q3), __half2float(__float2half_rn(d)), 8 * sum);
    }
}
static __device__ void prologue_q8_0(const GemvArgs& a, const ProStats& st, uint8_t* xq, float* xd, int* xs, int tid, int ntc) {
    const int nbk = (int)a.W.k / 32;
    prologue_q8_0_each(a, st, tid, ntc, [=](int b, int o, uint32_t w, float d, int s8) {
        *reinterpret_cast<uint32_t*>(xq + (o < 16 ? 0 : nbk * 16) + b * 16 + (o & 15)) = w;
        if (o == 0) { xd[b] = d; xs[b] = s8; }
    });
}

// Q8_K (quantize_row_q8_K_ref): per 256: iscale = -127/max (max = signed value of largest
// magnitude), q = min(127, nearest_int(iscale*x)), d = 1/iscale (f32), 16-element sums.
// One warp owns one super-block (8 elements per lane).
// Shared layout: xq natural order; xd[sb] = d; xs[K/16] = 16-element sums.
static __device__ void prologue_q8_K(const GemvArgs& a, const ProStats& st, uint8_t* xq, float* xd, int* xs, int tid, int ntc) {
    const int K = (int)a.W.k, nsb = K / 256;
    const int lane = tid & 31, warp = tid >> 5, nw = ntc >> 5;
    for (int sb = warp; sb < nsb; sb += nw) {
        const int i = sb * 256 + lane * 8;
        const float4 v0 = prologue_vals(a, st, i), v1 = prologue_vals(a, st, i + 4);
        const float e[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
        float amax = 0.f, mx = 0.f;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const float ax = fabsf(e[j]);
            if (ax > amax) { amax = ax; mx = e[j]; }
        }
        // first element of largest magnitude wins, as in the sequential scan
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const float oa = __shfl_xor_sync(0xffffffffu, amax, o), om = __shfl_xor_sync(0xffffffffu, mx, o);
            const bool other_first = (lane ^ o) < lane;
            if (oa > amax || (oa == amax && other_first)) { amax = oa; mx = om; }
        }
        int q[8];
        float d = 0.f;
        if (amax != 0.f) {
            const float iscale = __fdiv_rn(-127.f, mx);
#pragma unroll
            for (int j = 0; j < 8; j++) q[j] = min(127, __float2int_rn(__fmul_rn(iscale, e[j])));
            d = __fdiv_rn(1.0f, iscale);
        } else {
#pragma unroll
            for (int j = 0; j < 8; j++) q[j] = 0;
        }
        int sum = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) sum += q[j];
        sum += __shfl_xor_sync(0xffffffffu, sum, 1);
        uint2 w;
        w.x = pack4(q[0], q[1], q[2], q[3]);
        w.y = pack4(q[4], q[5], q[6], q[7]);
        *reinterpret_cast<uint2*>(xq + i) = w;
        if ((lane & 1) == 0) xs[sb * 16 + (lane >> 1)] = sum;
        if (lane == 0) xd[sb] = d;
    }
}

// ---------------------------------------------------------------------------------------------
// per-type register cache of the quantised activation slice and the row dot
// ---------------------------------------------------------------------------------------------
template <int TYPE> struct XRegs;

template <> struct XRegs<T_Q4_0> {
    uint4 a, b; float d; int s8;
    __device__ void load(const uint8_t* xq, const float* xd, const int* xs, int u, int nunits, bool valid, int = 0) {
        if (valid) {
            a = *reinterpret_cast<const uint4*>(xq + u * 16);
            b = *reinterpret_cast<const uint4*>(xq + nunits * 16 + u * 16);
            d = xd[u]; s8 = xs[u];
        } else { a = b = make_uint4(0, 0, 0, 0); d = 0.f; s8 = 0; }
    }
};
template <> struct XRegs<T_Q8_0> {
    uint4 a, b; float d;
    __device__ void load(const uint8_t* xq, const float* xd, const int*, int u, int nunits, bool valid, int = 0) {
        if (valid) {
            a = *reinterpret_cast<const uint4*>(xq + u * 16);
            b = *reinterpret_cast<const uint4*>(xq + nunits * 16 + u * 16);
            d = xd[u];
        } else { a = b = make_uint4(0, 0, 0, 0); d = 0.f; }
    }
};
template <> struct XRegs<T_Q6_K> {
    uint32_t w[32]; int bs32[8]; float d;
    __device__ void load(const uint8_t* xq, const float* xd, const int* xs, int u, int, bool valid, int = 0) {
        if (valid) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const uint4 t = *reinterpret_cast<const uint4*>(xq + u * 128 + i * 16);
                w[4 * i] = t.x; w[4 * i + 1] = t.y; w[4 * i + 2] = t.z; w[4 * i + 3] = t.w;
            }
#pragma unroll
            for (int g = 0; g < 8; g++) bs32[g] = 32 * xs[u * 8 + g];
            d = xd[u >> 1];
        } else {
#pragma unroll
            for (int i = 0; i < 32; i++) w[i] = 0;
#pragma unroll
            for (int g = 0; g < 8; g++) bs32[g] = 0;
            d = 0.f;
        }
    }
};
template <> struct XRegs<T_Q4_K> {
    uint32_t w[32]; int bs[4]; float d;
    __device__ void load(const uint8_t* xq, const float* xd, const int* xs, int u, int, bool valid, int = 0) {
        if (valid) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const uint4 t = *reinterpret_cast<const uint4*>(xq + u * 128 + i * 16);
                w[4 * i] = t.x; w[4 * i + 1] = t.y; w[4 * i + 2] = t.z; w[4 * i + 3] = t.w;
            }
#pragma unroll
            for (int j = 0; j < 4; j++) bs[j] = xs[u * 8 + 2 * j] + xs[u * 8 + 2 * j + 1];
            d = xd[u >> 1];
        } else {
#pragma unroll
            for (int i = 0; i < 32; i++) w[i] = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) bs[j] = 0;
            d = 0.f;
        }
    }
};

// Either K-quant, chosen per matrix at run time (warp-uniform): the same registers serve both layouts.
template <> struct XRegs<T_KQ> {
    uint32_t w[32]; int bs[8]; float d; int type;
    __device__ void load(const uint8_t* xq, const float* xd, const int* xs, int u, int, bool valid, int t) {
        type = t;
        if (valid) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const uint4 q = *reinterpret_cast<const uint4*>(xq + u * 128 + i * 16);
                w[4 * i] = q.x; w[4 * i + 1] = q.y; w[4 * i + 2] = q.z; w[4 * i + 3] = q.w;
            }
            if (t == T_Q6_K) {
#pragma unroll
                for (int g = 0; g < 8; g++) bs[g] = 32 * xs[u * 8 + g];
            } else {
#pragma unroll
                for (int j = 0; j < 4; j++) { bs[j] = xs[u * 8 + 2 * j] + xs[u * 8 + 2 * j + 1]; bs[4 + j] = 0; }
            }
            d = xd[u >> 1];
        } else {
#pragma unroll
            for (int i = 0; i < 32; i++) w[i] = 0;
#pragma unroll
            for (int g = 0; g < 8; g++) bs[g] = 0;
            d = 0.f;
        }
    }
};

// Pointers to one row of a stage, per plane.
struct RowPtrs { const uint8_t* p[4]; };

__device__ __forceinline__ uint32_t u4c(const uint4& v, int i) { return i == 0 ? v.x : i == 1 ? v.y : i == 2 ? v.z : v.w; }

// unsigned-by-signed byte dot (dp4a.u32.s32): d = c + sum_i a.u8[i] * b.s8[i]
__device__ __forceinline__ int dp4a_us(uint32_t a, int b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// ggml_vec_dot_q4_0_q8_0: sumf += sumi * d_w * d_a, sumi exact.  The high nibbles are used in
// place (w & 0xF0F0F0F0, as unsigned bytes = 16 * nibble): their dot is an exact multiple of 16,
// so one shift at the end replaces four; the two accumulators also halve the dp4a dependency chain.
__device__ __forceinline__ float unit_dot(const XRegs<T_Q4_0>& x, const RowPtrs& r, int u, int) {
    const uint4 w = *reinterpret_cast<const uint4*>(r.p[0] + u * 16);
    const float dw = h2f(*reinterpret_cast<const __half*>(r.p[1] + u * 2));
    int lo = 0, hi = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const uint32_t ww = u4c(w, i);
        lo = __dp4a((int)(ww & 0x0F0F0F0Fu), (int)u4c(x.a, i), lo);
        hi = dp4a_us(ww & 0xF0F0F0F0u, (int)u4c(x.b, i), hi);
    }
    return __fmul_rn(__fmul_rn((float)(lo + (hi >> 4) - x.s8), dw), x.d);
}
// ggml_vec_dot_q8_0_q8_0: sumf += sumi * (d_w * d_a); plane0 is chunk-interleaved [2][nunits][16]
__device__ __forceinline__ float unit_dot(const XRegs<T_Q8_0>& x, const RowPtrs& r, int u, int nunits) {
    const uint4 wa = *reinterpret_cast<const uint4*>(r.p[0] + u * 16);
    const uint4 wb = *reinterpret_cast<const uint4*>(r.p[0] + (nunits + u) * 16);
    const float dw = h2f(*reinterpret_cast<const __half*>(r.p[1] + u * 2));
    int s = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        s = __dp4a((int)u4c(wa, i), (int)u4c(x.a, i), s);
        s = __dp4a((int)u4c(wb, i), (int)u4c(x.b, i), s);
    }
    return __fmul_rn((float)s, __fmul_rn(dw, x.d));
}
// ggml_vec_dot_q6_K_q8_K for one half super-block (128 weights); planes are chunk-interleaved:
// ql = [4][nunits][16], qh = [2][nunits][16], scales = [nunits][8], d = [nunits/2] f16
__device__ __forceinline__ float unit_dot(const XRegs<T_Q6_K>& x, const RowPtrs& r, int u, int nunits) {
    uint4 A[2], B[2], H[2];
    A[0] = *reinterpret_cast<const uint4*>(r.p[0] + (0 * nunits + u) * 16);
    A[1] = *reinterpret_cast<const uint4*>(r.p[0] + (1 * nunits + u) * 16);
    B[0] = *reinterpret_cast<const uint4*>(r.p[0] + (2 * nunits + u) * 16);
    B[1] = *reinterpret_cast<const uint4*>(r.p[0] + (3 * nunits + u) * 16);
    H[0] = *reinterpret_cast<const uint4*>(r.p[1] + (0 * nunits + u) * 16);
    H[1] = *reinterpret_cast<const uint4*>(r.p[1] + (1 * nunits + u) * 16);
    const uint2 scw = *reinterpret_cast<const uint2*>(r.p[2] + u * 8);
    const float dw = h2f(*reinterpret_cast<const __half*>(r.p[3] + (u >> 1) * 2));
    int acc = 0;
#pragma unroll
    for (int hh = 0; hh < 2; hh++) {        // hh = l/16: which 16 of the 32 "l" positions
        int s0 = 0, s1 = 0, s2 = 0, s3 = 0;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t a = u4c(A[hh], i), b = u4c(B[hh], i), h = u4c(H[hh], i);
            const int wl = hh * 4 + i;
            s0 = __dp4a((int)((a & 0x0F0F0F0Fu) | ((h << 4) & 0x30303030u)), (int)x.w[wl], s0);
            s1 = __dp4a((int)((b & 0x0F0F0F0Fu) | ((h << 2) & 0x30303030u)), (int)x.w[8 + wl], s1);
            s2 = __dp4a((int)(((a >> 4) & 0x0F0F0F0Fu) | (h & 0x30303030u)), (int)x.w[16 + wl], s2);
            s3 = __dp4a((int)(((b >> 4) & 0x0F0F0F0Fu) | ((h >> 2) & 0x30303030u)), (int)x.w[24 + wl], s3);
        }
        // scale index = quad*2 + l/16; bytes of scw: [0..7]
        const int sc0 = (int)(int8_t)((hh == 0 ? scw.x : scw.x >> 8) & 0xFF);
        const int sc1 = (int)(int8_t)((hh == 0 ? scw.x >> 16 : scw.x >> 24) & 0xFF);
        const int sc2 = (int)(int8_t)((hh == 0 ? scw.y : scw.y >> 8) & 0xFF);
        const int sc3 = (int)(int8_t)((hh == 0 ? scw.y >> 16 : scw.y >> 24) & 0xFF);
        acc += sc0 * (s0 - x.bs32[0 + hh]) + sc1 * (s1 - x.bs32[2 + hh]) + sc2 * (s2 - x.bs32[4 + hh]) + sc3 * (s3 - x.bs32[6 + hh]);
    }
    return __fmul_rn(__fmul_rn(dw, x.d), (float)acc);
}
// ggml_vec_dot_q4_K_q8_K for one half super-block; qs = [4][nunits][16], scales = [nunits/2][12],
// {d,dmin} = [nunits/2] half2
__device__ __forceinline__ float unit_dot(const XRegs<T_Q4_K>& x, const RowPtrs& r, int u, int nunits) {
    uint4 Q[4];
#pragma unroll
    for (int c = 0; c < 4; c++) Q[c] = *reinterpret_cast<const uint4*>(r.p[0] + (c * nunits + u) * 16);
    const uint32_t* sp = reinterpret_cast<const uint32_t*>(r.p[1] + (u >> 1) * 12);
    const uint32_t w0 = sp[0], w1 = sp[1], w2 = sp[2];
    const __half2 dm = *reinterpret_cast<const __half2*>(r.p[2] + (u >> 1) * 4);
    // 6-bit scales/mins of sub-blocks 4h..4h+3 (h = u & 1); bytes: w0 = q[0..3], w1 = q[4..7], w2 = q[8..11]
    uint32_t sc4, mn4;   // four bytes each
    if ((u & 1) == 0) { sc4 = w0 & 0x3F3F3F3Fu; mn4 = w1 & 0x3F3F3F3Fu; }
    else { sc4 = (w2 & 0x0F0F0F0Fu) | ((w0 >> 2) & 0x30303030u); mn4 = ((w2 >> 4) & 0x0F0F0F0Fu) | ((w1 >> 2) & 0x30303030u); }
    int accs = 0, accm = 0;
#pragma unroll
    for (int pr = 0; pr < 2; pr++) {        // pr: first / second 32 bytes of this unit's qs
        int lo = 0, hi = 0;
#pragma unroll
        for (int c = 0; c < 2; c++)
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const uint32_t q = u4c(Q[pr * 2 + c], i);
                const int wi = c * 4 + i;
                lo = __dp4a((int)(q & 0x0F0F0F0Fu), (int)x.w[(2 * pr) * 8 + wi], lo);
                hi = __dp4a((int)((q >> 4) & 0x0F0F0F0Fu), (int)x.w[(2 * pr + 1) * 8 + wi], hi);
            }
        const int scl = (sc4 >> (16 * pr)) & 0xFF, sch = (sc4 >> (16 * pr + 8)) & 0xFF;
        const int mnl = (mn4 >> (16 * pr)) & 0xFF, mnh = (mn4 >> (16 * pr + 8)) & 0xFF;
        accs += scl * lo + sch * hi;
        accm += mnl * x.bs[2 * pr] + mnh * x.bs[2 * pr + 1];
    }
    const float d = __fmul_rn(__low2float(dm), x.d), dmin = __fmul_rn(__high2float(dm), x.d);
    return __fsub_rn(__fmul_rn(d, (float)accs), __fmul_rn(dmin, (float)accm));
}

__device__ __forceinline__ float unit_dot(const XRegs<T_KQ>& x, const RowPtrs& r, int u, int nunits) {
    if (x.type == T_Q6_K) {
        XRegs<T_Q6_K> t;
#pragma unroll
        for (int i = 0; i < 32; i++) t.w[i] = x.w[i];
#pragma unroll
        for (int g = 0; g < 8; g++) t.bs32[g] = x.bs[g];
        t.d = x.d;
        return unit_dot(t, r, u, nunits);
    }
    XRegs<T_Q4_K> t;
#pragma unroll
    for (int i = 0; i < 32; i++) t.w[i] = x.w[i];
#pragma unroll
    for (int j = 0; j < 4; j++) t.bs[j] = x.bs[j];
    t.d = x.d;
    return unit_dot(t, r, u, nunits);
}

// ---------------------------------------------------------------------------------------------
// epilogues
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long argmax_key(float v, uint32_t row) {
    uint32_t u = __float_as_uint(v);
    u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);      // order-preserving map to unsigned
    return ((unsigned long long)u << 32) | (unsigned long long)(0xFFFFFFFFu - row);   // ties -> lowest row
}

__device__ __forceinline__ float gelu_f16_table(float v) {
    // ggml_vec_gelu_f32: tanh form evaluated through an f16 lookup table
    if (v <= -10.0f) return 0.0f;
    if (v >= 10.0f) return v;
    const float t = __half2float(__float2half_rn(v));
    const float u = __fmul_rn(__fmul_rn(0.79788456080286535587989211986876f, t), __fadd_rn(1.0f, __fmul_rn(__fmul_rn(0.044715f, t), t)));
    const float th = (float)tanh((double)u);          // correctly rounded f32, as the oracle's canonical mode
    return __half2float(__float2half_rn(__fmul_rn(__fmul_rn(0.5f, t), __fadd_rn(1.0f, th))));
}

// What a thread of the persistent kernel keeps from layer to layer for "its" row pair (same rows in every layer): the
// residual values it wrote last and the RoPE (cos, sin) of its pair -- both would otherwise be an L2 round trip per layer.
struct EpiCache {
    float* res = nullptr;        // shared-memory slot of 2 floats, or null
    float2* rope = nullptr;      // shared-memory slot, or null
    bool res_valid = false, rope_valid = false;
};

static __device__ void epilogue_pair(const GemvArgs& a, int64_t r0, float v0, float v1, unsigned long long& best, int pos_value = -1,
                                     const int* page_table_override = nullptr, bool ll_out = false, const EpiCache ec = EpiCache()) {
    if (a.bias) { v0 = __fadd_rn(v0, a.bias[r0]); v1 = __fadd_rn(v1, a.bias[r0 + 1]); }
    int epi = a.epi;
    if (epi == EPI_QKV_ROPE_GELU) {
        const int64_t nqkv = (int64_t)a.n_q + 2 * a.n_kv;
        if (r0 >= nqkv) { a.y3[r0 - nqkv] = gelu_f16_table(v0); a.y3[r0 - nqkv + 1] = gelu_f16_table(v1); return; }
        epi = EPI_QKV_ROPE;
    }
    switch (epi) {
    case EPI_STORE: a.y[r0] = v0; a.y[r0 + 1] = v1; break;
    case EPI_RESIDUAL: {
        const bool cached = ec.res && ec.res_valid;
        const float n0 = __fadd_rn(cached ? ec.res[0] : a.y[r0], v0), n1 = __fadd_rn(cached ? ec.res[1] : a.y[r0 + 1], v1);
        a.y[r0] = n0; a.y[r0 + 1] = n1;
        if (ec.res) { ec.res[0] = n0; ec.res[1] = n1; }
        if (ll_out) {                                            // the next stage of the persistent kernel polls these
            ll_store2(a.y_ll + r0, n0, n1);
            ll_store2(a.y_ll_poison + r0, __uint_as_float(LL_POISON), __uint_as_float(LL_POISON));
        }
    } break;
    case EPI_RESIDUAL2:
        a.y[r0] = __fadd_rn(__fadd_rn(v0, a.y2[r0]), a.y[r0]);
        a.y[r0 + 1] = __fadd_rn(__fadd_rn(v1, a.y2[r0 + 1]), a.y[r0 + 1]);
        break;
    case EPI_SILU_MUL: {
        const float h = __fmul_rn(__fdiv_rn(v0, __fadd_rn(1.0f, (float)exp((double)(-v0)))), v1);
        a.y[r0 >> 1] = h;
        if (ll_out) { ll_store(a.y_ll + (r0 >> 1), h); ll_store(a.y_ll_poison + (r0 >> 1), __uint_as_float(LL_POISON)); }
    } break;
    case EPI_GELU: a.y[r0] = gelu_f16_table(v0); a.y[r0 + 1] = gelu_f16_table(v1); break;
    case EPI_LOGITS: {
        a.y[r0] = v0; a.y[r0 + 1] = v1;
        const unsigned long long k0 = argmax_key(v0, (uint32_t)r0), k1 = argmax_key(v1, (uint32_t)r0 + 1);
        best = max(best, max(k0, k1));
    } break;
    case EPI_QKV_ROPE: {
        // Rows are [q | k | v].  Inside a head the rotation partners sit in adjacent rows: for
        // "norm" rope that is the file order; for NeoX the loader permuted the rows so that
        // stored row 2p = element p and 2p+1 = element p + n_rot/2 (p < n_rot/2).
        const int hd = a.head_dim;
        const bool is_q = r0 < a.n_q, is_k = !is_q && r0 < a.n_q + a.n_kv;
        const int64_t rr = is_q ? r0 : is_k ? r0 - a.n_q : r0 - a.n_q - a.n_kv;
        const int e = (int)(rr % hd);       // even position inside the head (stored order)
        const int64_t hb = rr - e;          // head base
        int i0 = e, i1 = e + 1;             // destination element indices inside the head
        float o0 = v0, o1 = v1;
        if (!is_q && !is_k) {
            if (a.rope_neox && e < a.n_rot) { i0 = e >> 1; i1 = (e >> 1) + a.n_rot / 2; }
        } else if (e < a.n_rot) {
            const float2 cs = ec.rope && ec.rope_valid ? *ec.rope : a.rope_cs[(size_t)(pos_value >= 0 ? pos_value : *a.kv.pos) * (a.n_rot >> 1) + (e >> 1)];
            if (ec.rope && !ec.rope_valid) *ec.rope = cs;
            o0 = __fsub_rn(__fmul_rn(v0, cs.x), __fmul_rn(v1, cs.y));
            o1 = __fadd_rn(__fmul_rn(v0, cs.y), __fmul_rn(v1, cs.x));
            if (a.rope_neox) { i0 = e >> 1; i1 = (e >> 1) + a.n_rot / 2; }
        }
        if (is_q) {
            const float q0 = __fmul_rn(o0, a.q_scale), q1 = __fmul_rn(o1, a.q_scale);
            a.y[hb + i0] = q0;
            a.y[hb + i1] = q1;
            if (ll_out) {           // (the host sets y_ll for this epilogue only when i1 == i0 + 1: "norm" RoPE)
                ll_store2(a.y_ll + hb + i0, q0, q1);
                ll_store2(a.y_ll_poison + hb + i0, __uint_as_float(LL_POISON), __uint_as_float(LL_POISON));
            }
        } else {
            const int pos = pos_value >= 0 ? pos_value : *a.kv.pos;
            const int page = (page_table_override ? page_table_override : a.kv.page_table)[pos / a.kv.page_tokens];
            const size_t base = ((size_t)page * a.kv.page_tokens + pos % a.kv.page_tokens) * a.kv.kv_dim;
            __half* dst = (is_k ? a.kv.k_base : a.kv.v_base) + base + hb;
            dst[i0] = __float2half_rn(o0);
            dst[i1] = __float2half_rn(o1);
            if (ll_out) {           // the same two halves as one word, where the attention stage polls for them
                const uint32_t w = (uint32_t)__half_as_ushort(__float2half_rn(o0)) | ((uint32_t)__half_as_ushort(__float2half_rn(o1)) << 16);
                const size_t wi = (size_t)a.n_q + (is_k ? 0 : (a.n_kv >> 1)) + ((hb + i0) >> 1);
                ll_store_u32(reinterpret_cast<uint32_t*>(a.y_ll) + wi, w);
                ll_store_u32(reinterpret_cast<uint32_t*>(a.y_ll_poison) + wi, LL_POISON);
            }
        }
    } break;
    default: break;
    }
}


}  // namespace gemvdev
}  // namespace b2

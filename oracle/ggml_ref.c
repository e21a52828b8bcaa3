/*
 * oracle/ggml_ref.c -- TEST INFRASTRUCTURE (see ggml_ref.h for scope and parity status).
 *
 * Block formats, scalar reference quantisers and generic-C dot products of ggml,
 * restated from the published formats.  Where the in-container `gguf` python
 * package restates the same routine, the line range is cited; routines that only
 * exist upstream (ggml-quants.c in llama.cpp, not under /root/reference) are
 * tagged [upstream].
 */
#include "ggml_ref.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef _Float16 f16_t;

static inline float h2f(uint16_t h) { f16_t v; memcpy(&v, &h, 2); return (float)v; }
static inline uint16_t f2h(float f) { f16_t v = (f16_t)f; uint16_t h; memcpy(&h, &v, 2); return h; }

/* ---- block layouts (gguf/constants.py:4235-4245 for the sizes) ------------------ */
#pragma pack(push, 1)
typedef struct { uint16_t d; uint8_t qs[16]; } blk_q4_0;                       /* 18 B / 32 */
typedef struct { uint16_t d; int8_t qs[32]; } blk_q8_0;                        /* 34 B / 32 */
typedef struct { uint16_t d, dmin; uint8_t scales[12]; uint8_t qs[128]; } blk_q4_K; /* 144 B / 256 */
typedef struct { uint8_t ql[128]; uint8_t qh[64]; int8_t scales[16]; uint16_t d; } blk_q6_K; /* 210 B / 256 */
typedef struct { float d; int8_t qs[256]; int16_t bsums[16]; } blk_q8_K;       /* 292 B / 256 */
#pragma pack(pop)

size_t ref_row_size(int type, int64_t k) {
    switch (type) {
    case REF_F32: return (size_t)k * 4;
    case REF_F16: return (size_t)k * 2;
    case REF_Q4_0: return k % 32 ? 0 : (size_t)(k / 32) * sizeof(blk_q4_0);
    case REF_Q8_0: return k % 32 ? 0 : (size_t)(k / 32) * sizeof(blk_q8_0);
    case REF_Q4_K: return k % 256 ? 0 : (size_t)(k / 256) * sizeof(blk_q4_K);
    case REF_Q6_K: return k % 256 ? 0 : (size_t)(k / 256) * sizeof(blk_q6_K);
    case REF_Q8_K: return k % 256 ? 0 : (size_t)(k / 256) * sizeof(blk_q8_K);
    }
    return 0;
}

int ref_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* A launcher may export OMP_NUM_THREADS=1 for its workers (torch.distributed.run does); the timed CPU
 * baseline sets the count it wants explicitly. */
void ref_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* ---- dequantisation ------------------------------------------------------------- */
/* Q4_0: w[j] = (qs[j] & 15) - 8, w[j+16] = (qs[j] >> 4) - 8, times d  (gguf/quants.py:241-251) */
static void dq_q4_0(const blk_q4_0 *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / 32; i++) {
        const float d = h2f(x[i].d);
        for (int j = 0; j < 16; j++) {
            y[i * 32 + j] = (float)((x[i].qs[j] & 0x0F) - 8) * d;
            y[i * 32 + j + 16] = (float)((x[i].qs[j] >> 4) - 8) * d;
        }
    }
}
/* Q8_0 (gguf/quants.py:396-401) */
static void dq_q8_0(const blk_q8_0 *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / 32; i++) {
        const float d = h2f(x[i].d);
        for (int j = 0; j < 32; j++) y[i * 32 + j] = (float)x[i].qs[j] * d;
    }
}
/* 6-bit (scale, min) pair j of a Q4_K block (gguf/quants.py:479-502) */
static inline void k4_scale_min(int j, const uint8_t *q, uint8_t *sc, uint8_t *mn) {
    if (j < 4) {
        *sc = q[j] & 63;
        *mn = q[j + 4] & 63;
    } else {
        *sc = (uint8_t)((q[j + 4] & 0x0F) | ((q[j - 4] >> 6) << 4));
        *mn = (uint8_t)((q[j + 4] >> 4) | ((q[j] >> 6) << 4));
    }
}
/* Q4_K: w = d*sc*q - dmin*m (gguf/quants.py:504-522) */
static void dq_q4_K(const blk_q4_K *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / 256; i++) {
        const float d = h2f(x[i].d), dmin = h2f(x[i].dmin);
        const uint8_t *q = x[i].qs;
        int is = 0;
        for (int j = 0; j < 256; j += 64) {
            uint8_t sc, m;
            k4_scale_min(is + 0, x[i].scales, &sc, &m);
            const float d1 = d * sc, m1 = dmin * m;
            k4_scale_min(is + 1, x[i].scales, &sc, &m);
            const float d2 = d * sc, m2 = dmin * m;
            for (int l = 0; l < 32; l++) *y++ = d1 * (float)(q[l] & 0x0F) - m1;
            for (int l = 0; l < 32; l++) *y++ = d2 * (float)(q[l] >> 4) - m2;
            q += 32;
            is += 2;
        }
    }
}
/* Q6_K (gguf/quants.py:554-572) */
static void dq_q6_K(const blk_q6_K *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / 256; i++) {
        const float d = h2f(x[i].d);
        const uint8_t *ql = x[i].ql, *qh = x[i].qh;
        const int8_t *sc = x[i].scales;
        for (int n = 0; n < 256; n += 128) {
            for (int l = 0; l < 32; l++) {
                const int is = l / 16;
                const int8_t q1 = (int8_t)((ql[l] & 0x0F) | (((qh[l] >> 0) & 3) << 4)) - 32;
                const int8_t q2 = (int8_t)((ql[l + 32] & 0x0F) | (((qh[l] >> 2) & 3) << 4)) - 32;
                const int8_t q3 = (int8_t)((ql[l] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                const int8_t q4 = (int8_t)((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
                /* gguf/quants.py multiplies (d*scale) first, then by q */
                y[l + 0] = (d * (float)sc[is + 0]) * (float)q1;
                y[l + 32] = (d * (float)sc[is + 2]) * (float)q2;
                y[l + 64] = (d * (float)sc[is + 4]) * (float)q3;
                y[l + 96] = (d * (float)sc[is + 6]) * (float)q4;
            }
            y += 128; ql += 64; qh += 32; sc += 8;
        }
    }
}

void ref_dequantize_row(int type, const void *b, float *y, int64_t k) {
    switch (type) {
    case REF_F32: memcpy(y, b, (size_t)k * 4); break;
    case REF_F16: for (int64_t i = 0; i < k; i++) y[i] = h2f(((const uint16_t *)b)[i]); break;
    case REF_Q4_0: dq_q4_0((const blk_q4_0 *)b, y, k); break;
    case REF_Q8_0: dq_q8_0((const blk_q8_0 *)b, y, k); break;
    case REF_Q4_K: dq_q4_K((const blk_q4_K *)b, y, k); break;
    case REF_Q6_K: dq_q6_K((const blk_q6_K *)b, y, k); break;
    default: break;
    }
}

/* ---- quantisation ---------------------------------------------------------------- */
/* quantize_row_q4_0_ref [upstream]; gguf/quants.py:222-239 */
void ref_quantize_row_q4_0(const float *x, void *vy, int64_t k) {
    blk_q4_0 *y = (blk_q4_0 *)vy;
    for (int64_t i = 0; i < k / 32; i++) {
        float amax = 0.0f, max = 0.0f;
        for (int j = 0; j < 32; j++) {
            const float v = x[i * 32 + j];
            if (amax < fabsf(v)) { amax = fabsf(v); max = v; }
        }
        const float d = max / -8;
        const float id = d ? 1.0f / d : 0.0f;
        y[i].d = f2h(d);
        for (int j = 0; j < 16; j++) {
            const float x0 = x[i * 32 + j] * id, x1 = x[i * 32 + 16 + j] * id;
            int a = (int8_t)(x0 + 8.5f), b = (int8_t)(x1 + 8.5f);
            if (a > 15) a = 15;
            if (b > 15) b = 15;
            y[i].qs[j] = (uint8_t)(a | (b << 4));
        }
    }
}
/* quantize_row_q8_0_ref [upstream]; gguf/quants.py:381-393 (round half away from zero) */
void ref_quantize_row_q8_0(const float *x, void *vy, int64_t k) {
    blk_q8_0 *y = (blk_q8_0 *)vy;
    for (int64_t i = 0; i < k / 32; i++) {
        float amax = 0.0f;
        for (int j = 0; j < 32; j++) { const float a = fabsf(x[i * 32 + j]); if (a > amax) amax = a; }
        const float d = amax / 127;
        const float id = d ? 1.0f / d : 0.0f;
        y[i].d = f2h(d);
        for (int j = 0; j < 32; j++) y[i].qs[j] = (int8_t)roundf(x[i * 32 + j] * id);
    }
}
/* round-to-nearest-even via the 1.5*2^23 trick, as ggml's nearest_int [upstream] */
static inline int nearest_int(float f) {
    float v = f + 12582912.f;
    int i; memcpy(&i, &v, 4);
    return (i & 0x007fffff) - 0x00400000;
}
/* quantize_row_q8_K_ref [upstream]: per 256, f32 scale, iscale = -127/max, 16 partial sums */
void ref_quantize_row_q8_K(const float *x, void *vy, int64_t k) {
    blk_q8_K *y = (blk_q8_K *)vy;
    for (int64_t i = 0; i < k / 256; i++) {
        float max = 0, amax = 0;
        for (int j = 0; j < 256; j++) {
            const float ax = fabsf(x[j]);
            if (ax > amax) { amax = ax; max = x[j]; }
        }
        if (!amax) {
            y[i].d = 0;
            memset(y[i].qs, 0, 256);
            memset(y[i].bsums, 0, sizeof(y[i].bsums));
            x += 256;
            continue;
        }
        const float iscale = -127.f / max;
        for (int j = 0; j < 256; j++) {
            int v = nearest_int(iscale * x[j]);
            y[i].qs[j] = (int8_t)(v > 127 ? 127 : v);
        }
        for (int j = 0; j < 16; j++) {
            int s = 0;
            for (int l = 0; l < 16; l++) s += y[i].qs[j * 16 + l];
            y[i].bsums[j] = (int16_t)s;
        }
        y[i].d = 1 / iscale;
        x += 256;
    }
}

/* Synthetic Q4_K maker (not ggml's rmse search): per 32 sub-block affine
 * min/max -> 4 bit, then 6-bit scale/min against per-256 f16 d/dmin. */
void ref_make_q4_K(const float *x, void *vy, int64_t k) {
    blk_q4_K *y = (blk_q4_K *)vy;
    for (int64_t i = 0; i < k / 256; i++, x += 256) {
        float sc[8], mn[8], smax = 0, mmax = 0;
        for (int j = 0; j < 8; j++) {
            float lo = x[j * 32], hi = lo;
            for (int l = 1; l < 32; l++) { float v = x[j * 32 + l]; if (v < lo) lo = v; if (v > hi) hi = v; }
            if (lo > 0) lo = 0;           /* min is stored as a non-negative offset */
            sc[j] = (hi - lo) / 15.f; mn[j] = -lo;
            if (sc[j] > smax) smax = sc[j];
            if (mn[j] > mmax) mmax = mn[j];
        }
        const float d = smax / 63.f, dm = mmax / 63.f;
        y[i].d = f2h(d); y[i].dmin = f2h(dm);
        const float dq = h2f(y[i].d), dmq = h2f(y[i].dmin);
        uint8_t ls[8], lm[8];
        memset(y[i].scales, 0, 12);
        for (int j = 0; j < 8; j++) {
            int a = d > 0 ? nearest_int(sc[j] / d) : 0, b = dm > 0 ? nearest_int(mn[j] / dm) : 0;
            ls[j] = (uint8_t)(a > 63 ? 63 : a); lm[j] = (uint8_t)(b > 63 ? 63 : b);
            if (j < 4) { y[i].scales[j] = ls[j]; y[i].scales[j + 4] = lm[j]; }
            else {
                y[i].scales[j + 4] = (uint8_t)((ls[j] & 0xF) | ((lm[j] & 0xF) << 4));
                y[i].scales[j - 4] |= (uint8_t)((ls[j] >> 4) << 6);
                y[i].scales[j] |= (uint8_t)((lm[j] >> 4) << 6);
            }
        }
        uint8_t L[256];
        for (int j = 0; j < 8; j++) {
            const float dd = dq * ls[j], mm = dmq * lm[j];
            for (int l = 0; l < 32; l++) {
                int q = dd > 0 ? nearest_int((x[j * 32 + l] + mm) / dd) : 0;
                L[j * 32 + l] = (uint8_t)(q < 0 ? 0 : q > 15 ? 15 : q);
            }
        }
        uint8_t *q = y[i].qs;
        for (int j = 0; j < 256; j += 64) {
            for (int l = 0; l < 32; l++) q[l] = (uint8_t)(L[j + l] | (L[j + l + 32] << 4));
            q += 32;
        }
    }
}
/* Synthetic Q6_K maker: per 16 sub-block max-abs/32 scale -> int8 scales against f16 d. */
void ref_make_q6_K(const float *x, void *vy, int64_t k) {
    blk_q6_K *y = (blk_q6_K *)vy;
    for (int64_t i = 0; i < k / 256; i++, x += 256) {
        float sc[16], amax_s = 0, max_s = 0;
        for (int j = 0; j < 16; j++) {
            float amax = 0, mx = 0;
            for (int l = 0; l < 16; l++) { float v = x[j * 16 + l]; if (fabsf(v) > amax) { amax = fabsf(v); mx = v; } }
            sc[j] = mx / -32.f;
            if (fabsf(sc[j]) > amax_s) { amax_s = fabsf(sc[j]); max_s = sc[j]; }
        }
        if (amax_s == 0) { memset(&y[i], 0, sizeof(blk_q6_K)); continue; }
        const float iscale = -128.f / max_s;
        y[i].d = f2h(1 / iscale);
        const float d = h2f(y[i].d);
        int8_t L[256];
        for (int j = 0; j < 16; j++) {
            int s = nearest_int(iscale * sc[j]);
            y[i].scales[j] = (int8_t)(s > 127 ? 127 : s);
            const float dd = d * y[i].scales[j];
            for (int l = 0; l < 16; l++) {
                int q = dd != 0 ? nearest_int(x[j * 16 + l] / dd) : 0;
                q = q < -32 ? -32 : q > 31 ? 31 : q;
                L[j * 16 + l] = (int8_t)(q + 32);
            }
        }
        uint8_t *ql = y[i].ql, *qh = y[i].qh;
        for (int n = 0; n < 256; n += 128) {
            for (int l = 0; l < 32; l++) {
                const uint8_t q1 = L[n + l] & 0xF, q2 = L[n + l + 32] & 0xF, q3 = L[n + l + 64] & 0xF, q4 = L[n + l + 96] & 0xF;
                ql[l] = (uint8_t)(q1 | (q3 << 4));
                ql[l + 32] = (uint8_t)(q2 | (q4 << 4));
                qh[l] = (uint8_t)((L[n + l] >> 4) | ((L[n + l + 32] >> 4) << 2) | ((L[n + l + 64] >> 4) << 4) | ((L[n + l + 96] >> 4) << 6));
            }
            ql += 64; qh += 32;
        }
    }
}

/* ---- dot products ------------------------------------------------------------------ */
/* The integer part of every block dot is exact whatever the evaluation order, so the
 * AVX2 helper below (compiled in by `make SIMD=1`) only accelerates it; the float
 * accumulation keeps the order of ggml's generic C path. */
#if defined(__x86_64__) && defined(REF_SIMD)
#include <immintrin.h>
/* eight int32 partial sums of one block: sum over the vector == the block's sumi */
static inline __m256i q4_0_q8_0_p32(const blk_q4_0 *x, const blk_q8_0 *y) {
    const __m128i raw = _mm_loadu_si128((const __m128i *)x->qs);
    const __m256i nib = _mm256_and_si256(_mm256_set_m128i(_mm_srli_epi16(raw, 4), raw), _mm256_set1_epi8(0x0F));
    const __m256i w = _mm256_sub_epi8(nib, _mm256_set1_epi8(8));
    const __m256i a = _mm256_loadu_si256((const __m256i *)y->qs);
    const __m256i p16 = _mm256_maddubs_epi16(_mm256_sign_epi8(w, w), _mm256_sign_epi8(a, w));   /* |w| * (a*sign w) */
    return _mm256_madd_epi16(p16, _mm256_set1_epi16(1));
}
#define REF_HAVE_SIMD 1
#else
#define REF_HAVE_SIMD 0
#endif

/* ggml_vec_dot_q4_0_q8_0, generic path [upstream]: sumf += sumi * d_w * d_a, block by block */
float ref_vec_dot_q4_0_q8_0(int64_t n, const void *vx, const void *vy) {
    const blk_q4_0 *x = (const blk_q4_0 *)vx;
    const blk_q8_0 *y = (const blk_q8_0 *)vy;
    float sumf = 0;
    const int64_t nb = n / 32;
    int64_t i = 0;
#if REF_HAVE_SIMD
    /* eight blocks at a time: the integer sums are reduced with a hadd tree (exact), the
     * per-block float terms are formed with the same two roundings as the scalar code and
     * then added to sumf one after the other, in block order. */
    for (; i + 8 <= nb; i += 8) {
        __m256i p[8];
        for (int j = 0; j < 8; j++) p[j] = q4_0_q8_0_p32(&x[i + j], &y[i + j]);
        const __m256i h01 = _mm256_hadd_epi32(p[0], p[1]), h23 = _mm256_hadd_epi32(p[2], p[3]);
        const __m256i h45 = _mm256_hadd_epi32(p[4], p[5]), h67 = _mm256_hadd_epi32(p[6], p[7]);
        const __m256i h0123 = _mm256_hadd_epi32(h01, h23), h4567 = _mm256_hadd_epi32(h45, h67);
        /* lanes: lo128 = partial(0..3 | 4..7 of low halves), hi128 = same of high halves */
        const __m128i s0123 = _mm_add_epi32(_mm256_castsi256_si128(h0123), _mm256_extracti128_si256(h0123, 1));
        const __m128i s4567 = _mm_add_epi32(_mm256_castsi256_si128(h4567), _mm256_extracti128_si256(h4567, 1));
        float si[8], dx[8], dy[8];
        _mm_storeu_ps(si, _mm_cvtepi32_ps(s0123));
        _mm_storeu_ps(si + 4, _mm_cvtepi32_ps(s4567));
        for (int j = 0; j < 8; j++) { dx[j] = h2f(x[i + j].d); dy[j] = h2f(y[i + j].d); }
        for (int j = 0; j < 8; j++) sumf += si[j] * dx[j] * dy[j];
    }
#endif
    for (; i < nb; i++) {
        int sumi = 0;
        for (int j = 0; j < 16; j++) {
            const int v0 = (x[i].qs[j] & 0x0F) - 8, v1 = (x[i].qs[j] >> 4) - 8;
            sumi += v0 * y[i].qs[j] + v1 * y[i].qs[j + 16];
        }
        sumf += sumi * h2f(x[i].d) * h2f(y[i].d);
    }
    return sumf;
}
/* ggml_vec_dot_q8_0_q8_0, generic path [upstream] */
float ref_vec_dot_q8_0_q8_0(int64_t n, const void *vx, const void *vy) {
    const blk_q8_0 *x = (const blk_q8_0 *)vx, *y = (const blk_q8_0 *)vy;
    float sumf = 0;
    for (int64_t i = 0; i < n / 32; i++) {
        int sumi = 0;
        for (int j = 0; j < 32; j++) sumi += x[i].qs[j] * y[i].qs[j];
        sumf += sumi * (h2f(x[i].d) * h2f(y[i].d));
    }
    return sumf;
}
/* ggml_vec_dot_q4_K_q8_K, generic path [upstream]: 8 float lanes, mins folded via bsums */
float ref_vec_dot_q4_K_q8_K(int64_t n, const void *vx, const void *vy) {
    const blk_q4_K *x = (const blk_q4_K *)vx;
    const blk_q8_K *y = (const blk_q8_K *)vy;
    float sums[8] = {0}, sumf = 0;
    int8_t a8[256];
    for (int64_t i = 0; i < n / 256; i++) {
        const uint8_t *q4 = x[i].qs;
        const int8_t *q8 = y[i].qs;
        int32_t acc[8] = {0};
        int8_t *a = a8;
        for (int j = 0; j < 4; j++) {
            for (int l = 0; l < 32; l++) a[l] = (int8_t)(q4[l] & 0xF);
            a += 32;
            for (int l = 0; l < 32; l++) a[l] = (int8_t)(q4[l] >> 4);
            a += 32; q4 += 32;
        }
        uint8_t sc[8], mn[8];
        for (int j = 0; j < 8; j++) k4_scale_min(j, x[i].scales, &sc[j], &mn[j]);
        int sumi = 0;
        for (int j = 0; j < 16; j++) sumi += y[i].bsums[j] * mn[j / 2];
        a = a8;
        for (int j = 0; j < 8; j++) {
            const int32_t s = sc[j];
            for (int r = 0; r < 4; r++) {
                for (int l = 0; l < 8; l++) acc[l] += s * (int16_t)(q8[l] * a[l]);
                q8 += 8; a += 8;
            }
        }
        const float d = h2f(x[i].d) * y[i].d;
        for (int l = 0; l < 8; l++) sums[l] += d * acc[l];
        const float dmin = h2f(x[i].dmin) * y[i].d;
        sumf -= dmin * sumi;
    }
    for (int l = 0; l < 8; l++) sumf += sums[l];
    return sumf;
}
/* ggml_vec_dot_q6_K_q8_K, generic path [upstream] */
float ref_vec_dot_q6_K_q8_K(int64_t n, const void *vx, const void *vy) {
    const blk_q6_K *x = (const blk_q6_K *)vx;
    const blk_q8_K *y = (const blk_q8_K *)vy;
    float sums[8] = {0};
    int8_t a8[256];
    for (int64_t i = 0; i < n / 256; i++) {
        const uint8_t *q4 = x[i].ql, *qh = x[i].qh;
        const int8_t *q8 = y[i].qs;
        int32_t acc[8] = {0};
        int8_t *a = a8;
        for (int j = 0; j < 256; j += 128) {
            for (int l = 0; l < 32; l++) {
                a[l + 0] = (int8_t)((q4[l + 0] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                a[l + 32] = (int8_t)((q4[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                a[l + 64] = (int8_t)((q4[l + 0] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                a[l + 96] = (int8_t)((q4[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
            }
            a += 128; q4 += 64; qh += 32;
        }
        a = a8;
        for (int j = 0; j < 16; j++) {
            const int32_t s = x[i].scales[j];
            for (int r = 0; r < 2; r++) {
                for (int l = 0; l < 8; l++) acc[l] += s * (int16_t)(q8[l] * a[l]);
                q8 += 8; a += 8;
            }
        }
        const float d = h2f(x[i].d) * y[i].d;
        for (int l = 0; l < 8; l++) sums[l] += d * acc[l];
    }
    float sumf = 0;
    for (int l = 0; l < 8; l++) sumf += sums[l];
    return sumf;
}


/* ---- "canonical" (order-free) dots ------------------------------------------------------
 * ggml adds the per-block float terms in f32, in an order that depends on the ISA path (generic
 * C: block after block; AVX2: 8 lanes; NEON: two blocks at a time ...).  Rounding differences of
 * 1e-7 between two such orders are enough to flip int8 roundings of the NEXT activation
 * quantisation, after which two runs of the same model differ by ~1e-1 in the logits.  To have
 * something a massively parallel kernel can reproduce bit for bit, mode 2 keeps every term of
 * the generic path (exact integer block dot, same float scale products) but adds the terms in
 * f64, where the order no longer matters, and rounds the row sum to f32 once.  K-quants use
 * one term per half super-block (128 weights) with the exact integer total. */
static double canon_dot_q4_0(int64_t n, const blk_q4_0 *x, const blk_q8_0 *y) {
    double acc = 0;
    for (int64_t i = 0; i < n / 32; i++) {
        int sumi = 0;
        for (int j = 0; j < 16; j++) {
            const int v0 = (x[i].qs[j] & 0x0F) - 8, v1 = (x[i].qs[j] >> 4) - 8;
            sumi += v0 * y[i].qs[j] + v1 * y[i].qs[j + 16];
        }
        acc += (double)(((float)sumi * h2f(x[i].d)) * h2f(y[i].d));
    }
    return acc;
}
static double canon_dot_q8_0(int64_t n, const blk_q8_0 *x, const blk_q8_0 *y) {
    double acc = 0;
    for (int64_t i = 0; i < n / 32; i++) {
        int sumi = 0;
        for (int j = 0; j < 32; j++) sumi += x[i].qs[j] * y[i].qs[j];
        acc += (double)((float)sumi * (h2f(x[i].d) * h2f(y[i].d)));
    }
    return acc;
}
static double canon_dot_q6_K(int64_t n, const blk_q6_K *x, const blk_q8_K *y) {
    double acc = 0;
    for (int64_t i = 0; i < n / 256; i++) {
        const float d = h2f(x[i].d) * y[i].d;
        for (int half = 0; half < 2; half++) {
            const uint8_t *ql = x[i].ql + 64 * half, *qh = x[i].qh + 32 * half;
            const int8_t *sc = x[i].scales + 8 * half, *q8 = y[i].qs + 128 * half;
            int tot = 0;
            for (int l = 0; l < 32; l++) {
                const int is = l / 16;
                const int q1 = (int)((ql[l] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                const int q2 = (int)((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                const int q3 = (int)((ql[l] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                const int q4 = (int)((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
                tot += sc[is] * q1 * q8[l] + sc[is + 2] * q2 * q8[l + 32] + sc[is + 4] * q3 * q8[l + 64] + sc[is + 6] * q4 * q8[l + 96];
            }
            acc += (double)(d * (float)tot);
        }
    }
    return acc;
}
static double canon_dot_q4_K(int64_t n, const blk_q4_K *x, const blk_q8_K *y) {
    double acc = 0;
    for (int64_t i = 0; i < n / 256; i++) {
        const float d = h2f(x[i].d) * y[i].d, dmin = h2f(x[i].dmin) * y[i].d;
        for (int half = 0; half < 2; half++) {
            int accs = 0, accm = 0;
            for (int j = 0; j < 4; j++) {               /* sub-block 4*half + j: 32 weights */
                const int sb = 4 * half + j;
                uint8_t sc, mn;
                k4_scale_min(sb, x[i].scales, &sc, &mn);
                const uint8_t *q = x[i].qs + 32 * (sb / 2);
                const int8_t *q8 = y[i].qs + 32 * sb;
                int dot = 0, bs = 0;
                for (int l = 0; l < 32; l++) {
                    const int w = (sb & 1) ? (q[l] >> 4) : (q[l] & 0xF);
                    dot += w * q8[l];
                    bs += q8[l];
                }
                accs += sc * dot;
                accm += mn * bs;
            }
            acc += (double)(d * (float)accs - dmin * (float)accm);
        }
    }
    return acc;
}

/* ---- mode 3: the SAME arithmetic in the order of ggml's AVX2 kernels [upstream ggml_vec_dot_q4_0_q8_0 /
 * ggml_vec_dot_q8_0_q8_0, AVX2 branch]: per block d = d_w * d_a in f32, the 32 byte products folded into eight
 * int32 lanes (maddubs + madd: lane l = bytes 4l..4l+3 of [low nibbles | high nibbles]), acc[l] = fma(d, lane, acc[l]),
 * and hsum_float_8 at the end.  It exists to measure how far two legitimate ggml builds drift apart
 * (the noise floor any other implementation is judged against); nothing is pinned to it. */
static float avx2_hsum8(const float a[8]) {
    const float r0 = a[0] + a[4], r1 = a[1] + a[5], r2 = a[2] + a[6], r3 = a[3] + a[7];
    return (r0 + r2) + (r1 + r3);
}
static float avx2order_dot_q4_0(int64_t n, const blk_q4_0 *x, const blk_q8_0 *y) {
    float acc[8] = {0};
    for (int64_t i = 0; i < n / 32; i++) {
        const float d = h2f(x[i].d) * h2f(y[i].d);
        for (int l = 0; l < 8; l++) {
            int s = 0;
            for (int j = 0; j < 4; j++) {
                const int b = 4 * l + j;                    /* byte of the 32-byte vector [lo nibbles | hi nibbles] */
                const int w = b < 16 ? (x[i].qs[b] & 0x0F) - 8 : (x[i].qs[b - 16] >> 4) - 8;
                s += w * y[i].qs[b];
            }
            acc[l] = fmaf(d, (float)s, acc[l]);
        }
    }
    return avx2_hsum8(acc);
}
static float avx2order_dot_q8_0(int64_t n, const blk_q8_0 *x, const blk_q8_0 *y) {
    float acc[8] = {0};
    for (int64_t i = 0; i < n / 32; i++) {
        const float d = h2f(x[i].d) * h2f(y[i].d);
        for (int l = 0; l < 8; l++) {
            int s = 0;
            for (int j = 0; j < 4; j++) s += x[i].qs[4 * l + j] * y[i].qs[4 * l + j];
            acc[l] = fmaf(d, (float)s, acc[l]);
        }
    }
    return avx2_hsum8(acc);
}

/* vec_dot_type of each weight type (ggml type_traits [upstream]) */
static int act_type_for(int wtype) {
    switch (wtype) {
    case REF_Q4_0: case REF_Q8_0: return REF_Q8_0;
    case REF_Q4_K: case REF_Q6_K: return REF_Q8_K;
    case REF_F16: return REF_F16;
    default: return REF_F32;
    }
}

void ref_mul_mat_vec(int wtype, const void *W, int64_t rows, int64_t k, const float *x, float *y,
                     int faithful, int threads) {
    const size_t rs = ref_row_size(wtype, k);
    const char *Wb = (const char *)W;
#ifdef _OPENMP
    const int nt = threads > 0 ? threads : omp_get_max_threads();
#else
    const int nt = 1; (void)threads;
#endif
    (void)nt;
    if (!faithful || wtype == REF_F32) {
#pragma omp parallel num_threads(nt)
        {
            float *w = (float *)malloc((size_t)k * sizeof(float));
#pragma omp for schedule(static)
            for (int64_t r = 0; r < rows; r++) {
                ref_dequantize_row(wtype, Wb + r * rs, w, k);
                double s = 0;
                for (int64_t i = 0; i < k; i++) s += (double)w[i] * (double)x[i];
                y[r] = (float)s;
            }
            free(w);
        }
        return;
    }
    const int at = act_type_for(wtype);
    void *xq = malloc(at == REF_F16 ? (size_t)k * 2 : ref_row_size(at, k));
    if (at == REF_Q8_0) ref_quantize_row_q8_0(x, xq, k);
    else if (at == REF_Q8_K) ref_quantize_row_q8_K(x, xq, k);
    else for (int64_t i = 0; i < k; i++) ((uint16_t *)xq)[i] = f2h(x[i]);
#pragma omp parallel for schedule(static) num_threads(nt)
    for (int64_t r = 0; r < rows; r++) {
        const void *wr = Wb + r * rs;
        if (faithful == 2 && wtype != REF_F16) {
            switch (wtype) {
            case REF_Q4_0: y[r] = (float)canon_dot_q4_0(k, (const blk_q4_0 *)wr, (const blk_q8_0 *)xq); break;
            case REF_Q8_0: y[r] = (float)canon_dot_q8_0(k, (const blk_q8_0 *)wr, (const blk_q8_0 *)xq); break;
            case REF_Q4_K: y[r] = (float)canon_dot_q4_K(k, (const blk_q4_K *)wr, (const blk_q8_K *)xq); break;
            case REF_Q6_K: y[r] = (float)canon_dot_q6_K(k, (const blk_q6_K *)wr, (const blk_q8_K *)xq); break;
            default: y[r] = 0;
            }
            continue;
        }
        if (faithful == 3 && (wtype == REF_Q4_0 || wtype == REF_Q8_0)) {
            y[r] = wtype == REF_Q4_0 ? avx2order_dot_q4_0(k, (const blk_q4_0 *)wr, (const blk_q8_0 *)xq)
                                     : avx2order_dot_q8_0(k, (const blk_q8_0 *)wr, (const blk_q8_0 *)xq);
            continue;
        }
        switch (wtype) {
        case REF_Q4_0: y[r] = ref_vec_dot_q4_0_q8_0(k, wr, xq); break;
        case REF_Q8_0: y[r] = ref_vec_dot_q8_0_q8_0(k, wr, xq); break;
        case REF_Q4_K: y[r] = ref_vec_dot_q4_K_q8_K(k, wr, xq); break;
        case REF_Q6_K: y[r] = ref_vec_dot_q6_K_q8_K(k, wr, xq); break;
        case REF_F16: { /* ggml_vec_dot_f16 scalar form [upstream]: f32 product, f64 sum */
            double s = 0;
            const uint16_t *a = (const uint16_t *)wr, *b = (const uint16_t *)xq;
            for (int64_t i = 0; i < k; i++) s += (double)(h2f(a[i]) * h2f(b[i]));
            y[r] = (float)s;
        } break;
        default: y[r] = 0;
        }
    }
    free(xq);
}

/* ---- element ops -------------------------------------------------------------------- */
/* ggml rms_norm [upstream]: f64 sum of f32 squares, scale = 1/sqrtf(mean+eps); then * gain */
void ref_rms_norm(const float *x, const float *g, float *y, int64_t n, float eps) {
    double sum = 0;
    for (int64_t i = 0; i < n; i++) sum += (double)(x[i] * x[i]);
    const float mean = (float)(sum / n);
    const float scale = 1.0f / sqrtf(mean + eps);
    for (int64_t i = 0; i < n; i++) y[i] = (x[i] * scale) * g[i];
}
/* ggml norm [upstream]: mean, centred variance in f64, then gain and bias */
void ref_layer_norm(const float *x, const float *g, const float *b, float *y, int64_t n, float eps) {
    double sum = 0;
    for (int64_t i = 0; i < n; i++) sum += (double)x[i];
    const float mean = (float)(sum / n);
    double sum2 = 0;
    for (int64_t i = 0; i < n; i++) { const float v = x[i] - mean; y[i] = v; sum2 += (double)(v * v); }
    const float variance = (float)(sum2 / n);
    const float scale = 1.0f / sqrtf(variance + eps);
    for (int64_t i = 0; i < n; i++) y[i] = (y[i] * scale) * g[i] + (b ? b[i] : 0.0f);
}
/* ggml rope [upstream]: theta starts at pos and is multiplied by base^(-2/n_rot) per pair */
void ref_rope_ext(float *x, int n_head, int head_dim, int n_rot, int pos, float freq_base, int mode, float freq_scale, const float *factors) {
    /* ggml_rope_cache_init / rope_yarn with ext_factor = 0, attn_factor = 1 [upstream]: theta_base starts at pos and
     * is multiplied by base^(-2/n_rot) per pair; the angle used is freq_scale * (theta_base / freq_factor[pair]) */
    const float theta_scale = powf(freq_base, -2.0f / n_rot);
    if (freq_scale == 0.0f) freq_scale = 1.0f;
    for (int h = 0; h < n_head; h++) {
        float *v = x + (size_t)h * head_dim;
        float theta_base = (float)pos;
        for (int i0 = 0; i0 < n_rot; i0 += 2) {
            const float ff = factors ? factors[i0 / 2] : 1.0f;
            const float theta = freq_scale * (theta_base / ff);
            const float c = cosf(theta), s = sinf(theta);
            theta_base *= theta_scale;
            if (mode == 0) {
                const float a = v[i0], b = v[i0 + 1];
                v[i0] = a * c - b * s;
                v[i0 + 1] = a * s + b * c;
            } else {
                const int i = i0 / 2;
                const float a = v[i], b = v[i + n_rot / 2];
                v[i] = a * c - b * s;
                v[i + n_rot / 2] = a * s + b * c;
            }
        }
    }
}
void ref_rope(float *x, int n_head, int head_dim, int n_rot, int pos, float freq_base, int mode) {
    ref_rope_ext(x, n_head, head_dim, n_rot, pos, freq_base, mode, 1.0f, NULL);
}
/* ggml soft_max [upstream]: max, expf(x-max), f64 sum, scale by 1/sum */
void ref_softmax(float *x, int64_t n) {
    float mx = -INFINITY;
    for (int64_t i = 0; i < n; i++) if (x[i] > mx) mx = x[i];
    double sum = 0;
    for (int64_t i = 0; i < n; i++) { x[i] = expf(x[i] - mx); sum += (double)x[i]; }
    const float inv = (float)(1.0 / sum);
    for (int64_t i = 0; i < n; i++) x[i] *= inv;
}
/* mode 2 evaluates exp / tanh in f64 and rounds once: a correctly rounded f32 result, which is
 * what two different math libraries (glibc here, CUDA on the device) can agree on bit for bit */
static void softmax_canon(float *x, int64_t n) {
    float mx = -INFINITY;
    for (int64_t i = 0; i < n; i++) if (x[i] > mx) mx = x[i];
    double sum = 0;
    for (int64_t i = 0; i < n; i++) { x[i] = (float)exp((double)(x[i] - mx)); sum += (double)x[i]; }
    const float inv = (float)(1.0 / sum);
    for (int64_t i = 0; i < n; i++) x[i] *= inv;
}
static inline float silu_f(float v, int mode) {
    return mode == 2 ? v / (1.0f + (float)exp((double)(-v))) : v / (1.0f + expf(-v));
}
/* ggml gelu [upstream]: tanh form looked up through an f16 table => input and output rounded to f16 */
static inline float gelu_f(float v, int faithful) {
    const float A = 0.044715f, S = 0.79788456080286535587989211986876f;
    if (!faithful) return 0.5f * v * (1.0f + tanhf(S * v * (1.0f + A * v * v)));
    if (v <= -10.0f) return 0.0f;
    if (v >= 10.0f) return v;
    const float t = h2f(f2h(v));
    const float u = S * t * (1.0f + A * t * t);
    const float th = faithful == 2 ? (float)tanh((double)u) : tanhf(u);
    return h2f(f2h(0.5f * t * (1.0f + th)));
}

/* ---- forward pass ---------------------------------------------------------------------- */
struct ref_ctx {
    const ref_model *m;
    int faithful, threads;
    uint16_t **k16, **v16;     /* faithful: per layer [n_ctx][n_kv] f16 */
    float **k32, **v32;        /* ideal: f32 */
    float *x, *xn, *q, *k, *v, *att, *attn_out, *proj, *gate, *up, *ffn, *hidden, *scores;
};

static void *xcalloc(size_t n, size_t s) { void *p = calloc(n ? n : 1, s); if (!p) abort(); return p; }

ref_ctx *ref_ctx_new(const ref_model *m, int faithful, int threads) {
    ref_ctx *c = (ref_ctx *)xcalloc(1, sizeof(*c));
    c->m = m; c->faithful = faithful; c->threads = threads;
    const size_t nkv = (size_t)m->n_head_kv * m->head_dim, E = m->n_embd;
    if (faithful) { c->k16 = xcalloc(m->n_layer, sizeof(void *)); c->v16 = xcalloc(m->n_layer, sizeof(void *)); }
    else { c->k32 = xcalloc(m->n_layer, sizeof(void *)); c->v32 = xcalloc(m->n_layer, sizeof(void *)); }
    for (int l = 0; l < m->n_layer; l++) {
        if (faithful) { c->k16[l] = xcalloc(nkv * m->n_ctx, 2); c->v16[l] = xcalloc(nkv * m->n_ctx, 2); }
        else { c->k32[l] = xcalloc(nkv * m->n_ctx, 4); c->v32[l] = xcalloc(nkv * m->n_ctx, 4); }
    }
    const size_t qdim = (size_t)m->n_head * m->head_dim;
    c->x = xcalloc(E, 4); c->xn = xcalloc(E, 4); c->hidden = xcalloc(E, 4);
    c->q = xcalloc(qdim + 2 * nkv, 4); c->k = xcalloc(nkv, 4); c->v = xcalloc(nkv, 4);
    c->att = xcalloc(qdim, 4); c->attn_out = xcalloc(E, 4); c->proj = xcalloc(E, 4);
    c->gate = xcalloc(m->n_ff, 4); c->up = xcalloc(m->n_ff, 4); c->ffn = xcalloc(E, 4);
    c->scores = xcalloc((size_t)m->n_head * m->n_ctx, 4);
    return c;
}
void ref_ctx_free(ref_ctx *c) {
    if (!c) return;
    for (int l = 0; l < c->m->n_layer; l++) {
        if (c->k16) { free(c->k16[l]); free(c->v16[l]); }
        if (c->k32) { free(c->k32[l]); free(c->v32[l]); }
    }
    free(c->k16); free(c->v16); free(c->k32); free(c->v32);
    free(c->x); free(c->xn); free(c->hidden); free(c->q); free(c->k); free(c->v); free(c->att);
    free(c->attn_out); free(c->proj); free(c->gate); free(c->up); free(c->ffn); free(c->scores);
    free(c);
}
void ref_ctx_set_kv_f16(ref_ctx *c, int il, const uint16_t *k, const uint16_t *v, int n_tok) {
    const size_t nkv = (size_t)c->m->n_head_kv * c->m->head_dim;
    if (c->faithful) {
        memcpy(c->k16[il], k, nkv * n_tok * 2);
        memcpy(c->v16[il], v, nkv * n_tok * 2);
    } else {
        for (size_t i = 0; i < nkv * n_tok; i++) { c->k32[il][i] = h2f(k[i]); c->v32[il][i] = h2f(v[i]); }
    }
}
const float *ref_last_hidden(const ref_ctx *c) { return c->hidden; }

static void matvec(const ref_ctx *c, const ref_tensor *t, int64_t rows, int64_t k, const float *x, float *y) {
    ref_mul_mat_vec(t->type, t->data, rows, k, x, y, c->faithful, c->threads);
}
static void add_bias(float *y, const ref_tensor *b, int64_t n) {
    if (!b->data) return;
    const float *bb = (const float *)b->data;
    for (int64_t i = 0; i < n; i++) y[i] += bb[i];
}

/* ggml_vec_dot_f16 as ggml's 8-wide SIMD builds (AVX/AVX2/NEON-f32 style, GGML_F16_STEP = 32)
 * evaluate it [upstream]: 32 f32 lanes L = i mod 32, each an FMA chain over the elements it owns,
 * then the fixed reduction  (L,L+16) -> (L,L+8) -> (L,L+4) -> hadd -> hadd ; elements beyond the
 * last multiple of 32 are added in f64.  `stride*` let the same code walk a K row (stride 1) and
 * a V column (stride = row length).  The result is what the x86 servers the reference deploys on
 * compute, and -- unlike an f64 running sum over hundreds of cached positions -- it is cheap to
 * reproduce bit for bit on a 32-lane GPU warp. */
static float dot_f16_lanes32(int64_t n, const uint16_t *x, int64_t stride_x, const uint16_t *y, int64_t stride_y, int64_t n_simd) {
    float lane[32];
    for (int l = 0; l < 32; l++) lane[l] = 0.0f;
    for (int64_t i = 0; i < n_simd; i++)
        lane[i & 31] = fmaf(h2f(x[i * stride_x]), h2f(y[i * stride_y]), lane[i & 31]);
    for (int l = 0; l < 16; l++) lane[l] += lane[l + 16];
    for (int l = 0; l < 8; l++) lane[l] += lane[l + 8];
    for (int l = 0; l < 4; l++) lane[l] += lane[l + 4];
    const float res = (lane[0] + lane[1]) + (lane[2] + lane[3]);
    double sumf = (double)res;
    for (int64_t i = n_simd; i < n; i++) sumf += (double)(h2f(x[i * stride_x]) * h2f(y[i * stride_y]));
    return (float)sumf;
}

/* Attention for one new token over cache rows [0,pos]; llama.cpp's non-flash path
 * [upstream]: kq = K(f16).q(->f16), f32 result; soft_max_ext(kq*scale); probabilities
 * ->f16; out = V(f16)^T.p with f64 accumulate of f32 products. */
static void attend(ref_ctx *c, int il, int pos, float scale) {
    const ref_model *m = c->m;
    const int hd = m->head_dim, grp = m->n_head / m->n_head_kv;
    const size_t nkv = (size_t)m->n_head_kv * hd;
#pragma omp parallel for schedule(static) num_threads(c->threads > 0 ? c->threads : ref_num_threads())
    for (int h = 0; h < m->n_head; h++) {
        const int hk = h / grp;
        float *sc = c->scores + (size_t)h * m->n_ctx;
        const float *q = c->q + (size_t)h * hd;
        float *o = c->att + (size_t)h * hd;
        if (c->faithful == 2) {
            /* canonical: both dots in the 32-lane SIMD order above.  K rows: hd elements (the tail
             * beyond a multiple of 32, e.g. phi-2's 80, goes through the f64 leftover loop).  V
             * columns: llama.cpp pads the number of cached positions to a multiple of 32 with
             * masked (probability 0) entries, so every position sits in a SIMD lane and there is
             * no leftover loop. */
            uint16_t q16[512];
            uint16_t *p16 = (uint16_t *)malloc((size_t)(pos + 1) * 2);
            for (int d = 0; d < hd; d++) q16[d] = f2h(q[d]);
            for (int t = 0; t <= pos; t++) {
                const uint16_t *kr = c->k16[il] + (size_t)t * nkv + (size_t)hk * hd;
                sc[t] = dot_f16_lanes32(hd, kr, 1, q16, 1, hd & ~31) * scale;
            }
            softmax_canon(sc, pos + 1);
            for (int t = 0; t <= pos; t++) p16[t] = f2h(sc[t]);
            for (int d = 0; d < hd; d++)
                o[d] = dot_f16_lanes32(pos + 1, c->v16[il] + (size_t)hk * hd + d, (int64_t)nkv, p16, 1, pos + 1);
            free(p16);
        } else if (c->faithful) {
            uint16_t q16[512];
            for (int d = 0; d < hd; d++) q16[d] = f2h(q[d]);
            for (int t = 0; t <= pos; t++) {
                const uint16_t *kr = c->k16[il] + (size_t)t * nkv + (size_t)hk * hd;
                double s = 0;
                for (int d = 0; d < hd; d++) s += (double)(h2f(kr[d]) * h2f(q16[d]));
                sc[t] = (float)s * scale;
            }
            ref_softmax(sc, pos + 1);
            for (int d = 0; d < hd; d++) {
                double s = 0;
                for (int t = 0; t <= pos; t++)
                    s += (double)(h2f(c->v16[il][(size_t)t * nkv + (size_t)hk * hd + d]) * h2f(f2h(sc[t])));
                o[d] = (float)s;
            }
        } else {
            for (int t = 0; t <= pos; t++) {
                const float *kr = c->k32[il] + (size_t)t * nkv + (size_t)hk * hd;
                double s = 0;
                for (int d = 0; d < hd; d++) s += (double)kr[d] * (double)q[d];
                sc[t] = (float)s * scale;
            }
            ref_softmax(sc, pos + 1);
            for (int d = 0; d < hd; d++) {
                double s = 0;
                for (int t = 0; t <= pos; t++) s += (double)c->v32[il][(size_t)t * nkv + (size_t)hk * hd + d] * (double)sc[t];
                o[d] = (float)s;
            }
        }
    }
}

/* Canonical (mode 2) attention of one token over n cached rows, for the kernel tests:
 * q [n_head*hd] f32, k / v [n][n_head_kv*hd] f16, out [n_head*hd]. */
void ref_attention_canon(const float *q, const uint16_t *k, const uint16_t *v, int n, int n_head, int n_head_kv, int hd, float scale, float *out) {
    const int grp = n_head / n_head_kv;
    const size_t nkv = (size_t)n_head_kv * hd;
    float *sc = (float *)malloc((size_t)n * 4);
    uint16_t *p16 = (uint16_t *)malloc((size_t)n * 2);
    for (int h = 0; h < n_head; h++) {
        const int hk = h / grp;
        uint16_t q16[512];
        for (int d = 0; d < hd; d++) q16[d] = f2h(q[(size_t)h * hd + d]);
        for (int t = 0; t < n; t++) sc[t] = dot_f16_lanes32(hd, k + (size_t)t * nkv + (size_t)hk * hd, 1, q16, 1, hd & ~31) * scale;
        softmax_canon(sc, n);
        for (int t = 0; t < n; t++) p16[t] = f2h(sc[t]);
        for (int d = 0; d < hd; d++) out[(size_t)h * hd + d] = dot_f16_lanes32(n, v + (size_t)hk * hd + d, (int64_t)nkv, p16, 1, n);
    }
    free(sc); free(p16);
}

static void store_kv(ref_ctx *c, int il, int pos) {
    const size_t nkv = (size_t)c->m->n_head_kv * c->m->head_dim;
    if (c->faithful) {
        for (size_t i = 0; i < nkv; i++) {
            c->k16[il][(size_t)pos * nkv + i] = f2h(c->k[i]);
            c->v16[il][(size_t)pos * nkv + i] = f2h(c->v[i]);
        }
    } else {
        memcpy(c->k32[il] + (size_t)pos * nkv, c->k, nkv * 4);
        memcpy(c->v32[il] + (size_t)pos * nkv, c->v, nkv * 4);
    }
}

int ref_eval(ref_ctx *c, int32_t token, int pos, float *logits) {
    const ref_model *m = c->m;
    if (token < 0 || token >= m->n_vocab || pos < 0 || pos >= m->n_ctx) return -1;
    const int E = m->n_embd, hd = m->head_dim;
    const size_t qdim = (size_t)m->n_head * hd, nkv = (size_t)m->n_head_kv * hd;
    /* get_rows: dequantise one embedding row */
    ref_dequantize_row(m->tok_embd.type,
                       (const char *)m->tok_embd.data + (size_t)token * ref_row_size(m->tok_embd.type, E), c->x, E);
    for (int il = 0; il < m->n_layer; il++) {
        const ref_layer *L = &m->layers[il];
        if (m->arch == REF_ARCH_LLAMA) {
            /* llama.cpp build_llama [upstream] */
            ref_rms_norm(c->x, (const float *)L->attn_norm.data, c->xn, E, m->norm_eps);
            matvec(c, &L->wq, qdim, E, c->xn, c->q);
            matvec(c, &L->wk, nkv, E, c->xn, c->k);
            matvec(c, &L->wv, nkv, E, c->xn, c->v);
            ref_rope_ext(c->q, m->n_head, hd, m->n_rot, pos, m->rope_freq_base, 0, m->rope_freq_scale, m->rope_factors);
            ref_rope_ext(c->k, m->n_head_kv, hd, m->n_rot, pos, m->rope_freq_base, 0, m->rope_freq_scale, m->rope_factors);
            store_kv(c, il, pos);
            attend(c, il, pos, 1.0f / sqrtf((float)hd));
            matvec(c, &L->wo, E, qdim, c->att, c->proj);
            for (int i = 0; i < E; i++) c->x[i] += c->proj[i];
            ref_rms_norm(c->x, (const float *)L->ffn_norm.data, c->xn, E, m->norm_eps);
            matvec(c, &L->w_gate, m->n_ff, E, c->xn, c->gate);
            matvec(c, &L->w_up, m->n_ff, E, c->xn, c->up);
            for (int i = 0; i < m->n_ff; i++) c->gate[i] = silu_f(c->gate[i], c->faithful) * c->up[i];
            matvec(c, &L->w_down, E, m->n_ff, c->gate, c->ffn);
            for (int i = 0; i < E; i++) c->x[i] += c->ffn[i];
        } else {
            /* llama.cpp build_phi2 [upstream]: one LayerNorm feeds attention and MLP in parallel */
            ref_layer_norm(c->x, (const float *)L->attn_norm.data, (const float *)L->attn_norm_b.data, c->xn, E, m->norm_eps);
            matvec(c, &L->wq, qdim + 2 * nkv, E, c->xn, c->q);      /* fused qkv */
            add_bias(c->q, &L->bq, qdim + 2 * nkv);
            memcpy(c->k, c->q + qdim, nkv * 4);
            memcpy(c->v, c->q + qdim + nkv, nkv * 4);
            ref_rope_ext(c->q, m->n_head, hd, m->n_rot, pos, m->rope_freq_base, 2, m->rope_freq_scale, m->rope_factors);
            const float qs = 1.0f / sqrtf((float)hd);
            for (size_t i = 0; i < qdim; i++) c->q[i] *= qs;        /* q pre-scaled, kq_scale = 1 */
            ref_rope_ext(c->k, m->n_head_kv, hd, m->n_rot, pos, m->rope_freq_base, 2, m->rope_freq_scale, m->rope_factors);
            store_kv(c, il, pos);
            attend(c, il, pos, 1.0f);
            matvec(c, &L->wo, E, qdim, c->att, c->proj);
            add_bias(c->proj, &L->bo, E);
            matvec(c, &L->w_up, m->n_ff, E, c->xn, c->up);
            add_bias(c->up, &L->b_up, m->n_ff);
            for (int i = 0; i < m->n_ff; i++) c->up[i] = gelu_f(c->up[i], c->faithful);
            matvec(c, &L->w_down, E, m->n_ff, c->up, c->ffn);
            add_bias(c->ffn, &L->b_down, E);
            for (int i = 0; i < E; i++) c->x[i] = (c->ffn[i] + c->proj[i]) + c->x[i];
        }
    }
    if (m->arch == REF_ARCH_LLAMA) ref_rms_norm(c->x, (const float *)m->out_norm.data, c->hidden, E, m->norm_eps);
    else ref_layer_norm(c->x, (const float *)m->out_norm.data, (const float *)m->out_norm_b.data, c->hidden, E, m->norm_eps);
    if (logits) {
        matvec(c, &m->output, m->n_vocab, E, c->hidden, logits);
        add_bias(logits, &m->output_b, m->n_vocab);
    }
    return 0;
}

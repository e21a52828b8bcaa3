// qmat_dev.cuh -- device-side element access into a split-plane weight matrix (QMat).
//
// Inside a row, planes that are read 16 bytes per lane by the GEMV kernel are stored
// "chunk-interleaved": a unit (a Q8_0 block, or a 128-weight half super-block of a K-quant)
// owns C consecutive 16-byte chunks in the file; in HBM chunk c of unit u sits at
// ((c * n_units) + u) * 16, so that lanes working on consecutive units read consecutive
// 16-byte words (conflict-free LDS.128 after the bulk copy into shared memory).
#pragma once
#include "common.cuh"
#include "formats.cuh"

namespace b2 {

__device__ __forceinline__ float h2f_bits(uint16_t h) { return __half2float(__ushort_as_half(h)); }

// 6-bit (scale, min) pair j of a Q4_K block (gguf/quants.py:479-502)
__device__ __forceinline__ void q4k_scale_min(int j, const uint8_t* q, int& sc, int& mn) {
    if (j < 4) {
        sc = q[j] & 63;
        mn = q[j + 4] & 63;
    } else {
        sc = (q[j + 4] & 0x0F) | ((q[j - 4] >> 6) << 4);
        mn = (q[j + 4] >> 4) | ((q[j] >> 6) << 4);
    }
}

// byte offset inside a row of a chunk-interleaved plane
__host__ __device__ __forceinline__ size_t il_off(int64_t n_units, int64_t unit, int byte_in_unit) {
    return ((size_t)(byte_in_unit >> 4) * n_units + unit) * 16 + (byte_in_unit & 15);
}

__device__ __forceinline__ float qmat_elem(const QMat& m, int64_t row, int64_t col) {
    switch (m.type) {
    case T_Q4_0: {
        const int64_t b = col / 32; const int j = (int)(col % 32);
        const uint8_t byte = m.plane[0][(size_t)row * m.stride[0] + b * 16 + (j & 15)];
        const int q = j < 16 ? (byte & 0x0F) : (byte >> 4);
        const uint16_t d = *reinterpret_cast<const uint16_t*>(m.plane[1] + (size_t)row * m.stride[1] + b * 2);
        return __fmul_rn((float)(q - 8), h2f_bits(d));
    }
    case T_Q8_0: {
        const int64_t b = col / 32, nu = m.k / 32;
        const int8_t q = (int8_t)m.plane[0][(size_t)row * m.stride[0] + il_off(nu, b, (int)(col % 32))];
        const uint16_t d = *reinterpret_cast<const uint16_t*>(m.plane[1] + (size_t)row * m.stride[1] + b * 2);
        return __fmul_rn((float)q, h2f_bits(d));
    }
    case T_Q4_K: {
        const int64_t b = col / 256, nu = m.k / 128; const int j = (int)(col % 256), sub = j / 32, l = j % 32;
        int sc, mn;
        q4k_scale_min(sub, m.plane[1] + (size_t)row * m.stride[1] + b * 12, sc, mn);
        const int qoff = (sub / 2) * 32 + l;            // byte inside the block's qs[128]
        const uint8_t byte = m.plane[0][(size_t)row * m.stride[0] + il_off(nu, b * 2 + qoff / 64, qoff % 64)];
        const int q = (sub & 1) ? (byte >> 4) : (byte & 0x0F);
        const uint16_t* dm = reinterpret_cast<const uint16_t*>(m.plane[2] + (size_t)row * m.stride[2] + b * 4);
        const float d = __fmul_rn(h2f_bits(dm[0]), (float)sc), mm = __fmul_rn(h2f_bits(dm[1]), (float)mn);
        return __fsub_rn(__fmul_rn(d, (float)q), mm);
    }
    case T_Q6_K: {
        const int64_t b = col / 256, nu = m.k / 128;
        const int j = (int)(col % 256), half = j / 128, r = j % 128, quad = r / 32, l = r % 32;
        const int64_t u = b * 2 + half;
        const uint8_t lo = m.plane[0][(size_t)row * m.stride[0] + il_off(nu, u, (quad & 1) * 32 + l)];
        const int q4 = (quad & 2) ? (lo >> 4) : (lo & 0x0F);
        const int q2 = (m.plane[1][(size_t)row * m.stride[1] + il_off(nu, u, l)] >> (2 * quad)) & 3;
        const int q = (int)(int8_t)(q4 | (q2 << 4)) - 32;
        const int sc = (int8_t)m.plane[2][(size_t)row * m.stride[2] + b * 16 + half * 8 + quad * 2 + l / 16];
        const uint16_t d = *reinterpret_cast<const uint16_t*>(m.plane[3] + (size_t)row * m.stride[3] + b * 2);
        return __fmul_rn(__fmul_rn(h2f_bits(d), (float)sc), (float)q);
    }
    }
    return 0.f;
}

}  // namespace b2

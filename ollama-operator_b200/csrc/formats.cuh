// formats.cuh -- GGUF block formats as they sit in the file, and the split-plane layout the
// engine keeps them in once resident in HBM.
//
// File formats (gguf/constants.py:4235-4245, gguf/quants.py:241-251,396-401,504-522,554-572):
//   Q4_0  32 w / 18 B  {f16 d; u8 qs[16]}            w[j] = (qs[j]&15)-8, w[j+16] = (qs[j]>>4)-8
//   Q8_0  32 w / 34 B  {f16 d; i8 qs[32]}
//   Q4_K 256 w / 144 B {f16 d; f16 dmin; u8 scales[12]; u8 qs[128]}
//   Q6_K 256 w / 210 B {u8 ql[128]; u8 qh[64]; i8 scales[16]; f16 d}
//
// None of these is 16-byte aligned, so at load time every weight matrix is split into planes
// (same bytes, same order inside a block) whose rows start on 16-byte boundaries; a tile of
// consecutive rows is then one contiguous range per plane and can be fetched with one bulk
// async copy per plane:
//   Q4_0: plane0 = qs  [rows][K/32][16 B]      plane1 = d  [rows][K/32] f16
//   Q8_0: plane0 = qs  [rows][K/32][32 B]      plane1 = d  [rows][K/32] f16
//   Q4_K: plane0 = qs  [rows][K/256][128 B]    plane1 = scales [rows][K/256][12 B]
//         plane2 = {d,dmin} [rows][K/256] half2
//   Q6_K: plane0 = ql  [rows][K/256][128 B]    plane1 = qh [rows][K/256][64 B]
//         plane2 = scales [rows][K/256][16 B]  plane3 = d  [rows][K/256] f16
// Row strides are rounded up to 16 B (only matters for toy shapes).
#pragma once
#include <stddef.h>
#include <stdint.h>

namespace b2 {

enum GgmlType : int { T_F32 = 0, T_F16 = 1, T_Q4_0 = 2, T_Q8_0 = 8, T_Q4_K = 12, T_Q6_K = 14 };
constexpr int T_KQ = 100;      // persistent-kernel variant for K-quant models: every matrix is Q4_K or Q6_K (Q4_K_M mixes them)

#pragma pack(push, 1)
struct BlockQ4_0 { uint16_t d; uint8_t qs[16]; };
struct BlockQ8_0 { uint16_t d; int8_t qs[32]; };
struct BlockQ4_K { uint16_t d, dmin; uint8_t scales[12]; uint8_t qs[128]; };
struct BlockQ6_K { uint8_t ql[128]; uint8_t qh[64]; int8_t scales[16]; uint16_t d; };
#pragma pack(pop)
static_assert(sizeof(BlockQ4_0) == 18 && sizeof(BlockQ8_0) == 34 && sizeof(BlockQ4_K) == 144 && sizeof(BlockQ6_K) == 210, "block sizes");

inline int type_block_elems(int t) { return (t == T_Q4_0 || t == T_Q8_0) ? 32 : (t == T_Q4_K || t == T_Q6_K) ? 256 : 1; }
inline int type_block_bytes(int t) {
    switch (t) {
    case T_F32: return 4;
    case T_F16: return 2;
    case T_Q4_0: return 18;
    case T_Q8_0: return 34;
    case T_Q4_K: return 144;
    case T_Q6_K: return 210;
    }
    return 0;
}
inline bool type_is_quant(int t) { return t == T_Q4_0 || t == T_Q8_0 || t == T_Q4_K || t == T_Q6_K; }
inline size_t type_row_bytes(int t, int64_t k) { return (size_t)(k / type_block_elems(t)) * type_block_bytes(t); }
inline int type_num_planes(int t) { return t == T_Q4_0 || t == T_Q8_0 ? 2 : t == T_Q4_K ? 3 : t == T_Q6_K ? 4 : 0; }
// bytes of plane p per block
inline int type_plane_block_bytes(int t, int p) {
    static const int q40[4] = {16, 2, 0, 0}, q80[4] = {32, 2, 0, 0}, q4k[4] = {128, 12, 4, 0}, q6k[4] = {128, 64, 16, 2};
    switch (t) {
    case T_Q4_0: return q40[p];
    case T_Q8_0: return q80[p];
    case T_Q4_K: return q4k[p];
    case T_Q6_K: return q6k[p];
    }
    return 0;
}

// A quantised weight matrix resident in HBM in split-plane form.
struct QMat {
    int type = 0;
    int64_t rows = 0, k = 0;
    int n_planes = 0;
    uint8_t* plane[4] = {nullptr, nullptr, nullptr, nullptr};
    uint32_t stride[4] = {0, 0, 0, 0};   // bytes per row in each plane (multiple of 16)
    size_t bytes() const {
        size_t s = 0;
        for (int p = 0; p < n_planes; p++) s += (size_t)rows * stride[p];
        return s;
    }
};

inline void qmat_shape(QMat& m, int type, int64_t rows, int64_t k) {
    m.type = type; m.rows = rows; m.k = k;
    m.n_planes = type_num_planes(type);
    const int64_t nb = k / type_block_elems(type);
    for (int p = 0; p < 4; p++) {
        const size_t b = p < m.n_planes ? (size_t)nb * type_plane_block_bytes(type, p) : 0;
        m.stride[p] = (uint32_t)((b + 15) / 16 * 16);
    }
}

}  // namespace b2

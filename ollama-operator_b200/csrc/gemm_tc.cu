// gemm_tc.cu -- Q4_0 weights x up to 128 Q8_0 activation columns per pass on the 5th-generation tensor
// cores (tcgen05.mma kind::i8, accumulators in tensor memory), for chunked prefill and batched decode.
//
// Arithmetic.  ggml's mul_mat quantises every activation row to Q8_0 and adds, per 32-weight block,
// ((float)sumi * d_w) * d_a with an EXACT integer block sum.  That per-block scaling is what has to be kept:
// a tensor-core accumulator may therefore only run over K = 32 (one block) before it is read back.  kind::i8
// has exactly that K, so every MMA starts from a zero accumulator:
//     D[128 rows x NT columns] (s32, tensor memory)  =  A[128 x 32] (s8: nibble - 8)  x  B[32 x NT] (s8: q_a)
// and the s32 block sums are drained with tcgen05.ld and scaled in registers.  Two accumulation modes:
//     ACC64  the engine's canonical arithmetic (terms added in f64, one rounding): bit-identical to the
//            one-token kernels, at the cost of one f32->f64 conversion per term (the XU pipe bounds it);
//     ACC32  f32, block after block: ggml's generic C order (oracle mode 4), ~1.7x faster.
//
// Pipeline of one CTA = one tile of 128 weight rows x NT columns, over all of K in chunks of 8 blocks:
//   warp 0      producer: TMA tensor-map boxes (nibbles 128 x 128 B with the 128-byte hardware swizzle, scales
//               128 x 16 B, the s8 activation columns as two [NT x 128 B] swizzled boxes = the UMMA operand
//               layout as it lands) plus bulk copies of the activation scales, mbarrier complete_tx;
//   warps 2-5   unpack: thread r turns row r's 8 x 16 B of nibbles into 256 B of s8 in the K-major 128-byte
//               swizzled UMMA layout (a fence.proxy.async hands it to the tensor core);
//   warp 1      one thread issues tcgen05.mma per block into one of two TMEM accumulator buffers and commits
//               to mbarriers (buffer full / stage free); it also owns the TMEM allocation;
//   warps 6..   epilogue: tcgen05.ld of the block sums (32 lanes x 32 columns per load), release of the TMEM
//               buffer, scaling with d_w (per row) and d_a (per column, broadcast from shared memory),
//               accumulation in registers; at the end the GEMV epilogues (RoPE + KV append, SiLU, residual...)
//               through a shared-memory transpose so that one thread holds a row pair.
#include <cuda.h>
#include <cudaTypedefs.h>
#include <math.h>

#include <algorithm>
#include <type_traits>

#include "common.cuh"
#include "formats.cuh"
#include "gemv_dev.cuh"
#include "kernels.h"

namespace b2 {

namespace {
using namespace gemvdev;

constexpr int TC_M = 128;                 // weight rows per CTA
constexpr int TC_KC = 8;                  // 32-weight blocks per K chunk (256 weights = two 128-byte atoms of s8)
constexpr int TC_NS = 3;                  // pipeline stages
constexpr int TC_NB = 8;                  // accumulator buffers in tensor memory: the MMAs run up to eight blocks ahead of the drain
constexpr int TC_UNPACK_WARPS = 2;             // each thread unpacks two rows (12 warps in all leave the epilogue 168 registers)
constexpr int TC_FIRST_EPI_WARP = 2 + TC_UNPACK_WARPS;

template <int NT> struct TcCfg {
    static constexpr int NEPI = NT / 8;                              // epilogue warps: 4 TMEM lane quarters x (NT / 32) column groups
    static constexpr int CPT = 32;                                   // accumulator columns per epilogue thread (one tcgen05.ld.32x32b.x32)
    static_assert(NT == 32 || NT == 64, "column tile");
    static constexpr int THREADS = (TC_FIRST_EPI_WARP + NEPI) * 32;
    static constexpr int OFF_RAW = 0;                                // nibbles  [128 rows][128 B], swizzled
    static constexpr int OFF_A = OFF_RAW + TC_M * 128;               // s8 A     2 atoms x [128 rows][128 B]
    static constexpr int OFF_B = OFF_A + 2 * TC_M * 128;             // s8 B     2 atoms x [NT cols][128 B]
    static constexpr int OFF_SC = OFF_B + 2 * NT * 128;              // f16 d_w  [128 rows][8]
    static constexpr int OFF_DA = OFF_SC + TC_M * TC_KC * 2;         // f32 d_a  [8 blocks][NT]
    static constexpr int STAGE = (OFF_DA + TC_KC * NT * 4 + 1023) / 1024 * 1024;
    static constexpr int TX_BYTES = TC_M * 128 + 2 * NT * 128 + TC_M * TC_KC * 2 + TC_KC * NT * 4;
    static constexpr int TMEM_COLS = TC_NB * NT;                     // 256 or 512 columns (a power of two)
    static constexpr size_t SMEM = 1024 + (size_t)TC_NS * STAGE + 512;
    static_assert(OFF_A % 1024 == 0 && OFF_B % 1024 == 0 && OFF_SC % 128 == 0 && OFF_DA % 16 == 0, "stage layout alignment");
    static_assert((size_t)TC_M * (NT + 1) * 8 <= (size_t)TC_NS * STAGE, "output staging reuses the pipeline stages");
    static_assert(SMEM <= 227 * 1024, "shared memory budget");
};

// ---- tcgen05 / TMA wrappers ----------------------------------------------------------------------------
__device__ __forceinline__ void tma_box_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2, %3}], [%4], %5;" ::"r"(
            smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// mbarrier arrive once every tcgen05 operation this thread issued so far has completed
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] = A[smem desc] x B[smem desc], s8 x s8 -> s32, fresh accumulator (enable_input_d = 0)
__device__ __forceinline__ void tc_mma_i8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, 0, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc)
        : "memory");
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, int (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
          "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
          "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]),
          "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
// Wait for this thread's outstanding tcgen05.ld.  The loaded registers are passed through as in/out operands so that
// the compiler cannot schedule a use of them above the wait (the load is asynchronous until here).
__device__ __forceinline__ void tc_wait_ld(int (&r)[32]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]), "+r"(r[9]), "+r"(r[10]),
                   "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]), "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]),
                   "+r"(r[21]), "+r"(r[22]), "+r"(r[23]), "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]),
                   "+r"(r[31])
                 :
                 : "memory");
}

// Shared-memory matrix descriptor of a K-major operand in the 128-byte swizzled layout: rows of 128 bytes,
// groups of 8 rows 1024 bytes apart (stride byte offset), descriptor version 1 (sm_100), layout type 2.
__device__ __forceinline__ uint64_t tc_desc_sw128(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
// Instruction descriptor: D = s32, A = B = signed 8-bit, both K-major, M = 128, N = nt
__host__ __device__ constexpr uint32_t tc_idesc_i8(int nt) {
    return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(nt >> 3) << 17) | ((uint32_t)(TC_M >> 4) << 24);
}

// A wait that cannot hang the GPU: a lost arrival (a bug, not a load condition) traps after about a second.
__device__ __forceinline__ void mbar_wait_bounded(uint64_t* bar, uint32_t parity) {
    for (uint32_t spins = 0; !mbar_try_wait(bar, parity); spins++)
        if (spins > (1u << 26)) __trap();
}

// four packed nibbles-as-bytes (0..15) -> four signed bytes (value - 8), without carries between the bytes
__device__ __forceinline__ uint32_t sub8_bytes(uint32_t x) {
    return (x ^ 0x08080808u) | (((~x) & 0x08080808u) * 30u);        // x < 8: (x + 8) | 0xF0 = x - 8 as a signed byte
}

template <int NT, bool ACC64>
__global__ void __launch_bounds__(TcCfg<NT>::THREADS, 1)
gemm_tc_kernel(const __grid_constant__ TcMaps wmaps, const __grid_constant__ CUtensorMap xmap, const float* __restrict__ xd, const int xd_stride,
               const int rows, const int nbk, const int B, const GemvArgs* __restrict__ colargs, const int* __restrict__ col_pos,
               const int* __restrict__ col_pt, const int pt_stride, const int debug) {
    using C = TcCfg<NT>;
    using acc_t = typename std::conditional<ACC64, double, float>::type;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);      // swizzled operands need 1024-byte alignment
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)TC_NS * C::STAGE);
    uint64_t *full = bars, *aready = bars + TC_NS, *sfree = bars + 2 * TC_NS, *dfull = bars + 3 * TC_NS, *dfree = bars + 3 * TC_NS + TC_NB;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * TC_NS + 2 * TC_NB);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int row0 = blockIdx.x * TC_M, col0 = blockIdx.y * NT;
    const int ncols = min(NT, B - col0);
    const int n_chunks = nbk / TC_KC;

    if (tid == 0) {
        for (int s = 0; s < TC_NS; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&aready[s], TC_UNPACK_WARPS);
            mbar_init(&sfree[s], 1 + C::NEPI);
        }
        for (int i = 0; i < TC_NB; i++) {
            mbar_init(&dfull[i], 1);
            mbar_init(&dfree[i], C::NEPI);
        }
        mbar_fence_init();
    }
    if (warp == 1) {        // tensor memory: two accumulator buffers of NT columns
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"((uint32_t)C::TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_trigger();

    if (warp == 0) {
        // ===== producer =====
        if (lane == 0) {
            const uint64_t pol_w = l2_policy_evict_first(), pol_x = l2_policy_evict_last();
            auto issue_w = [&](int it) {
                const int s = it % TC_NS;
                uint8_t* st = smem + (size_t)s * C::STAGE;
                mbar_expect_tx(&full[s], (debug & 8) ? C::TX_BYTES - TC_KC * NT * 4 : C::TX_BYTES);
                tma_box_2d(st + C::OFF_RAW, &wmaps.q, it * (TC_KC * 16), row0, &full[s], pol_w);
                tma_box_2d(st + C::OFF_SC, &wmaps.d, it * (TC_KC * 2), row0, &full[s], pol_w);
            };
            auto issue_x = [&](int it) {
                const int s = it % TC_NS;
                uint8_t* st = smem + (size_t)s * C::STAGE;
                tma_box_2d(st + C::OFF_B, &xmap, it * (TC_KC * 32), col0, &full[s], pol_x);
                tma_box_2d(st + C::OFF_B + NT * 128, &xmap, it * (TC_KC * 32) + 128, col0, &full[s], pol_x);
                if (!(debug & 8))
#pragma unroll
                for (int b = 0; b < TC_KC; b++)
                    bulk_g2s(st + C::OFF_DA + b * NT * 4, xd + (size_t)(it * TC_KC + b) * xd_stride + col0, NT * 4, &full[s], pol_x);
            };
            // the weights do not depend on the preceding kernel: their first boxes go out before griddepcontrol.wait
            const int n_pre = min(n_chunks, TC_NS);
            for (int it = 0; it < n_pre; it++) issue_w(it);
            pdl_wait();
            for (int it = 0; it < n_pre; it++) issue_x(it);
            for (int it = n_pre; it < n_chunks; it++) {
                mbar_wait_bounded(&sfree[it % TC_NS], ((it / TC_NS) - 1) & 1);
                issue_w(it);
                issue_x(it);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            constexpr uint32_t idesc = tc_idesc_i8(NT);
            for (int it = 0; it < n_chunks; it++) {
                const int s = it % TC_NS;
                const uint32_t par = (it / TC_NS) & 1;
                mbar_wait_bounded(&full[s], par);           // activation columns (and the raw weights) have landed
                mbar_wait_bounded(&aready[s], par);         // the s8 weight operand is written
                tc_fence_after();
                const uint32_t a_base = smem_u32(smem + (size_t)s * C::STAGE + C::OFF_A), b_base = smem_u32(smem + (size_t)s * C::STAGE + C::OFF_B);
#pragma unroll
                for (int b = 0; b < TC_KC; b++) {
                    const int g = it * TC_KC + b, buf = g % TC_NB;
                    if (g >= TC_NB && !(debug & 32)) { mbar_wait_bounded(&dfree[buf], ((g / TC_NB) - 1) & 1); tc_fence_after(); }
                    const uint64_t ad = tc_desc_sw128(a_base + (b >> 2) * (TC_M * 128) + (b & 3) * 32);
                    const uint64_t bd = tc_desc_sw128(b_base + (b >> 2) * (NT * 128) + (b & 3) * 32);
                    tc_mma_i8(tmem_base + buf * NT, ad, bd, idesc);
                    if (debug & 16) { tc_mma_i8(tmem_base + buf * NT, ad, bd, idesc); tc_mma_i8(tmem_base + buf * NT, ad, bd, idesc); tc_mma_i8(tmem_base + buf * NT, ad, bd, idesc); }
                    tc_commit(&dfull[buf]);
                }
                tc_commit(&sfree[s]);                       // this stage's operands are no longer read by the tensor core
            }
        }
    } else if (warp < TC_FIRST_EPI_WARP) {
        // ===== unpack: nibbles -> signed bytes in the UMMA operand layout =====
        for (int it = 0; it < n_chunks; it++) {
            const int s = it % TC_NS;
            mbar_wait_bounded(&full[s], (it / TC_NS) & 1);
          if (!(debug & 4))
#pragma unroll
          for (int rr = 0; rr < TC_M / (TC_UNPACK_WARPS * 32); rr++) {
            const int r = rr * (TC_UNPACK_WARPS * 32) + (warp - 2) * 32 + lane, r7 = r & 7;
            const uint32_t row_off = (uint32_t)(r >> 3) * 1024u + (uint32_t)r7 * 128u;
            const uint8_t* raw = smem + (size_t)s * C::STAGE + C::OFF_RAW + row_off;
            uint8_t* arow = smem + (size_t)s * C::STAGE + C::OFF_A + row_off;
#pragma unroll
            for (int c = 0; c < TC_KC; c++) {
                const uint4 w = *reinterpret_cast<const uint4*>(raw + ((c ^ r7) << 4));
                uint4 lo, hi;
                lo.x = sub8_bytes(w.x & 0x0F0F0F0Fu); hi.x = sub8_bytes((w.x >> 4) & 0x0F0F0F0Fu);
                lo.y = sub8_bytes(w.y & 0x0F0F0F0Fu); hi.y = sub8_bytes((w.y >> 4) & 0x0F0F0F0Fu);
                lo.z = sub8_bytes(w.z & 0x0F0F0F0Fu); hi.z = sub8_bytes((w.z >> 4) & 0x0F0F0F0Fu);
                lo.w = sub8_bytes(w.w & 0x0F0F0F0Fu); hi.w = sub8_bytes((w.w >> 4) & 0x0F0F0F0Fu);
                uint8_t* atom = arow + (c >> 2) * (TC_M * 128);
                const int ch = 2 * (c & 3);                   // block c of the chunk: elements 0..15 then 16..31, 16 bytes each
                *reinterpret_cast<uint4*>(atom + ((ch ^ r7) << 4)) = lo;
                *reinterpret_cast<uint4*>(atom + (((ch + 1) ^ r7) << 4)) = hi;
            }
          }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy writes -> visible to the tensor core's reads
            __syncwarp();
            if (lane == 0) mbar_arrive(&aready[s]);
        }
    } else {
        // ===== epilogue =====
        pdl_wait();
        const int e = warp - TC_FIRST_EPI_WARP, q = warp & 3, cg = e >> 2;     // a warp reaches the TMEM lanes of quarter (warp index % 4)
        const int rl = q * 32 + lane, colbase = cg * C::CPT;
        const int etid = e * 32 + lane;
        constexpr int ENT = C::NEPI * 32;
        acc_t acc[C::CPT];
#pragma unroll
        for (int j = 0; j < C::CPT; j++) acc[j] = (acc_t)0;
        const int nblk = n_chunks * TC_KC;                              // even
        const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)colbase;
        // scale the 32 block sums of block g (row rl, columns colbase..+31) and add them to the accumulators
        auto accumulate = [&](const int (&v)[32], int g) {
            const int s = (g / TC_KC) % TC_NS, b = g % TC_KC;
            if (b == 0) mbar_wait_bounded(&full[s], ((g / TC_KC) / TC_NS) & 1);        // this chunk's scales are in shared memory
            const float dw = __half2float(reinterpret_cast<const __half*>(smem + (size_t)s * C::STAGE + C::OFF_SC)[rl * TC_KC + b]);
            const float* da = reinterpret_cast<const float*>(smem + (size_t)s * C::STAGE + C::OFF_DA) + b * NT + colbase;
            if (!(debug & 1)) {
#pragma unroll
            for (int j4 = 0; j4 < 8; j4++) {
                const float4 d4 = *reinterpret_cast<const float4*>(da + j4 * 4);
                const float dd[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
                for (int jj = 0; jj < 4; jj++) {
                    const float t = __fmul_rn(__fmul_rn((float)v[j4 * 4 + jj], dw), dd[jj]);
                    acc_t& a = acc[j4 * 4 + jj];
                    if (ACC64) a += (acc_t)t;
                    else a = (acc_t)__fadd_rn((float)a, t);
                }
            }
            }
            if (b == TC_KC - 1) {                                       // done with this stage's scales
                __syncwarp();
                if (lane == 0) mbar_arrive(&sfree[s]);
            }
        };
        // The drain of block g + 1 (tcgen05.ld is asynchronous until tcgen05.wait::ld) runs under the arithmetic of block g:
        // two register sets, two TMEM buffers, the loop unrolled by two so that both are static.
        int vA[32], vB[32];
        mbar_wait_bounded(&dfull[0], 0);
        tc_fence_after();
        tc_ld32(tbase, vA);
        for (int g = 0; g < nblk; g += 2) {
            const int b0 = g % TC_NB, b1 = (g + 1) % TC_NB, b2 = (g + 2) % TC_NB;
            tc_wait_ld(vA);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&dfree[b0]);                     // the buffer may take the MMA of block g + TC_NB
            mbar_wait_bounded(&dfull[b1], ((g + 1) / TC_NB) & 1);
            tc_fence_after();
            if (!(debug & 2)) tc_ld32(tbase + b1 * NT, vB);
            accumulate(vA, g);
            tc_wait_ld(vB);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&dfree[b1]);
            if (g + 2 < nblk) {
                mbar_wait_bounded(&dfull[b2], ((g + 2) / TC_NB) & 1);
                tc_fence_after();
                if (!(debug & 2)) tc_ld32(tbase + b2 * NT, vA);
            }
            accumulate(vB, g + 1);
        }
        // ---- the GEMV epilogues: transpose through shared memory so that one thread holds a row pair of a column ----
        // (every stage has been consumed: all tensor-core reads and all TMA writes of this CTA are complete)
        named_bar_sync(1, ENT);
        const bool partial = colargs[0].epi == EPI_PARTIAL;      // tensor-parallel K slice: the f64 sums leave unrounded
        float* outf = reinterpret_cast<float*>(smem);
        double* outd = reinterpret_cast<double*>(smem);
#pragma unroll
        for (int j = 0; j < C::CPT; j++) {
            if (ACC64 && partial) outd[(size_t)rl * NT + colbase + j] = (double)acc[j];
            else outf[(size_t)rl * (NT + 1) + colbase + j] = (float)acc[j];
        }
        named_bar_sync(1, ENT);
        for (int i = etid; i < (TC_M / 2) * ncols; i += ENT) {
            const int p = i / ncols, c = i - p * ncols;
            const int64_t r0 = (int64_t)row0 + 2 * p;
            if (r0 < rows) {
                const GemvArgs& a = colargs[col0 + c];
                if (ACC64 && partial) {
                    a.y64[a.epi_row0 + r0] = outd[(size_t)(2 * p) * NT + c];
                    a.y64[a.epi_row0 + r0 + 1] = outd[(size_t)(2 * p + 1) * NT + c];
                } else {
                    unsigned long long best = 0ull;
                    epilogue_pair(a, a.epi_row0 + r0, outf[(size_t)(2 * p) * (NT + 1) + c], outf[(size_t)(2 * p + 1) * (NT + 1) + c], best, col_pos[col0 + c],
                                  col_pt + (size_t)(col0 + c) * pt_stride);
                    if (a.epi == EPI_LOGITS && best != 0ull) atomicMax(a.argmax_key, best);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)C::TMEM_COLS) : "memory");
    }
}

// ---- activation columns: the GEMV prologue (norm + Q8_0 quantiser, same bits) writing plain K-major s8 rows ------
constexpr int TQ_SLICE = 1024;           // elements one CTA quantises
__global__ void __launch_bounds__(256) quant_cols_tc_kernel(GemvArgs a, int64_t x_stride, int8_t* q, float* d, int64_t k_cap, int cols_cap) {
    __shared__ double dscratch[32];
    const int c = blockIdx.x, tid = threadIdx.x;
    a.x += (size_t)c * x_stride;
    pdl_trigger();
    pdl_wait();          // the activations come from the preceding kernel, which may also still be reading the buffer we overwrite
    const ProStats st = prologue_stats(a, dscratch, tid, 256);
    const int k0 = blockIdx.y * TQ_SLICE, k1 = min((int)a.W.k, k0 + TQ_SLICE);
    prologue_q8_0_each(a, st, tid, 256, [=](int blk, int o, uint32_t w, float dd, int) {
        *reinterpret_cast<uint32_t*>(q + (size_t)c * k_cap + blk * 32 + o) = w;
        if (o == 0) d[(size_t)blk * cols_cap + c] = dd;
    }, k0, k1);
}

PFN_cuTensorMapEncodeTiled tc_encode_fn() {
    static PFN_cuTensorMapEncodeTiled fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled>(p);
    }
    return fn;
}

int encode_2d(CUtensorMap* out, void* base, uint64_t inner_bytes, uint64_t n_rows, uint64_t row_stride, uint32_t box_inner, uint32_t box_rows, bool swizzle128,
              const char* what) {
    PFN_cuTensorMapEncodeTiled enc = tc_encode_fn();
    if (!enc) B2_FAIL(B2_ERR_CUDA, "gemm_tc: cuTensorMapEncodeTiled is not available from this driver");
    const cuuint64_t dim[2] = {(cuuint64_t)inner_bytes, (cuuint64_t)n_rows}, str[1] = {(cuuint64_t)row_stride};
    const cuuint32_t box[2] = {box_inner, box_rows}, estr[2] = {1, 1};
    const CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dim, str, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) B2_FAIL(B2_ERR_CUDA, "gemm_tc: tensor map (%s) rejected, CUresult %d", what, (int)r);
    return B2_OK;
}

template <int NT, bool ACC64>
int gemm_tc_launch_t(const TcMaps& maps, const TcCols& x, int rows, int nbk, int B, const GemvArgs* colargs, const int* pos, const int* pt, int pt_stride,
                     cudaStream_t st) {
    using C = TcCfg<NT>;
    auto kern = gemm_tc_kernel<NT, ACC64>;
    static PerDeviceSmem configured;
    if (configured.raise(C::SMEM)) B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((rows + TC_M - 1) / TC_M, (B + NT - 1) / NT);
    cfg.blockDim = dim3(C::THREADS);
    cfg.dynamicSmemBytes = C::SMEM;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = batch_pdl_enabled() ? 1 : 0;
    const CUtensorMap& xm = NT == 64 ? x.map64 : x.map32;
    static const int debug = getenv("B2_TC_DEBUG") ? atoi(getenv("B2_TC_DEBUG")) : 0;      // tuning experiments (results are wrong when set)
    B2_CUDA(cudaLaunchKernelEx(&cfg, kern, maps, xm, (const float*)x.d, x.cols_cap, rows, nbk, B, colargs, pos, pt, pt_stride, debug));
    return B2_OK;
}

}  // namespace

bool gemm_tc_supported(const QMat& W) { return W.type == T_Q4_0 && W.k % 256 == 0 && (W.rows & 1) == 0 && W.stride[0] % 16 == 0 && W.stride[1] % 16 == 0; }

int gemm_tc_make_maps(const QMat& W, TcMaps* out) {
    if (!gemm_tc_supported(W)) B2_FAIL(B2_ERR_UNSUPPORTED, "gemm_tc: type %d shape %lld x %lld has no tensor-core kernel", W.type, (long long)W.rows, (long long)W.k);
    B2_TRY(encode_2d(&out->q, W.plane[0], (uint64_t)W.k / 2, (uint64_t)W.rows, W.stride[0], 128, TC_M, true, "nibbles"));
    B2_TRY(encode_2d(&out->d, W.plane[1], (uint64_t)W.k / 16, (uint64_t)W.rows, W.stride[1], TC_KC * 2, TC_M, false, "scales"));
    return B2_OK;
}

// q: [cols_cap][k_cap] s8, d: [k_cap / 32][cols_cap] f32, both caller-allocated (q 128-byte aligned)
int tc_cols_init(TcCols& c, int8_t* q, float* d, int64_t k_cap, int cols_cap) {
    if (k_cap % 256 || cols_cap % 32 || cols_cap > TC_COLS_MAX) B2_FAIL(B2_ERR_ARG, "tc_cols_init: k_cap %lld, cols_cap %d", (long long)k_cap, cols_cap);
    c.q = q; c.d = d; c.k_cap = k_cap; c.cols_cap = cols_cap;
    B2_TRY(encode_2d(&c.map32, q, (uint64_t)k_cap, (uint64_t)cols_cap, (uint64_t)k_cap, 128, 32, true, "activation columns / 32"));
    B2_TRY(encode_2d(&c.map64, q, (uint64_t)k_cap, (uint64_t)cols_cap, (uint64_t)k_cap, 128, std::min(64, cols_cap), true, "activation columns / 64"));
    B2_TRY(encode_2d(&c.map128, q, (uint64_t)k_cap, (uint64_t)cols_cap, (uint64_t)k_cap, 128, std::min(128, cols_cap), true, "activation columns / 128"));
    return B2_OK;
}

int quant_cols_tc_launch(const float* x, int64_t x_stride, int64_t K, int B, int pro, const float* gain, const float* nbias, float eps, const TcCols& out,
                         cudaStream_t st) {
    if (K > out.k_cap || B > out.cols_cap || K % 32) B2_FAIL(B2_ERR_ARG, "quant_cols_tc: K=%lld B=%d beyond the column buffer (%lld x %d)", (long long)K, B, (long long)out.k_cap, out.cols_cap);
    GemvArgs a;
    a.W.k = K; a.pro = pro; a.x = x; a.gain = gain; a.nbias = nbias; a.eps = eps;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(B, (unsigned)((K + TQ_SLICE - 1) / TQ_SLICE));
    cfg.blockDim = dim3(256);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = batch_pdl_enabled() ? 1 : 0;
    B2_CUDA(cudaLaunchKernelEx(&cfg, quant_cols_tc_kernel, a, x_stride, out.q, out.d, out.k_cap, out.cols_cap));
    return B2_OK;
}

// Y[:, c] = epilogue_c(W . q8col_c) for c < B.  acc32: ggml's generic f32 block-after-block accumulation instead of the
// canonical f64 one (faster; not available for tensor-parallel partial sums).
int gemm_tc_launch(const QMat& W, const TcMaps& maps, const TcCols& x, int B, const GemvArgs* d_colargs, const int* d_pos, const int* d_pt, int pt_stride,
                   cudaStream_t st, bool partial_out, bool acc32) {
    if (!gemm_tc_supported(W)) B2_FAIL(B2_ERR_UNSUPPORTED, "gemm_tc: type %d shape %lld x %lld has no tensor-core kernel", W.type, (long long)W.rows, (long long)W.k);
    if (B < 1 || B > x.cols_cap) B2_FAIL(B2_ERR_ARG, "gemm_tc: %d columns (1..%d)", B, x.cols_cap);
    if (W.k > x.k_cap) B2_FAIL(B2_ERR_ARG, "gemm_tc: K=%lld beyond the activation buffer (%lld)", (long long)W.k, (long long)x.k_cap);
    if (partial_out && acc32) B2_FAIL(B2_ERR_UNSUPPORTED, "gemm_tc: tensor-parallel partial sums need the f64 accumulation");
    const int rows = (int)W.rows, nbk = (int)(W.k / 32);
    if (acc32) {
        if (B > 32) return gemm_tc_launch_t<64, false>(maps, x, rows, nbk, B, d_colargs, d_pos, d_pt, pt_stride, st);
        return gemm_tc_launch_t<32, false>(maps, x, rows, nbk, B, d_colargs, d_pos, d_pt, pt_stride, st);
    }
    if (B > 32) return gemm_tc_launch_t<64, true>(maps, x, rows, nbk, B, d_colargs, d_pos, d_pt, pt_stride, st);
    return gemm_tc_launch_t<32, true>(maps, x, rows, nbk, B, d_colargs, d_pos, d_pt, pt_stride, st);
}

}  // namespace b2

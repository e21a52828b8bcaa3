// kernels.h -- internal interface between the engine (host C++) and the CUDA kernels.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "formats.cuh"

namespace b2 {

// ---- dequant.cu ---------------------------------------------------------------------------
int dequant_raw(int type, const void* d_blocks, int64_t n_elem, float* d_out, cudaStream_t st);
int repack(int type, const void* d_raw, int64_t rows, int64_t k, const QMat& dst, int64_t dst_row0, int64_t dst_row_step,
           cudaStream_t st);
int dequant_rows(const QMat& m, int64_t row0, int64_t nrows, float* d_out, cudaStream_t st);

// ---- gemv.cu: y = epilogue(W . quantise(prologue(x))) for one activation vector --------------
enum GemvPro : int { PRO_PLAIN = 0, PRO_RMSNORM = 1, PRO_LAYERNORM = 2 };
enum GemvEpi : int {
    EPI_STORE = 0,       // y[r] = v
    EPI_RESIDUAL = 1,    // y[r] += v            (y is the residual stream, updated in place)
    EPI_SILU_MUL = 2,    // rows interleaved (gate_i, up_i): y[i] = silu(v[2i]) * v[2i+1]
    EPI_QKV_ROPE = 3,    // rows [q | k | v]: rope on q,k; q -> y (f32), k,v -> f16 KV cache page
    EPI_LOGITS = 4,      // y[r] = v and a running arg-max into *argmax_key
    EPI_GELU = 5,        // y[r] = gelu(v)
    EPI_RESIDUAL2 = 6,   // y[r] = (v + y2[r]) + y[r]   (phi-2 parallel block: ffn + attn + x)
    EPI_PARTIAL = 7,     // y64[r] = the f64 sum, unrounded (tensor-parallel K slice)
    EPI_QKV_ROPE_GELU = 8,   // phi-2 parallel block: rows [q | k | v | ffn_up] share the normed input; the first
                             // n_q + 2 n_kv rows as EPI_QKV_ROPE, the rest as GELU into y3
};

struct KvDesc {                 // where the new token's K and V go (one layer, one sequence slot)
    __half* k_base = nullptr;   // [n_pages][page_tokens][kv_dim]
    __half* v_base = nullptr;
    const int* page_table = nullptr;   // logical page -> physical page, for this slot
    const int* pos = nullptr;          // device scalar: position of the token being decoded
    int page_tokens = 0, kv_dim = 0;
};

struct GemvArgs {
    QMat W;
    int pro = PRO_PLAIN, epi = EPI_STORE;
    const float* x = nullptr;      // [K]
    // persistent kernel only: hand-over copies of a vector passed from one stage to the next.  Such a buffer holds the
    // bit pattern LL_POISON (a NaN no arithmetic produces) in every word until the producing stage stores the value, so a
    // reader that polls the words needs no grid barrier.  Two or three buffers rotate; whoever writes row r of one restores
    // the poison in row r of the buffer of the previous write, whose readers are provably done by then (mega.cu has the argument).
    const float* x_ll = nullptr;   // [K] hand-over copy of x, written by the previous stage
    float* y_ll = nullptr;         // [rows] the same for this stage's output (EPI_RESIDUAL, EPI_SILU_MUL) ...
                                   // EPI_QKV_ROPE: [n_q floats | n_kv/2 words of K | n_kv/2 words of V] -- the query and, as half2
                                   // words, the new cache rows, for the attention stage that follows (AttnArgs::qkv_ll)
    float* y_ll_poison = nullptr;  // ... and the buffer of the alternate write, re-poisoned row by row
    const float* gain = nullptr;   // [K]  norm gain
    const float* nbias = nullptr;  // [K]  layer-norm bias
    float eps = 0.f;
    const float* bias = nullptr;   // [rows] optional, added before the epilogue
    float* y = nullptr;
    const float* y2 = nullptr;
    double* y64 = nullptr;         // EPI_PARTIAL
    float* y3 = nullptr;           // EPI_QKV_ROPE_GELU: the FFN hidden vector
    int64_t epi_row0 = 0;          // added to the matrix row before the epilogue (split QKV matrices)
    // EPI_QKV_ROPE
    int n_q = 0, n_kv = 0, head_dim = 0, n_rot = 0, rope_neox = 0;
    float q_scale = 1.f;
    const float2* rope_cs = nullptr;   // [max_ctx][n_rot/2] (cos, sin) table, built on the host
    KvDesc kv;
    // EPI_LOGITS
    unsigned long long* argmax_key = nullptr;
};

struct GemvTune { int nb = 0, g = 0, rpg = 0, ns = 0, ctas_per_sm = 0; };  // 0 = heuristic

int gemv_launch(const GemvArgs& a, cudaStream_t st, bool pdl, const GemvTune* tune = nullptr);
// bytes of weight data one launch streams (the algorithmic bytes of the kernel)
size_t gemv_weight_bytes(const QMat& W);

// ---- attn.cu: flash-decode over the paged f16 KV cache ---------------------------------------
struct AttnArgs {
    const float* q = nullptr;          // [n_head][head_dim] f32 (rounded to f16 inside, as ggml does)
    float* out = nullptr;              // [n_head][head_dim] f32
    float* out_ll = nullptr;           // persistent kernel: hand-over copy of out, and the alternate buffer to re-poison (see GemvArgs::x_ll)
    float* out_ll_poison = nullptr;
    const float* qkv_ll = nullptr;     // persistent kernel: hand-over copy of q and of the current position's K/V rows (GemvArgs::y_ll
                                       // of the fused QKV stage); the attention polls it instead of waiting at a grid barrier
    const __half* k_base = nullptr;    // [n_pages][page_tokens][kv_dim]
    const __half* v_base = nullptr;
    const int* page_table = nullptr;
    const int* pos = nullptr;          // attends to cache rows [0, *pos]
    int n_head = 0, n_head_kv = 0, head_dim = 0, page_tokens = 0, kv_dim = 0;
    int max_ctx = 0;                   // upper bound of *pos + 1 (sizes the score buffer)
    float scale = 1.f;
};
int attn_decode_launch(const AttnArgs& a, cudaStream_t st, bool pdl, int cluster = 0);

// ---- tp.cu: tensor-parallel all-reduce over peer memory (one process, one engine shard per GPU) ------------
constexpr int TP_MAX = 8;
struct TpPeers {
    int n = 1, rank = 0;
    double* part[2][TP_MAX] = {};     // [buffer][device]: every shard's partial sums (peer-mapped pointers)
    unsigned* flags[TP_MAX] = {};     // [device]: flags[d][r] = last epoch shard r published to device d
    unsigned* epoch = nullptr;        // local: epochs completed so far
    unsigned* ticket = nullptr;       // local: CTA exit counter of the reduce kernel
};
// x[i] += (float)(sum over shards of part[buf][d][i]), the same bits on every shard (fixed order, f64):
// publishes this shard's epoch to all peers, waits for theirs, then reads their partial sums over NVLink.
int tp_allreduce_residual_launch(const TpPeers& p, int buf, float* x, int n, cudaStream_t st, bool pdl);

// ---- mega.cu: the persistent decode kernel (all layers of a step in one launch) ---------------------
#ifndef B2_MEGA_NCW
#define B2_MEGA_NCW 15
#endif
constexpr int MEGA_NCW = B2_MEGA_NCW;      // consumer warps per CTA
struct alignas(16) MegaStage {    // one stage of the layer pipeline, resident in device memory (copied to shared memory by 16-byte bulk copies)
    int kind = 0;                 // 0 = GEMV, 1 = attention, 2 = tensor-parallel exchange (all-reduce + residual)
    GemvArgs g;
    int S = 1;                    // warps along K per row
    int P = 2;                    // warps sharing a row pair (each takes every P-th row)
    int G = 1;                    // row groups per CTA (G * S * P <= MEGA_NCW)
    int RPG = 2, RT = 2;          // rows per group per tile, rows per tile (= 2 * PT)
    int PT = 1;                   // row pairs per tile (one ring slot); PT * S owner warps release a slot
    int nunits = 0, nbeff = 4;    // units per row, units per lane actually used
    // filled in by mega_finish_stage once the grid is known, so that the kernel divides nothing per stage:
    int pairs_base = 0, pairs_rem = 0;   // row pairs per CTA; the first pairs_rem CTAs take one more
    int inv_S = 65536, inv_PT = 65536;   // ceil(2^16 / S), ceil(2^16 / PT): x / S == (x * inv_S) >> 16 for the small x used
    uint32_t arrive_units = 0;           // MEGA_EMPTY_BASE / (G * S): every working warp arrives on a tile's `empty` barrier
    int ll_in = 0;                       // the input vector arrives through a hand-over buffer (g.x_ll): poll it instead of waiting at the grid barrier
    int next_attn = 0;                   // an attention stage follows; `at` holds its arguments (for the KV prefetch)
    int epi_cache = 0;                   // 1: EPI_RESIDUAL on the residual stream, 2: the fused QKV matrix -- per-thread state kept across layers (gemvdev::EpiCache)
    AttnArgs at;
    int att_max_local = 0;
    int tp_buf = 0, tp_n = 0;     // kind 2: which partial-sum buffer, vector length
    float* tp_x = nullptr;        //         the residual stream the sum is added to
};
struct MegaLaunch {
    const MegaStage* stages = nullptr;   // device array
    int s_begin = 0, s_end = 0;
    unsigned* grid_count = nullptr;      // zeroed before every launch
    struct LlInit { float* p[6] = {}; int n[6] = {}; };
    LlInit ll_init;                      // hand-over buffers the first writers of a launch use: poisoned by the kernel before it (step_begin)
    unsigned long long* trace = nullptr; // optional [grid][stages][4] globaltimer stamps (debug)
    const int* pos = nullptr;            // device scalar: position of the token being decoded
    const int* page_table = nullptr;     // this slot's page table
    int n_pages = 0;
    int type = 0, grp = 1, head_dim = 0, cluster = 4, grid = 0;
    int nb = 4;                          // units of the activation vector a lane keeps in registers (kernel variant)
    uint32_t off_red = 0, off_xq = 0, off_xd = 0, off_xs = 0, off_att = 0, off_ring = 0, stage_bytes = 0;
    int ns = 0, att_max_local = 0;
    uint32_t pf_bytes = 0;               // L2 prefetch run-ahead per CTA, bytes (0 = off)
    uint32_t ring_room = 0;              // shared memory left for the ring after the other buffers (set by mega_configure)
    int debug = 0;                       // tuning experiments (see mega.cu), results are wrong when set
    size_t smem_bytes = 0;
    TpPeers tp;                          // peers of the exchange stages (tp.n == 1: none)
};
bool mega_supported(int type, int grp, int head_dim);
void mega_finish_stage(MegaStage& st, int grid);
int mega_plan_stage(MegaStage& st, int ncw, int nb_max, int tile_target = 0);      // tile_target: bytes per weight tile (0 = default)
double mega_plan_cost(const QMat& W, int nb_max);      // estimated consumer cost of a matrix with at most nb_max units per lane
int mega_configure(MegaLaunch& L, int64_t k_max, size_t tile_bytes_max, int rt_s_max, int max_ctx);
int mega_launch(const MegaLaunch& L, cudaStream_t stream);

// ---- step.cu: per-step bookkeeping kernels ----------------------------------------------------
struct SeqState {        // one per sequence slot, resident on the device
    int pos;             // position of the token about to be evaluated
    int cur_tok;         // that token
    int n_forced;        // tokens still to be taken from `forced` (prompt) instead of arg-max
    int forced_idx;
    int n_out;           // tokens emitted into out_toks so far
    int pad[3];
};
struct StepBeginArgs {
    SeqState* st = nullptr;
    QMat embd;
    float* x = nullptr;            // [n_embd]
    int* pos_out = nullptr;        // device scalar consumed by the layer kernels
    unsigned long long* argmax_key = nullptr;
    unsigned* grid_count = nullptr;   // persistent kernel's barrier counter, cleared every step (may be null)
    MegaLaunch::LlInit ll_init;       // ... and the hand-over buffers its first writers use, poisoned every step
    const SeqState* st_host = nullptr;   // when set: the step's input state, read straight from pinned host memory (no copy node)
};
int step_begin_launch(const StepBeginArgs& a, cudaStream_t st, bool pdl);
int delay_launch(cudaStream_t st, unsigned long long ns);      // a kernel that waits `ns` (profiling aid)
struct StepEndArgs {
    SeqState* st = nullptr;
    const unsigned long long* argmax_key = nullptr;
    const int* forced = nullptr;   // prompt tokens
    int* out_toks = nullptr;       // ring of emitted tokens
    int out_cap = 0;
    int use_argmax = 1;            // 0: prompt step, the head did not run
    SeqState* st_host = nullptr;   // when set: the resulting state is also written to pinned host memory (no copy node)
};
int step_end_launch(const StepEndArgs& a, cudaStream_t st, bool pdl);

}  // namespace b2

// ---- batch.cu / gemm.cu: several columns (tokens) per weight pass: batched decode and chunked prefill ----
namespace b2 {
// Programmatic dependent launch for the batched-step kernels (quantise, GEMM, attention): on unless the model was
// loaded with B2_LOAD_NO_PDL; the engine sets it before it enqueues a batched step.
bool& batch_pdl_enabled();
constexpr int BATCH_MAX = 32;          // columns per pass (four 8-column MMA tiles)
constexpr int GEMM_KC = 16;            // 32-weight blocks per K chunk (one pipeline stage of the batched kernel)
constexpr int GEMM_XCOL = 2 * GEMM_KC * 16 + 16;   // bytes of one column's K chunk, padded (conflict-free B-fragment loads)
// Q8_0-quantised activation columns in global memory, chunk-major so that one K chunk of all columns
// is one contiguous range already in the batched kernel's shared-memory layout:
//   q  [n_chunks][BATCH_MAX][GEMM_XCOL]: per column and chunk two planes of 16 bytes per block
//                                         (elements 0..15 | 16..31 of each block), then 16 bytes of padding
//   d  [n_chunks][GEMM_KC][BATCH_MAX]:   f16-rounded scales;  s8: same shape, -8 * sum(q) per block
struct Q8Cols {
    uint8_t* q = nullptr;
    float* d = nullptr;
    int* s8 = nullptr;
    int64_t k_cap = 0;         // capacity in elements per column
};
inline int64_t q8cols_chunks(int64_t k) { return (k / 32 + GEMM_KC - 1) / GEMM_KC; }
inline size_t q8cols_q_bytes(int64_t k) { return (size_t)q8cols_chunks(k) * BATCH_MAX * GEMM_XCOL; }
inline size_t q8cols_d_bytes(int64_t k) { return (size_t)q8cols_chunks(k) * GEMM_KC * BATCH_MAX * 4; }
struct alignas(64) GemmMaps { CUtensorMap q, d; };     // TMA descriptors of one Q4_0 matrix: nibble plane, scale plane
// x[c] = embedding row of tok[c]; arg-max accumulators cleared
int batch_begin_launch(const QMat& embd, float* xb, int64_t x_stride, const int* d_tok, unsigned long long* keys, int B, cudaStream_t st);
// norm (pro) + Q8_0 quantisation of B activation columns x[c*x_stride ..] of length K
int quant_cols_launch(const float* x, int64_t x_stride, int64_t K, int B, int pro, const float* gain, const float* nbias, float eps, const Q8Cols& out,
                      cudaStream_t st);
// Y[:, c] = epilogue_c(W . q8col_c) for c < B: integer tensor-core MMAs per 32-weight block, exact, with
// the same per-block float terms and f64 accumulation as the GEMV (bit-identical results per column).
// colargs: device array [B] of GemvArgs (epilogue operands of each column); pos / page tables per column.
bool gemm_supported(const QMat& W);
int gemm_make_maps(const QMat& W, GemmMaps* out);
int gemm_w4a8_launch(const QMat& W, const GemmMaps& maps, const Q8Cols& x, int B, const GemvArgs* d_colargs, const int* d_pos, const int* d_pt, int pt_stride,
                     cudaStream_t st, bool partial_out = false);      // partial_out: the columns' epilogue is EPI_PARTIAL
// attention for B columns: d_args[c] describes column c (q, out, page table, position pointer)
int attn_decode_batch_launch(const AttnArgs* d_args, const AttnArgs& shape, int B, cudaStream_t st);
// ---- gemm_tc.cu: the same product on tcgen05 tensor cores (kind::i8, TMEM accumulators), up to 128 columns per pass ----
constexpr int TC_COLS_MAX = 512;      // columns of one pass (a prompt chunk); a CTA takes 32, 64 or 128 of them
struct alignas(64) TcMaps { CUtensorMap q, d; };      // TMA descriptors of one Q4_0 matrix: nibble plane (swizzled 128 x 128 B boxes), scale plane
struct alignas(64) TcCols {                           // Q8_0-quantised activation columns, K-major: q [cols_cap][k_cap] s8, d [k_cap / 32][cols_cap] f32
    CUtensorMap map32, map64, map128;                 // boxes of 32 / 64 / 128 columns x 128 bytes of K, 128-byte swizzle
    int8_t* q = nullptr;
    float* d = nullptr;
    int64_t k_cap = 0;
    int cols_cap = 0;
};
bool gemm_tc_supported(const QMat& W);
int gemm_tc_make_maps(const QMat& W, TcMaps* out);
int tc_cols_init(TcCols& c, int8_t* q, float* d, int64_t k_cap, int cols_cap);
int quant_cols_tc_launch(const float* x, int64_t x_stride, int64_t K, int B, int pro, const float* gain, const float* nbias, float eps, const TcCols& out,
                         cudaStream_t st);
int gemm_tc_launch(const QMat& W, const TcMaps& maps, const TcCols& x, int B, const GemvArgs* d_colargs, const int* d_pos, const int* d_pt, int pt_stride,
                   cudaStream_t st, bool partial_out = false, bool acc32 = false);
// next token of every column from its arg-max accumulator
int batch_end_launch(const unsigned long long* keys, int* d_next, int B, cudaStream_t st);
}  // namespace b2

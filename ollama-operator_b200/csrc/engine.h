// engine.h -- host-side model / sequence state behind the C ABI.
#pragma once
#include <string>
#include <vector>

#include "../../include/b2gguf.h"
#include "common.cuh"
#include "formats.cuh"
#include "gguf.h"
#include "kernels.h"

namespace b2 {

constexpr int PAGE_TOKENS = 64;    // tokens per KV-cache page

struct Arena {                     // one cudaMalloc that all resident weights are carved from
    void* base = nullptr;
    size_t cap = 0, used = 0;
    int reserve(size_t bytes);
    void* take(size_t bytes);
    void release();
};

struct HParams {
    int arch = 0;                  // 0 llama, 1 phi2
    int n_layer = 0, n_embd = 0, n_head = 0, n_head_kv = 0, head_dim = 0, n_ff = 0, n_vocab = 0, n_ctx = 0, n_rot = 0;
    float norm_eps = 1e-5f, rope_freq_base = 10000.f;
};

struct QkvSeg { QMat W; int64_t row0 = 0; };     // a run of q|k|v rows that share a tensor type

struct LayerW {
    std::vector<QkvSeg> qkv;
    QMat w_o, w_up, w_down;        // llama: w_up holds gate and up interleaved row by row
    float *attn_norm = nullptr, *attn_norm_b = nullptr, *ffn_norm = nullptr;
    float *b_qkv = nullptr, *b_o = nullptr, *b_up = nullptr, *b_down = nullptr;
    bool fused_up = false;         // phi-2: ffn_up rows follow q|k|v in qkv[0].W (one stage for the parallel block's shared input), b_qkv covers both
};

struct StepOp { int kind = 0; int cls = 0; GemvArgs g; AttnArgs a; int tp_buf = -1; };     // kind 0 = GEMV, 1 = attention; tp_buf >= 0: all-reduce follows

struct Slot {
    bool used = false;
    int len = 0;                   // tokens in the KV cache
    std::vector<int> pages;        // physical pages, in logical order
    bool table_synced = false;
    SeqState host{};               // mirror of the device state after the last call
    SeqState* d_state = nullptr;
    int* d_page_table = nullptr;
    float *x = nullptr, *q = nullptr, *att = nullptr, *att2 = nullptr, *h = nullptr, *logits = nullptr;
    int* d_pos = nullptr;
    unsigned long long* d_key = nullptr;
    int *d_out = nullptr, *d_forced = nullptr;
    unsigned* d_grid_count = nullptr;
    float *x_ll[3] = {}, *att_ll[2] = {}, *h_ll[2] = {}, *qkv_ll[2] = {};      // hand-over copies of x, att and h for the persistent kernel (GemvArgs::x_ll)
    std::vector<StepOp> ops;       // the layer pipeline with this slot's buffers bound
    bool mega_ok = false;          // the persistent kernel can run this model
    MegaLaunch mega;
    MegaStage* d_stages = nullptr;
    cudaGraphExec_t g_full = nullptr, g_nohead = nullptr;
    cudaGraphExec_t g_decode = nullptr;      // b2_decode of this one sequence: token in, step, state out as ONE graph launch
    SeqState *zc_in = nullptr, *zc_out = nullptr;      // set while that graph is captured: state read from / written to pinned host memory by the step kernels
    int gl_full = 0, gl_nohead = 0;
    bool warm_full = false, warm_nohead = false;
};

// Batched steps (several sequences per decode call, or a chunk of prompt tokens): one set of
// column buffers for the whole model, per-column operand tables resident on the device.
struct BatchCtx {
    bool ok = false;                 // every layer matrix has a batched kernel
    void* mem = nullptr;
    float *x = nullptr, *q = nullptr, *att = nullptr, *att2 = nullptr, *h = nullptr, *logits = nullptr;   // [BATCH_MAX][...]
    Q8Cols q8;
    int cap = BATCH_MAX;             // columns per step the buffers hold: BATCH_MAX, or TC_COLS_MAX when the tensor-core path serves the model
    bool tc_ok = false;              // chunks of more than BATCH_MAX columns go through gemm_tc.cu (tcgen05)
    bool tc_acc32 = false;           // ... with ggml's generic f32 accumulation order instead of the canonical f64 one
    TcCols tc;
    std::vector<TcMaps> tcmaps;
    int *d_tok = nullptr, *d_pos = nullptr, *d_pt = nullptr, *d_next = nullptr;    // d_pt: [BATCH_MAX][pages_per_seq]
    unsigned long long* d_keys = nullptr;
    GemvArgs* d_colargs = nullptr;   // [ops][BATCH_MAX]
    AttnArgs* d_attargs = nullptr;   // [ops][BATCH_MAX] (attention ops only)
    std::vector<StepOp> ops;         // the layer pipeline bound to column 0
    std::vector<GemmMaps> maps;      // TMA descriptors of every matrix of the pipeline
    int* h_stage = nullptr;          // pinned upload regions: tok[BATCH_MAX] | pos[BATCH_MAX] | page tables
    int n_regions = 0, region_ints = 0, next_region = 0;
    int* h_next = nullptr;           // pinned [BATCH_MAX]
};

struct StepProfile { std::vector<cudaEvent_t> ev; std::vector<int> cls; };

struct Model {
    HParams hp;                       // the whole model's hyper-parameters
    int tp_rank = 0, tp_size = 1;     // this shard of a tensor-parallel group (one shard per GPU)
    int n_head_l = 0, n_head_kv_l = 0, n_ff_l = 0;     // heads / FFN rows held by this shard
    TpPeers tp, tpb;                  // peer-mapped partial-sum buffers and flags (tp_size > 1): single-token steps, batched steps
    void* tp_mem = nullptr;
    int device = 0;
    bool tp_skip_reduce = false;
    bool neox = false, use_pdl = true, use_graph = true, use_mega = true, use_batch = true;
    int max_batch = 1, max_ctx = 0, pages_per_seq = 0, n_pages = 0, out_cap = 0;
    float2* rope_table = nullptr;     // [max_ctx][n_rot/2]
    Arena arena;
    void* staging = nullptr;
    QMat w_embd, w_head;
    float *out_norm = nullptr, *out_norm_b = nullptr, *head_b = nullptr;
    std::vector<LayerW> layers;
    __half *kv_k = nullptr, *kv_v = nullptr;
    size_t layer_kv_elems = 0;
    std::vector<int> free_pages;
    void* slot_mem = nullptr;
    SeqState* pin_state = nullptr;     // pinned host staging: [slot][up, down]
    std::vector<Slot> slots;
    BatchCtx bt;
    int batch_min_seqs = 3;           // b2_decode switches to the batched kernels from this many sequences
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    size_t weight_bytes = 0, step_weight_bytes = 0;
    uint64_t steps = 0, launches = 0, last_steps = 0, last_launches = 0;
    double last_ms = 0;

    ~Model();
    int load(const char* path, const b2_load_opts* o, int tp_rank = 0, int tp_size = 1);
    int tp_alloc();                                     // this shard's exchange buffers (before tp_connect)
    int tp_connect(Model* const* shards);               // peer-map every shard's buffers, then build the step pipelines
    int finish_load();
    int seq_new(int* out);
    int seq_free(int seq);
    int seq_truncate(int seq, int n_keep);
    int seq_assume(int seq, int n);
    int check_seq(int seq) const;
    int prefill(int seq, const int32_t* toks, int n, float* logits);
    int decode(int n_seq, const int* seqs, const int32_t* last_tok, int32_t* next_tok, float* logits);
    int decode_greedy(int seq, int32_t first_tok, int n_steps, int32_t* out_toks);
    int kv_read(int seq, int layer, uint16_t* k_out, uint16_t* v_out, int n_tok);
    int profile_step(int seq, int32_t tok, float* class_ms, int* class_launches);
    int mega_trace(int seq, int32_t tok, unsigned long long* out, int cap, int* n_cta, int* n_stage);

  private:
    int upload_quant(const GgufTensor* t, int64_t rows, int64_t k, const QMat& dst, int64_t dst_row0, int64_t dst_row_step, int64_t src_row0);
    int upload_quant_kslice(const GgufTensor* t, int64_t rows, int64_t k_full, int64_t k0, int64_t kl, const QMat& dst);
    int upload_f32(const GgufFile& f, const std::string& name, int64_t n, float** out, bool optional);
    int upload_rows_headwise(const GgufTensor* t, int64_t n_heads, int64_t k, const QMat& dst, int64_t dst_row0, int64_t src_row0);
    void build_ops(Slot& s);
    int setup_mega(Slot& s);
    int ensure_pages(Slot& s, int upto);
    int enqueue_step(Slot& s, bool with_head, int* n_launch, StepProfile* prof);
    int run_steps(Slot& s, bool with_head, int n);
    int setup_batch();
    int batch_step(int B, const int* col_slot, const int* col_pos, const int32_t* col_tok, int head_from);
    int begin_call();
    int end_call(Slot* s);
};

}  // namespace b2

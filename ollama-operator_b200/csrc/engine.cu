// engine.cu -- model residency, paged KV cache, sequence slots and the decode-step schedule.
//
// This is the part of the `ollama serve` process (the container of reference
// pkg/model/pod.go:14-66) that sits between the HTTP handlers and the kernels: llama.cpp's
// llama_decode over build_llama / build_phi2 graphs [upstream], rebuilt around CUDA graphs:
// one decode step is a fixed chain of ~5 kernels per layer (fused norm+QKV+RoPE+KV-append,
// cluster attention, wo+residual, norm+gate/up+SiLU, down+residual), captured once per
// sequence slot and replayed; token feedback, position and RoPE table stay on the device.
#include "engine.h"

#include <math.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>

namespace b2 {

// ---- error state ----------------------------------------------------------------------------
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
const char* last_error() { return g_err; }

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// ---- device arena -----------------------------------------------------------------------------
int Arena::reserve(size_t bytes) {
    B2_CUDA(cudaMalloc(&base, bytes));
    cap = bytes; used = 0;
    return B2_OK;
}
void* Arena::take(size_t bytes) {
    const size_t off = align_up(used, 256);
    if (off + bytes > cap) return nullptr;
    used = off + bytes;
    return (uint8_t*)base + off;
}
void Arena::release() {
    if (base) cudaFree(base);
    base = nullptr; cap = used = 0;
}

static size_t qmat_arena_bytes(int type, int64_t rows, int64_t k) {
    QMat m;
    qmat_shape(m, type, rows, k);
    size_t s = 0;
    for (int p = 0; p < m.n_planes; p++) s += align_up((size_t)rows * m.stride[p], 256);
    return s + 256;
}
static int qmat_alloc(Arena& ar, QMat& m, int type, int64_t rows, int64_t k) {
    qmat_shape(m, type, rows, k);
    for (int p = 0; p < m.n_planes; p++) {
        m.plane[p] = (uint8_t*)ar.take((size_t)rows * m.stride[p]);
        if (!m.plane[p]) B2_FAIL(B2_ERR_NOMEM, "weight arena exhausted");
    }
    return B2_OK;
}

// ---- Model ---------------------------------------------------------------------------------------
Model::~Model() {
    for (auto& s : slots) {
        if (s.g_full) cudaGraphExecDestroy(s.g_full);
        if (s.g_nohead) cudaGraphExecDestroy(s.g_nohead);
        if (s.g_decode) cudaGraphExecDestroy(s.g_decode);
        if (s.d_stages) cudaFree(s.d_stages);
    }
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
    if (stream) cudaStreamDestroy(stream);
    if (staging) cudaFree(staging);
    if (kv_k) cudaFree(kv_k);
    if (kv_v) cudaFree(kv_v);
    if (slot_mem) cudaFree(slot_mem);
    if (tp_mem) cudaFree(tp_mem);
    if (pin_state) cudaFreeHost(pin_state);
    if (bt.mem) cudaFree(bt.mem);
    if (bt.h_stage) cudaFreeHost(bt.h_stage);
    if (bt.h_next) cudaFreeHost(bt.h_next);
    if (rope_table) cudaFree(rope_table);
    arena.release();
}

struct TensorPlan { const GgufTensor* t; int64_t rows, k; };

static int need(const GgufFile& f, const std::string& name, int64_t rows, int64_t k, TensorPlan* out, bool quant_ok = true) {
    const GgufTensor* t = f.find(name);
    if (!t) B2_FAIL(B2_ERR_FORMAT, "gguf: tensor %s missing", name.c_str());
    const int64_t tk = t->ne.size() >= 1 ? t->ne[0] : 0, tr = t->ne.size() >= 2 ? t->ne[1] : 1;
    if (tk != k || tr != rows) B2_FAIL(B2_ERR_FORMAT, "gguf: tensor %s has shape [%lld x %lld], expected [%lld x %lld]", name.c_str(),
                                       (long long)tr, (long long)tk, (long long)rows, (long long)k);
    if (quant_ok ? !type_is_quant(t->type) : t->type != T_F32)
        B2_FAIL(B2_ERR_UNSUPPORTED, "gguf: tensor %s has ggml type %d (%s expected)", name.c_str(), t->type,
                quant_ok ? "Q4_0/Q8_0/Q4_K/Q6_K" : "F32");
    out->t = t; out->rows = rows; out->k = k;
    return B2_OK;
}

// upload one file tensor into the staging buffer and repack it into `dst`
int Model::upload_quant(const GgufTensor* t, int64_t rows, int64_t k, const QMat& dst, int64_t dst_row0, int64_t dst_row_step,
                        int64_t src_row0) {
    const size_t rb = type_row_bytes(t->type, k);
    B2_CUDA(cudaMemcpyAsync(staging, t->data + (size_t)src_row0 * rb, (size_t)rows * rb, cudaMemcpyHostToDevice, stream));
    B2_TRY(repack(t->type, staging, rows, k, dst, dst_row0, dst_row_step, stream));
    B2_CUDA(cudaStreamSynchronize(stream));     // staging is reused by the next tensor
    return B2_OK;
}
// Tensor-parallel K slice: columns [k0, k0 + kl) of every row (whole blocks), gathered by a strided copy
int Model::upload_quant_kslice(const GgufTensor* t, int64_t rows, int64_t k_full, int64_t k0, int64_t kl, const QMat& dst) {
    const int be = type_block_elems(t->type);
    if (k0 % be || kl % be) B2_FAIL(B2_ERR_UNSUPPORTED, "tensor-parallel split of K=%lld at %lld does not fall on a %d-element block boundary", (long long)k_full, (long long)k0, be);
    const size_t rb_full = type_row_bytes(t->type, k_full), rb = type_row_bytes(t->type, kl), off = type_row_bytes(t->type, k0);
    B2_CUDA(cudaMemcpy2DAsync(staging, rb, t->data + off, rb_full, rb, (size_t)rows, cudaMemcpyHostToDevice, stream));
    B2_TRY(repack(t->type, staging, rows, kl, dst, 0, 1, stream));
    B2_CUDA(cudaStreamSynchronize(stream));
    return B2_OK;
}
int Model::upload_f32(const GgufFile& f, const std::string& name, int64_t n, float** out, bool optional) {
    const GgufTensor* t = f.find(name);
    *out = nullptr;
    if (!t) {
        if (optional) return B2_OK;
        B2_FAIL(B2_ERR_FORMAT, "gguf: tensor %s missing", name.c_str());
    }
    int64_t tn = 1;
    for (int64_t d : t->ne) tn *= d;
    if (t->type != T_F32 || tn != n) B2_FAIL(B2_ERR_FORMAT, "gguf: tensor %s must be F32 with %lld elements", name.c_str(), (long long)n);
    float* d = (float*)arena.take((size_t)n * 4);
    if (!d) B2_FAIL(B2_ERR_NOMEM, "weight arena exhausted");
    B2_CUDA(cudaMemcpy(d, t->data, (size_t)n * 4, cudaMemcpyHostToDevice));
    *out = d;
    return B2_OK;
}

// NeoX rope pairs element p with p + n_rot/2.  The loader stores the rows of every q/k/v head so
// that partners are adjacent (stored 2p <- p, 2p+1 <- p + n_rot/2), which lets the GEMV
// epilogue rotate a pair held by one thread; see EPI_QKV_ROPE in gemv.cu.
int Model::upload_rows_headwise(const GgufTensor* t, int64_t n_heads, int64_t k, const QMat& dst, int64_t dst_row0, int64_t src_row0) {
    const int hd = hp.head_dim, nr = hp.n_rot;
    if (!neox) return upload_quant(t, n_heads * hd, k, dst, dst_row0, 1, src_row0);
    for (int64_t h = 0; h < n_heads; h++) {
        const int64_t s = src_row0 + h * hd, d = dst_row0 + h * hd;
        B2_TRY(upload_quant(t, nr / 2, k, dst, d, 2, s));
        B2_TRY(upload_quant(t, nr / 2, k, dst, d + 1, 2, s + nr / 2));
        if (hd > nr) B2_TRY(upload_quant(t, hd - nr, k, dst, d + nr, 1, s + nr));
    }
    return B2_OK;
}

int Model::load(const char* path, const b2_load_opts* o, int rank, int size) {
    tp_rank = rank; tp_size = size;
    B2_CUDA(cudaGetDevice(&device));
    GgufFile f;
    B2_TRY(f.open(path));
    std::string arch;
    if (!f.get_str("general.architecture", &arch)) B2_FAIL(B2_ERR_FORMAT, "gguf: general.architecture missing");
    if (arch == "llama") hp.arch = 0;
    else if (arch == "phi2") hp.arch = 1;
    else B2_FAIL(B2_ERR_UNSUPPORTED, "architecture '%s' not supported (llama, phi2)", arch.c_str());
    neox = hp.arch == 1;
    auto u = [&](const char* key, int* dst, bool required, int dflt) -> int {
        uint64_t v;
        if (f.get_u64(arch + "." + key, &v)) { *dst = (int)v; return B2_OK; }
        if (required) B2_FAIL(B2_ERR_FORMAT, "gguf: key %s.%s missing", arch.c_str(), key);
        *dst = dflt;
        return B2_OK;
    };
    B2_TRY(u("block_count", &hp.n_layer, true, 0));
    B2_TRY(u("embedding_length", &hp.n_embd, true, 0));
    B2_TRY(u("feed_forward_length", &hp.n_ff, true, 0));
    B2_TRY(u("attention.head_count", &hp.n_head, true, 0));
    B2_TRY(u("attention.head_count_kv", &hp.n_head_kv, false, 0));
    B2_TRY(u("context_length", &hp.n_ctx, false, 2048));
    if (hp.n_head_kv <= 0) hp.n_head_kv = hp.n_head;
    if (hp.n_layer <= 0 || hp.n_embd <= 0 || hp.n_head <= 0 || hp.n_ff <= 0 || hp.n_embd % hp.n_head)
        B2_FAIL(B2_ERR_FORMAT, "gguf: implausible hyper-parameters");
    hp.head_dim = hp.n_embd / hp.n_head;
    B2_TRY(u("rope.dimension_count", &hp.n_rot, false, hp.head_dim));
    if (hp.n_rot <= 0 || hp.n_rot > hp.head_dim || (hp.n_rot & 1) || (hp.head_dim & 1)) B2_FAIL(B2_ERR_UNSUPPORTED, "rope dimension count %d", hp.n_rot);
    double d;
    hp.norm_eps = 1e-5f;
    if (f.get_f64(arch + ".attention.layer_norm_rms_epsilon", &d) || f.get_f64(arch + ".attention.layer_norm_epsilon", &d)) hp.norm_eps = (float)d;
    hp.rope_freq_base = 10000.f;
    if (f.get_f64(arch + ".rope.freq_base", &d)) hp.rope_freq_base = (float)d;
    const GgufTensor* te = f.find("token_embd.weight");
    if (!te || te->ne.size() != 2 || te->ne[0] != hp.n_embd) B2_FAIL(B2_ERR_FORMAT, "gguf: token_embd.weight missing or mis-shaped");
    hp.n_vocab = (int)te->ne[1];

    max_batch = o && o->max_batch > 0 ? o->max_batch : 1;
    max_ctx = o && o->max_ctx > 0 ? std::min(o->max_ctx, hp.n_ctx) : hp.n_ctx;
    {   // The decode attention keeps the scores of every cached position (per query head of a KV group) plus a row-offset
        // table in shared memory: (grp + 1) * 4 bytes per position out of ~200 KB.  Refuse a context it cannot serve at
        // load time (and before the KV cache is allocated) instead of failing every later call; a file that only
        // advertises a long context (Llama-3.1: 131072) is clamped when the caller did not ask for a size.
        const int grp_q = hp.n_head / hp.n_head_kv;
        const int ctx_limit = (int)((196 * 1024) / ((size_t)(grp_q + 1) * 4)) / 64 * 64;
        if (o && o->max_ctx > 0) {
            if (max_ctx > ctx_limit)
                B2_FAIL(B2_ERR_UNSUPPORTED, "max_ctx %d is beyond what the decode attention serves for %d query heads per KV head (%d tokens)", max_ctx, grp_q, ctx_limit);
        } else {
            max_ctx = std::min(max_ctx, std::min(ctx_limit, 8192));
        }
    }
    if (tp_size > 1) {
        if (hp.arch != 0) B2_FAIL(B2_ERR_UNSUPPORTED, "tensor parallelism is built for the llama graph only");
        if (hp.n_head % tp_size || hp.n_head_kv % tp_size || hp.n_ff % tp_size || (hp.n_ff / tp_size) % 32)
            B2_FAIL(B2_ERR_UNSUPPORTED, "tp_size=%d does not divide %d heads / %d KV heads / n_ff %d into whole blocks", tp_size, hp.n_head, hp.n_head_kv, hp.n_ff);
    }
    n_head_l = hp.n_head / tp_size; n_head_kv_l = hp.n_head_kv / tp_size; n_ff_l = hp.n_ff / tp_size;
    const int flags = o ? o->flags : 0;
    use_pdl = !(flags & B2_LOAD_NO_PDL) && !getenv("B2_NO_PDL");
    use_graph = !(flags & B2_LOAD_NO_GRAPH) && !getenv("B2_NO_GRAPH");
    use_mega = !(flags & B2_LOAD_NO_MEGA) && !getenv("B2_NO_MEGA");
    use_batch = !(flags & B2_LOAD_NO_BATCH) && !getenv("B2_NO_BATCH");
    tp_skip_reduce = getenv("B2_TP_SKIP_REDUCE") != nullptr;      // timing experiment only: results are wrong

    const int E = hp.n_embd, hd = hp.head_dim, nq_full = hp.n_head * hd, nkv_full = hp.n_head_kv * hd, F_full = hp.n_ff, V = hp.n_vocab;
    const int nq = n_head_l * hd, nkv = n_head_kv_l * hd, F = n_ff_l;       // this shard's part
    if ((nq | nkv | F | V | E) & 1) B2_FAIL(B2_ERR_UNSUPPORTED, "odd matrix dimensions are not supported");

    // ---- plan: which file tensors feed which resident matrices ---------------------------------
    struct LayerPlan { TensorPlan q, k, v, qkv, o, gate, up, down; };
    std::vector<LayerPlan> lp(hp.n_layer);
    TensorPlan embd, head;
    B2_TRY(need(f, "token_embd.weight", V, E, &embd));
    const bool tied = f.find("output.weight") == nullptr;
    B2_TRY(need(f, tied ? "token_embd.weight" : "output.weight", V, E, &head));
    size_t arena_bytes = qmat_arena_bytes(embd.t->type, V, E) + (tied ? 0 : qmat_arena_bytes(head.t->type, V, E));
    size_t max_raw = std::max(embd.t->bytes, head.t->bytes);
    size_t f32_bytes = (size_t)(2 * E + V) * 4 + 1024;
    for (int l = 0; l < hp.n_layer; l++) {
        const std::string p = "blk." + std::to_string(l) + ".";
        LayerPlan& L = lp[l];
        if (hp.arch == 0) {
            B2_TRY(need(f, p + "attn_q.weight", nq_full, E, &L.q));
            B2_TRY(need(f, p + "attn_k.weight", nkv_full, E, &L.k));
            B2_TRY(need(f, p + "attn_v.weight", nkv_full, E, &L.v));
            B2_TRY(need(f, p + "ffn_gate.weight", F_full, E, &L.gate));
            arena_bytes += qmat_arena_bytes(L.q.t->type, nq, E) + qmat_arena_bytes(L.k.t->type, nkv, E) + qmat_arena_bytes(L.v.t->type, nkv, E);
            arena_bytes += qmat_arena_bytes(L.gate.t->type, 2 * (int64_t)F, E);
            max_raw = std::max({max_raw, L.q.t->bytes, L.k.t->bytes, L.v.t->bytes, L.gate.t->bytes});
        } else {
            B2_TRY(need(f, p + "attn_qkv.weight", nq + 2 * nkv, E, &L.qkv));
            arena_bytes += qmat_arena_bytes(L.qkv.t->type, nq + 2 * nkv, E);
            max_raw = std::max(max_raw, L.qkv.t->bytes);
        }
        B2_TRY(need(f, p + "attn_output.weight", E, nq_full, &L.o));
        B2_TRY(need(f, p + "ffn_up.weight", F_full, E, &L.up));
        B2_TRY(need(f, p + "ffn_down.weight", E, F_full, &L.down));
        if (hp.arch == 0 && L.gate.t->type != L.up.t->type) B2_FAIL(B2_ERR_UNSUPPORTED, "layer %d: ffn_gate and ffn_up differ in type", l);
        if (hp.arch == 1) arena_bytes += qmat_arena_bytes(L.up.t->type, F, E);
        arena_bytes += qmat_arena_bytes(L.o.t->type, E, nq) + qmat_arena_bytes(L.down.t->type, E, F);
        max_raw = std::max({max_raw, L.o.t->bytes, L.up.t->bytes, L.down.t->bytes});
        f32_bytes += (size_t)(4 * E + 2 * (nq + 2 * nkv) + 2 * F) * 4 + 4096;
    }
    B2_TRY(arena.reserve(arena_bytes + f32_bytes + (1 << 20)));
    B2_CUDA(cudaMalloc(&staging, max_raw));
    B2_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    B2_CUDA(cudaEventCreate(&ev0));
    B2_CUDA(cudaEventCreate(&ev1));

    // ---- upload + repack -------------------------------------------------------------------------
    B2_TRY(qmat_alloc(arena, w_embd, embd.t->type, V, E));
    B2_TRY(upload_quant(embd.t, V, E, w_embd, 0, 1, 0));
    if (tied) w_head = w_embd;
    else {
        B2_TRY(qmat_alloc(arena, w_head, head.t->type, V, E));
        B2_TRY(upload_quant(head.t, V, E, w_head, 0, 1, 0));
    }
    B2_TRY(upload_f32(f, "output_norm.weight", E, &out_norm, false));
    B2_TRY(upload_f32(f, "output_norm.bias", E, &out_norm_b, true));
    B2_TRY(upload_f32(f, "output.bias", V, &head_b, true));
    step_weight_bytes = gemv_weight_bytes(w_head) + (size_t)E * 4;
    layers.resize(hp.n_layer);
    for (int l = 0; l < hp.n_layer; l++) {
        const std::string p = "blk." + std::to_string(l) + ".";
        LayerPlan& P = lp[l];
        LayerW& L = layers[l];
        B2_TRY(upload_f32(f, p + "attn_norm.weight", E, &L.attn_norm, false));
        B2_TRY(upload_f32(f, p + "attn_norm.bias", E, &L.attn_norm_b, true));
        if (hp.arch == 0) {
            // q | k | v fused as far as the tensor types allow (Q4_K_M mixes Q4_K q,k with Q6_K v)
            const TensorPlan* parts[3] = {&P.q, &P.k, &P.v};
            const int64_t prow[3] = {nq, nkv, nkv};
            int64_t row0 = 0;
            for (int i = 0; i < 3;) {
                int j = i;
                int64_t rows = 0;
                while (j < 3 && parts[j]->t->type == parts[i]->t->type) rows += prow[j++];
                QkvSeg seg;
                seg.row0 = row0;
                B2_TRY(qmat_alloc(arena, seg.W, parts[i]->t->type, rows, E));
                int64_t r = 0;
                for (int q = i; q < j; q++) {
                    B2_TRY(upload_rows_headwise(parts[q]->t, prow[q] / hd, E, seg.W, r, (int64_t)tp_rank * prow[q]));
                    r += prow[q];
                }
                L.qkv.push_back(seg);
                row0 += rows;
                i = j;
            }
            B2_TRY(upload_f32(f, p + "ffn_norm.weight", E, &L.ffn_norm, false));
            B2_TRY(qmat_alloc(arena, L.w_up, P.gate.t->type, 2 * (int64_t)F, E));      // rows interleaved (gate_i, up_i)
            B2_TRY(upload_quant(P.gate.t, F, E, L.w_up, 0, 2, (int64_t)tp_rank * F));
            B2_TRY(upload_quant(P.up.t, F, E, L.w_up, 1, 2, (int64_t)tp_rank * F));
        } else {
            // attention and MLP of a phi-2 block read the same normed input: when the tensor types agree,
            // ffn_up's rows follow q|k|v in ONE matrix (one stage, one prologue, one barrier less per layer)
            const int64_t nqkv = nq + 2 * nkv;
            L.fused_up = P.qkv.t->type == P.up.t->type && !getenv("B2_NO_FUSED_UP");
            QkvSeg seg;
            seg.row0 = 0;
            B2_TRY(qmat_alloc(arena, seg.W, P.qkv.t->type, nqkv + (L.fused_up ? F : 0), E));
            B2_TRY(upload_rows_headwise(P.qkv.t, nqkv / hd, E, seg.W, 0, 0));
            if (L.fused_up) B2_TRY(upload_quant(P.up.t, F, E, seg.W, nqkv, 1, 0));
            L.qkv.push_back(seg);
            // bias rows follow the same permutation
            const GgufTensor* tb = f.find(p + "attn_qkv.bias");
            const GgufTensor* tub = f.find(p + "ffn_up.bias");
            if (tub && (tub->type != T_F32 || tub->ne.empty() || tub->ne[0] != F)) B2_FAIL(B2_ERR_FORMAT, "gguf: %sffn_up.bias mis-shaped", p.c_str());
            if (tb || (L.fused_up && tub)) {
                if (tb && (tb->type != T_F32 || tb->ne.empty() || tb->ne[0] != nqkv)) B2_FAIL(B2_ERR_FORMAT, "gguf: %sattn_qkv.bias mis-shaped", p.c_str());
                std::vector<float> b((size_t)nqkv + (L.fused_up ? F : 0), 0.f);
                if (tb) {
                    const float* src = (const float*)tb->data;
                    for (int64_t h = 0; h < nqkv / hd; h++)
                        for (int e = 0; e < hd; e++) {
                            const int dsti = e < hp.n_rot / 2 ? 2 * e : e < hp.n_rot ? 2 * (e - hp.n_rot / 2) + 1 : e;
                            b[h * hd + dsti] = src[h * hd + e];
                        }
                }
                if (L.fused_up && tub) memcpy(b.data() + nqkv, tub->data, (size_t)F * 4);
                L.b_qkv = (float*)arena.take(b.size() * 4);
                if (!L.b_qkv) B2_FAIL(B2_ERR_NOMEM, "weight arena exhausted");
                B2_CUDA(cudaMemcpy(L.b_qkv, b.data(), b.size() * 4, cudaMemcpyHostToDevice));
            }
            if (!L.fused_up) {
                B2_TRY(qmat_alloc(arena, L.w_up, P.up.t->type, F, E));
                B2_TRY(upload_quant(P.up.t, F, E, L.w_up, 0, 1, 0));
                B2_TRY(upload_f32(f, p + "ffn_up.bias", F, &L.b_up, true));
            }
            B2_TRY(upload_f32(f, p + "ffn_down.bias", E, &L.b_down, true));
            B2_TRY(upload_f32(f, p + "attn_output.bias", E, &L.b_o, true));
        }
        B2_TRY(qmat_alloc(arena, L.w_o, P.o.t->type, E, nq));
        B2_TRY(qmat_alloc(arena, L.w_down, P.down.t->type, E, F));
        if (tp_size == 1) {
            B2_TRY(upload_quant(P.o.t, E, nq, L.w_o, 0, 1, 0));
            B2_TRY(upload_quant(P.down.t, E, F, L.w_down, 0, 1, 0));
        } else {          // K slices: this shard's heads of the attention output, its rows of the FFN hidden vector
            B2_TRY(upload_quant_kslice(P.o.t, E, nq_full, (int64_t)tp_rank * nq, nq, L.w_o));
            B2_TRY(upload_quant_kslice(P.down.t, E, F_full, (int64_t)tp_rank * F, F, L.w_down));
        }
        for (auto& s : L.qkv) step_weight_bytes += gemv_weight_bytes(s.W);
        step_weight_bytes += gemv_weight_bytes(L.w_o) + (L.fused_up ? 0 : gemv_weight_bytes(L.w_up)) + gemv_weight_bytes(L.w_down) + (size_t)2 * E * 4;
    }
    weight_bytes = arena.used;
    cudaFree(staging);
    staging = nullptr;

    // ---- KV cache pages + slots --------------------------------------------------------------------
    pages_per_seq = (max_ctx + PAGE_TOKENS - 1) / PAGE_TOKENS;
    n_pages = pages_per_seq * max_batch;
    layer_kv_elems = (size_t)n_pages * PAGE_TOKENS * nkv;
    B2_CUDA(cudaMalloc(&kv_k, layer_kv_elems * hp.n_layer * sizeof(__half)));
    B2_CUDA(cudaMalloc(&kv_v, layer_kv_elems * hp.n_layer * sizeof(__half)));
    B2_CUDA(cudaMemset(kv_k, 0, layer_kv_elems * hp.n_layer * sizeof(__half)));
    B2_CUDA(cudaMemset(kv_v, 0, layer_kv_elems * hp.n_layer * sizeof(__half)));
    for (int i = n_pages - 1; i >= 0; i--) free_pages.push_back(i);

    // per-slot device block
    out_cap = max_ctx;
    size_t off = 0;
    auto carve = [&](size_t bytes) { const size_t o2 = align_up(off, 256); off = o2 + bytes; return o2; };
    const size_t o_state = carve(sizeof(SeqState)), o_pt = carve((size_t)pages_per_seq * 4), o_x = carve((size_t)E * 4), o_q = carve((size_t)nq * 4),
                 o_att = carve((size_t)nq * 4), o_att2 = carve((size_t)E * 4), o_h = carve((size_t)F * 4), o_logits = carve((size_t)V * 4),
                 o_pos = carve(4), o_key = carve(8), o_gc = carve(4), o_out = carve((size_t)out_cap * 4),
                 o_forced = carve((size_t)max_ctx * 4), o_xll = carve((size_t)E * 12), o_attll = carve((size_t)nq * 8), o_hll = carve((size_t)F * 8), o_qkvll = carve((size_t)(nq + nkv) * 8);
    const size_t slot_bytes = align_up(off, 256);
    B2_CUDA(cudaMalloc(&slot_mem, slot_bytes * max_batch));
    B2_CUDA(cudaMemset(slot_mem, 0, slot_bytes * max_batch));
    slots.resize(max_batch);
    B2_CUDA(cudaMallocHost(&pin_state, (size_t)2 * max_batch * sizeof(SeqState)));
    for (int i = 0; i < max_batch; i++) {
        uint8_t* b = (uint8_t*)slot_mem + slot_bytes * i;
        Slot& s = slots[i];
        s.d_state = (SeqState*)(b + o_state); s.d_page_table = (int*)(b + o_pt); s.x = (float*)(b + o_x); s.q = (float*)(b + o_q);
        s.att = (float*)(b + o_att); s.att2 = (float*)(b + o_att2); s.h = (float*)(b + o_h); s.logits = (float*)(b + o_logits);
        s.d_pos = (int*)(b + o_pos); s.d_key = (unsigned long long*)(b + o_key);
        s.d_out = (int*)(b + o_out); s.d_forced = (int*)(b + o_forced); s.d_grid_count = (unsigned*)(b + o_gc);
        s.x_ll[0] = (float*)(b + o_xll); s.x_ll[1] = s.x_ll[0] + E; s.x_ll[2] = s.x_ll[1] + E;
        s.att_ll[0] = (float*)(b + o_attll); s.att_ll[1] = s.att_ll[0] + nq;
        s.h_ll[0] = (float*)(b + o_hll); s.h_ll[1] = s.h_ll[0] + F;
        s.qkv_ll[0] = (float*)(b + o_qkvll); s.qkv_ll[1] = s.qkv_ll[0] + nq + nkv;      // [nq floats | nkv/2 words K | nkv/2 words V]
    }
    // RoPE (cos, sin) for every position, computed on the host exactly as ggml does (ggml_rope_cache_init with
    // ext_factor 0: theta_base starts at pos and is multiplied by freq_base^(-2/n_rot) once per pair; the angle is
    // freq_scale * (theta_base / freq_factor[pair]); libm cosf/sinf).  `rope_freqs.weight` (Llama-3.1/3.2) holds the
    // per-pair factors, `<arch>.rope.scaling.type = linear` divides positions by the scaling factor; yarn is refused.
    {
        const std::string arch = hp.arch == 0 ? "llama" : "phi2";
        std::string stype;
        float freq_scale = 1.0f;
        if (f.get_str(arch + ".rope.scaling.type", &stype) && stype != "none") {
            if (stype != "linear") B2_FAIL(B2_ERR_UNSUPPORTED, "rope scaling type '%s' is not supported (none, linear)", stype.c_str());
            double fac = 1.0;
            if (f.get_f64(arch + ".rope.scaling.factor", &fac) && fac > 0.0) freq_scale = 1.0f / (float)fac;
        }
        const int half = hp.n_rot / 2;
        std::vector<float> factors;
        if (const GgufTensor* t = f.find("rope_freqs.weight")) {
            if (t->type != T_F32 || t->ne.size() != 1 || t->ne[0] != half) B2_FAIL(B2_ERR_FORMAT, "rope_freqs.weight must be f32[%d]", half);
            factors.assign((const float*)t->data, (const float*)t->data + half);
        }
        const float theta_scale = powf(hp.rope_freq_base, -2.0f / hp.n_rot);
        std::vector<float2> tab((size_t)max_ctx * half);
        for (int pos = 0; pos < max_ctx; pos++) {
            float theta_base = (float)pos;
            for (int p = 0; p < half; p++) {
                const float ff = factors.empty() ? 1.0f : factors[p];
                const float theta = freq_scale * (theta_base / ff);
                tab[(size_t)pos * half + p] = make_float2(cosf(theta), sinf(theta));
                theta_base *= theta_scale;
            }
        }
        B2_CUDA(cudaMalloc(&rope_table, tab.size() * sizeof(float2)));
        B2_CUDA(cudaMemcpy(rope_table, tab.data(), tab.size() * sizeof(float2), cudaMemcpyHostToDevice));
    }
    if (tp_size > 1) return tp_alloc();       // the pipelines are built by tp_connect once every shard is loaded
    return finish_load();
}

int Model::finish_load() {
    for (Slot& s : slots) {
        build_ops(s);
        B2_TRY(setup_mega(s));
    }
    if (use_batch) B2_TRY(setup_batch());
    return B2_OK;
}

// Exchange buffers of this shard, in one allocation: [2][E] partial sums of a single-token step,
// flags, epoch and ticket, then [2][BATCH_MAX][E] partial sums of a batched step.
int Model::tp_alloc() {
    const size_t part_bytes = (size_t)hp.n_embd * sizeof(double), total = 2 * part_bytes + 512 + 2 * BATCH_MAX * part_bytes;
    B2_CUDA(cudaMalloc(&tp_mem, total));
    B2_CUDA(cudaMemset(tp_mem, 0, total));
    return B2_OK;
}

int Model::tp_connect(Model* const* shards) {
    B2_CUDA(cudaSetDevice(device));
    const size_t part_bytes = (size_t)hp.n_embd * sizeof(double);
    tp = TpPeers();
    tp.n = tp_size; tp.rank = tp_rank;
    for (int d = 0; d < tp_size; d++) {
        if (d != tp_rank) {
            int can = 0;
            B2_CUDA(cudaDeviceCanAccessPeer(&can, device, shards[d]->device));
            if (!can) B2_FAIL(B2_ERR_UNSUPPORTED, "tensor parallelism needs peer access between GPU %d and GPU %d", device, shards[d]->device);
            const cudaError_t e = cudaDeviceEnablePeerAccess(shards[d]->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) B2_CUDA(e);
            cudaGetLastError();
        }
        uint8_t* b = (uint8_t*)shards[d]->tp_mem;
        tp.part[0][d] = (double*)b;
        tp.part[1][d] = (double*)(b + part_bytes);
        tp.flags[d] = (unsigned*)(b + 2 * part_bytes);
    }
    tp.epoch = (unsigned*)((uint8_t*)tp_mem + 2 * part_bytes + 256);
    tp.ticket = tp.epoch + 1;
    // batched steps: same flags and epoch sequence (every shard runs the same schedule), their own partial-sum buffers
    tpb = tp;
    for (int d = 0; d < tp_size; d++) {
        uint8_t* b = (uint8_t*)shards[d]->tp_mem + 2 * part_bytes + 512;
        tpb.part[0][d] = (double*)b;
        tpb.part[1][d] = (double*)(b + BATCH_MAX * part_bytes);
    }
    return finish_load();
}

// ---- sequences -------------------------------------------------------------------------------------
int Model::seq_new(int* out) {
    for (int i = 0; i < max_batch; i++)
        if (!slots[i].used) {
            slots[i].used = true;
            slots[i].len = 0;
            slots[i].host = SeqState{};
            *out = i;
            return B2_OK;
        }
    B2_FAIL(B2_ERR_NOMEM, "all %d sequence slots are in use", max_batch);
}
int Model::check_seq(int seq) const {
    if (seq < 0 || seq >= max_batch || !slots[seq].used) B2_FAIL(B2_ERR_STATE, "sequence %d is not open", seq);
    return B2_OK;
}
int Model::seq_free(int seq) {
    B2_TRY(check_seq(seq));
    Slot& s = slots[seq];
    for (int p : s.pages) free_pages.push_back(p);
    s.pages.clear();
    s.used = false;
    s.len = 0;
    return B2_OK;
}
int Model::seq_truncate(int seq, int n_keep) {
    B2_TRY(check_seq(seq));
    Slot& s = slots[seq];
    if (n_keep < 0 || n_keep > s.len) B2_FAIL(B2_ERR_ARG, "truncate: n_keep=%d outside [0,%d]", n_keep, s.len);
    s.len = n_keep;
    const size_t keep_pages = (size_t)(n_keep + PAGE_TOKENS - 1) / PAGE_TOKENS;
    while (s.pages.size() > keep_pages) { free_pages.push_back(s.pages.back()); s.pages.pop_back(); }
    return B2_OK;
}
// Tooling: declare cache rows [0, n) valid without evaluating a prompt (K/V are whatever the pages hold,
// zeros after load) -- lets a benchmark reach a long context without paying for the prefill.
int Model::seq_assume(int seq, int n) {
    B2_TRY(check_seq(seq));
    Slot& s = slots[seq];
    if (n < s.len) B2_FAIL(B2_ERR_ARG, "seq_assume: %d is shorter than the sequence (%d)", n, s.len);
    B2_TRY(ensure_pages(s, n));
    B2_CUDA(cudaStreamSynchronize(stream));
    s.len = n;
    return B2_OK;
}
// make sure pages exist for positions [0, upto) and the device page table matches
int Model::ensure_pages(Slot& s, int upto) {
    if (upto > max_ctx) B2_FAIL(B2_ERR_NOMEM, "sequence would grow to %d tokens, max_ctx is %d", upto, max_ctx);
    const size_t want = (size_t)(upto + PAGE_TOKENS - 1) / PAGE_TOKENS;
    const size_t had = s.pages.size();
    while (s.pages.size() < want) {
        if (free_pages.empty()) B2_FAIL(B2_ERR_NOMEM, "KV cache pages exhausted");
        s.pages.push_back(free_pages.back());
        free_pages.pop_back();
    }
    if (s.pages.size() != had || !s.table_synced) {
        B2_CUDA(cudaMemcpyAsync(s.d_page_table, s.pages.data(), s.pages.size() * 4, cudaMemcpyHostToDevice, stream));
        s.table_synced = true;
    }
    return B2_OK;
}

// ---- the layer pipeline of one sequence slot: 5 ops per layer (qkv may be split by tensor type) -----
void Model::build_ops(Slot& s) {
    const int hd = hp.head_dim, nq = n_head_l * hd, nkv = n_head_kv_l * hd;
    s.ops.clear();
    for (int l = 0; l < hp.n_layer; l++) {
        const LayerW& L = layers[l];
        __half* kb = kv_k + layer_kv_elems * l;
        __half* vb = kv_v + layer_kv_elems * l;
        for (const QkvSeg& seg : L.qkv) {
            StepOp op; op.kind = 0; op.cls = 1;
            GemvArgs& a = op.g;
            a.W = seg.W; a.pro = hp.arch == 0 ? PRO_RMSNORM : PRO_LAYERNORM; a.epi = EPI_QKV_ROPE;
            a.x = s.x; a.gain = L.attn_norm; a.nbias = L.attn_norm_b; a.eps = hp.norm_eps;
            a.bias = L.b_qkv; a.y = s.q; a.epi_row0 = seg.row0;
            if (L.fused_up) { a.epi = EPI_QKV_ROPE_GELU; a.y3 = s.h; }
            a.n_q = nq; a.n_kv = nkv; a.head_dim = hd; a.n_rot = hp.n_rot; a.rope_neox = neox ? 1 : 0;
            a.q_scale = hp.arch == 1 ? 1.0f / sqrtf((float)hd) : 1.0f;
            a.rope_cs = rope_table;
            a.kv.k_base = kb; a.kv.v_base = vb; a.kv.page_table = s.d_page_table; a.kv.pos = s.d_pos;
            a.kv.page_tokens = PAGE_TOKENS; a.kv.kv_dim = nkv;
            s.ops.push_back(op);
        }
        {
            StepOp op; op.kind = 1; op.cls = 2;
            AttnArgs& a = op.a;
            a.q = s.q; a.out = s.att; a.k_base = kb; a.v_base = vb; a.page_table = s.d_page_table; a.pos = s.d_pos;
            a.n_head = n_head_l; a.n_head_kv = n_head_kv_l; a.head_dim = hd; a.page_tokens = PAGE_TOKENS; a.kv_dim = nkv;
            a.max_ctx = max_ctx;
            a.scale = hp.arch == 1 ? 1.0f : 1.0f / sqrtf((float)hd);
            s.ops.push_back(op);
        }
        StepOp o, u, d;
        o.kind = u.kind = d.kind = 0; o.cls = 3; u.cls = 4; d.cls = 5;
        if (hp.arch == 0) {
            o.g.W = L.w_o; o.g.pro = PRO_PLAIN; o.g.epi = EPI_RESIDUAL; o.g.x = s.att; o.g.y = s.x;
            u.g.W = L.w_up; u.g.pro = PRO_RMSNORM; u.g.epi = EPI_SILU_MUL; u.g.x = s.x; u.g.gain = L.ffn_norm; u.g.eps = hp.norm_eps; u.g.y = s.h;
            d.g.W = L.w_down; d.g.pro = PRO_PLAIN; d.g.epi = EPI_RESIDUAL; d.g.x = s.h; d.g.y = s.x;
            if (tp_size > 1) {      // K-split projections: unrounded partial sums, then the all-reduce adds the residual
                o.g.epi = EPI_PARTIAL; o.g.y64 = tp.part[0][tp_rank]; o.tp_buf = 0;
                d.g.epi = EPI_PARTIAL; d.g.y64 = tp.part[1][tp_rank]; d.tp_buf = 1;
            }
        } else {
            o.g.W = L.w_o; o.g.pro = PRO_PLAIN; o.g.epi = EPI_STORE; o.g.x = s.att; o.g.bias = L.b_o; o.g.y = s.att2;
            u.g.W = L.w_up; u.g.pro = PRO_LAYERNORM; u.g.epi = EPI_GELU; u.g.x = s.x; u.g.gain = L.attn_norm; u.g.nbias = L.attn_norm_b;
            u.g.eps = hp.norm_eps; u.g.bias = L.b_up; u.g.y = s.h;
            d.g.W = L.w_down; d.g.pro = PRO_PLAIN; d.g.epi = EPI_RESIDUAL2; d.g.x = s.h; d.g.bias = L.b_down; d.g.y = s.x; d.g.y2 = s.att2;
        }
        s.ops.push_back(o);
        if (!L.fused_up) s.ops.push_back(u);
        s.ops.push_back(d);
    }
}

// Persistent-kernel path: possible when every layer matrix has the same supported type.
int Model::setup_mega(Slot& s) {
    s.mega_ok = false;
    if (!use_mega) return B2_OK;
    int type = -1;
    for (const StepOp& op : s.ops)
        if (op.kind == 0) {
            const int t = op.g.W.type == T_Q4_K || op.g.W.type == T_Q6_K ? T_KQ : op.g.W.type;      // Q4_K_M mixes the two K-quants
            if (type < 0) type = t;
            if (t != type) {          // other mixtures: per-matrix kernels
                if (getenv("B2_MEGA_VERBOSE")) fprintf(stderr, "b2: persistent kernel not used: tensor types %d and %d in one model\n", type, t);
                return B2_OK;
            }
        }
    const int grp = hp.n_head / hp.n_head_kv;
    if (!mega_supported(type, grp, hp.head_dim) || (tp_size > 1 && hp.head_dim == 80)) {
        if (getenv("B2_MEGA_VERBOSE")) fprintf(stderr, "b2: persistent kernel not used: no variant for type %d, %d query heads per KV head, head_dim %d\n", type, grp, hp.head_dim);
        return B2_OK;
    }
    std::vector<MegaStage> st;
    int64_t k_max = 0;
    size_t tile_max = 0;
    int rt_s_max = 0;
    MegaLaunch& L = s.mega;
    L = MegaLaunch();
    L.type = type; L.grp = grp; L.head_dim = hp.head_dim;
    L.cluster = hp.head_dim % 32 == 0 ? 4 : 2;      // the output dims of a head are split eight per warp over the cluster's CTAs (phi-2: 80 = 2 x 40)
    L.tp = tp;
    if (const char* v = getenv("B2_MEGA_CLUSTER")) L.cluster = atoi(v);
    {   // kernel variant: 4 or 5 activation units per lane (K = 5120 needs 5 to keep one warp per row)
        double c4 = 0, c5 = 0;
        for (const StepOp& op : s.ops)
            if (op.kind == 0) { c4 += mega_plan_cost(op.g.W, 4); c5 += mega_plan_cost(op.g.W, 5); }
        L.nb = c5 < 0.9 * c4 ? 5 : 4;
        if (const char* v = getenv("B2_MEGA_NB")) L.nb = atoi(v) == 5 ? 5 : 4;
    }
    for (size_t i = 0; i < s.ops.size(); i++) {
        MegaStage m;
        m.kind = s.ops[i].kind;
        if (s.ops[i].kind == 0) {
            m.g = s.ops[i].g;
            if (mega_plan_stage(m, MEGA_NCW, L.nb) != B2_OK) return B2_OK;
            const QMat& W = m.g.W;
            size_t rb = 0;
            for (int p = 0; p < W.n_planes; p++) rb += W.stride[p];
            k_max = std::max(k_max, W.k);
            tile_max = std::max(tile_max, rb * m.RT);
            // per-CTA row-sum buffer: (rows of the CTA + slack) x S doubles; sized for the smallest grid we might get
            rt_s_max = std::max(rt_s_max, (int)((W.rows / 64 + 4) * m.S));
        } else {
            m.at = s.ops[i].a;
        }
        st.push_back(m);
        if (s.ops[i].tp_buf >= 0) {          // K-split projection: the exchange with the other shards follows as a stage of its own
            MegaStage x;
            x.kind = 2; x.tp_buf = s.ops[i].tp_buf; x.tp_n = hp.n_embd; x.tp_x = s.x;
            st.push_back(x);
        }
    }
    // Stage hand-overs through tagged values instead of a grid barrier (mega.cu, `ll_in`): where a GEMV stage of the fast
    // prologue path reads exactly the vector the stage before it wrote (attention -> wo, wo -> up, down -> next qkv).
    if (tp_size <= 1 && !getenv("B2_MEGA_NO_LL")) {
        const int nkv_l = n_head_kv_l * hp.head_dim;
        int wx = 0, wa = 0, wh = 0, wq = 0;  // writes of x / att / h / qkv so far in the step: write w uses buffer w % NB and re-poisons that of write w - 1
        for (size_t i = 1; i < st.size(); i++) {
            MegaStage& c = st[i];
            MegaStage& p = st[i - 1];
            if (c.kind == 1 && p.kind == 0 && p.g.epi == EPI_QKV_ROPE && !p.g.rope_neox && p.g.epi_row0 == 0 && !p.g.bias && p.g.y == c.at.q &&
                p.g.W.rows == (int64_t)p.g.n_q + 2 * p.g.n_kv && p.g.n_q == hp.n_head * hp.head_dim && p.g.n_kv == nkv_l && hp.head_dim % 2 == 0 &&
                !getenv("B2_MEGA_NO_LL_QKV")) {
                // qkv -> attention: the query and the current position's K/V rows (written in adjacent pairs by "norm" RoPE)
                p.g.y_ll = s.qkv_ll[wq & 1]; p.g.y_ll_poison = s.qkv_ll[(wq + 1) & 1]; c.at.qkv_ll = s.qkv_ll[wq & 1]; c.ll_in = 1; wq++;
                continue;
            }
            if (c.kind != 0) continue;
            const GemvArgs& g = c.g;
            const bool fast = (g.W.type == T_Q4_0 || g.W.type == T_Q8_0) && (g.pro == PRO_PLAIN || g.pro == PRO_RMSNORM) && g.W.k % 256 == 0 &&
                              g.W.k / 4 <= (int64_t)8 * MEGA_NCW * 32;          // the kernel's fast prologue path (mega.cu)
            // (a stage outside the fast path -- K-quants, very long vectors -- polls the buffer complete and then runs its generic prologue on it)
            if (!fast && !((g.pro == PRO_PLAIN || g.pro == PRO_RMSNORM) && g.W.k % 4 == 0 && !getenv("B2_MEGA_NO_LL_GENERIC"))) continue;
            if (p.kind == 1 && p.at.out == g.x && g.W.k == (int64_t)hp.n_head * hp.head_dim) {
                p.at.out_ll = s.att_ll[wa & 1]; p.at.out_ll_poison = s.att_ll[(wa + 1) & 1]; c.g.x_ll = s.att_ll[wa & 1]; c.ll_in = 1; wa++;
            } else if (p.kind == 0 && p.g.epi == EPI_RESIDUAL && p.g.y == g.x && p.g.epi_row0 == 0 && p.g.W.rows == g.W.k && g.W.k == hp.n_embd) {
                p.g.y_ll = s.x_ll[wx % 3]; p.g.y_ll_poison = s.x_ll[(wx + 2) % 3]; c.g.x_ll = s.x_ll[wx % 3]; c.ll_in = 1; wx++;
            } else if (p.kind == 0 && p.g.epi == EPI_SILU_MUL && p.g.y == g.x && p.g.epi_row0 == 0 && p.g.W.rows == 2 * g.W.k && g.W.k == hp.n_ff &&
                       getenv("B2_MEGA_LL_H")) {      // (measured: 650 tok/s with h handed over this way, 659 with the grid barrier kept there)
                p.g.y_ll = s.h_ll[wh & 1]; p.g.y_ll_poison = s.h_ll[(wh + 1) & 1]; c.g.x_ll = s.h_ll[wh & 1]; c.ll_in = 1; wh++;
            }
        }
        if (wx) { L.ll_init.p[0] = s.x_ll[0]; L.ll_init.p[1] = s.x_ll[1]; L.ll_init.n[0] = L.ll_init.n[1] = hp.n_embd; }
        if (wa) { L.ll_init.p[2] = s.att_ll[0]; L.ll_init.n[2] = hp.n_head * hp.head_dim; }
        if (wh) { L.ll_init.p[3] = s.h_ll[0]; L.ll_init.n[3] = hp.n_ff; }
        if (wq) { L.ll_init.p[4] = s.qkv_ll[0]; L.ll_init.n[4] = hp.n_head * hp.head_dim + nkv_l; }
    }
    for (size_t i = 0; i + 1 < st.size(); i++)
        if (st[i].kind == 0 && st[i + 1].kind == 1) { st[i].next_attn = 1; st[i].at = st[i + 1].at; }
    // per-thread epilogue state kept across layers: only where every such stage has the same rows (one split) and nobody else writes x
    if (tp_size <= 1 && !getenv("B2_MEGA_NO_EPI_CACHE"))
        for (MegaStage& m : st) {
            if (m.kind != 0 || m.g.epi_row0 != 0 || m.g.bias) continue;
            if (m.g.epi == EPI_RESIDUAL && m.g.y == s.x && m.g.W.rows == hp.n_embd) m.epi_cache = 1;
            else if (m.g.epi == EPI_QKV_ROPE && m.g.W.rows == (int64_t)m.g.n_q + 2 * m.g.n_kv) m.epi_cache = 2;
        }
    int rc_cfg = mega_configure(L, k_max, tile_max, rt_s_max, max_ctx);
    if (L.ns < 3 && L.ring_room / 3 >= (44u << 10) && !getenv("B2_MEGA_TILE_KB")) {
        // A long context leaves the ring less room (score and row-offset buffers of the attention).  Three slots of smaller
        // tiles stream better than two of the default size (7B: 712 against 680 tok/s), down to tiles of ~44 KB.
        const int target = (int)(L.ring_room / 3 / 128 * 128);
        tile_max = 0;
        for (MegaStage& m : st)
            if (m.kind == 0) {
                if (mega_plan_stage(m, MEGA_NCW, L.nb, target) != B2_OK) return B2_OK;
                size_t rb = 0;
                for (int p = 0; p < m.g.W.n_planes; p++) rb += m.g.W.stride[p];
                tile_max = std::max(tile_max, rb * m.RT);
            }
        rc_cfg = mega_configure(L, k_max, tile_max, rt_s_max, max_ctx);
    }
    if (rc_cfg != B2_OK) {
        if (getenv("B2_MEGA_VERBOSE")) fprintf(stderr, "b2: persistent kernel not used: %s\n", last_error());
        cudaGetLastError();          // does not fit: keep the per-matrix path, and leave no stale CUDA error behind
        return B2_OK;
    }
    for (auto& x : st) {
        x.att_max_local = L.att_max_local;
        mega_finish_stage(x, L.grid);
        // (x * inv) >> 16 is exact only for small x: the per-CTA pair count must stay far below 2^16 / PT
        if (x.kind == 0 && x.pairs_base + 1 + x.PT >= 4096) return B2_OK;
    }
    B2_CUDA(cudaMalloc(&s.d_stages, st.size() * sizeof(MegaStage)));
    B2_CUDA(cudaMemcpy(s.d_stages, st.data(), st.size() * sizeof(MegaStage), cudaMemcpyHostToDevice));
    L.stages = s.d_stages; L.s_begin = 0; L.s_end = (int)st.size(); L.grid_count = s.d_grid_count;
    L.pos = s.d_pos; L.page_table = s.d_page_table; L.n_pages = pages_per_seq;
    L.pf_bytes = 256u << 10;                        // x 132 CTAs = 34 MB of the 126 MB L2
    if (const char* v = getenv("B2_MEGA_PF_KB")) L.pf_bytes = (uint32_t)atoi(v) << 10;
    if (const char* v = getenv("B2_MEGA_DEBUG")) L.debug = atoi(v);
    if (getenv("B2_MEGA_VERBOSE"))
        fprintf(stderr, "b2: persistent kernel: %d CTAs (clusters of %d), %d ring slots of %u B after %u B of buffers, %zu B of shared memory\n", L.grid, L.cluster, L.ns,
                L.stage_bytes, L.off_ring, L.smem_bytes);
    if (pages_per_seq > 256) return B2_OK;          // the kernel keeps the page table in shared memory
    s.mega_ok = true;
    return B2_OK;
}

// ---- one decode step, as a chain of kernel launches on `stream` -----------------------------------
int Model::enqueue_step(Slot& s, bool with_head, int* n_launch, StepProfile* prof) {
    auto mark = [&](int cls) -> int {
        if (!prof) return B2_OK;
        cudaEvent_t e;
        B2_CUDA(cudaEventCreate(&e));
        B2_CUDA(cudaEventRecord(e, stream));
        prof->ev.push_back(e);
        prof->cls.push_back(cls);
        return B2_OK;
    };
    B2_TRY(mark(-1));
    int nl = 0;
    const bool pdl = use_pdl;
    {
        StepBeginArgs a;
        a.st = s.d_state; a.embd = w_embd; a.x = s.x; a.pos_out = s.d_pos; a.argmax_key = s.d_key; a.grid_count = s.d_grid_count;
        if (s.mega_ok) a.ll_init = s.mega.ll_init;
        a.st_host = s.zc_in;
        B2_TRY(step_begin_launch(a, stream, pdl)); nl++; B2_TRY(mark(0));
    }
    if (s.mega_ok) {
        B2_TRY(mega_launch(s.mega, stream)); nl++; B2_TRY(mark(8));
    } else {
        for (const StepOp& op : s.ops) {
            if (op.kind == 0) B2_TRY(gemv_launch(op.g, stream, pdl));
            else B2_TRY(attn_decode_launch(op.a, stream, pdl));
            nl++;
            if (op.tp_buf >= 0 && !tp_skip_reduce) { B2_TRY(tp_allreduce_residual_launch(tp, op.tp_buf, s.x, hp.n_embd, stream, pdl)); nl++; }
            B2_TRY(mark(op.cls));
        }
    }
    if (with_head) {
        GemvArgs h;
        h.W = w_head; h.pro = hp.arch == 0 ? PRO_RMSNORM : PRO_LAYERNORM; h.epi = EPI_LOGITS;
        h.x = s.x; h.gain = out_norm; h.nbias = out_norm_b; h.eps = hp.norm_eps; h.bias = head_b; h.y = s.logits; h.argmax_key = s.d_key;
        B2_TRY(gemv_launch(h, stream, pdl && !s.mega_ok)); nl++; B2_TRY(mark(6));
    }
    {
        StepEndArgs a;
        a.st = s.d_state; a.argmax_key = s.d_key; a.forced = s.d_forced; a.out_toks = s.d_out; a.out_cap = out_cap; a.use_argmax = with_head ? 1 : 0;
        a.st_host = s.zc_out;
        B2_TRY(step_end_launch(a, stream, pdl)); nl++; B2_TRY(mark(7));
    }
    *n_launch = nl;
    return B2_OK;
}

// Run `n` steps of one kind.  The first step of each kind runs eagerly (it also sets the
// kernels' shared-memory attributes); after that a captured graph of the step is replayed.
int Model::run_steps(Slot& s, bool with_head, int n) {
    cudaGraphExec_t& g = with_head ? s.g_full : s.g_nohead;
    int& gl = with_head ? s.gl_full : s.gl_nohead;
    bool& warmed = with_head ? s.warm_full : s.warm_nohead;
    for (int i = 0; i < n; i++) {
        if (!use_graph || !warmed) {
            int nl = 0;
            B2_TRY(enqueue_step(s, with_head, &nl, nullptr));
            warmed = true;
            gl = nl;
            launches += nl; last_launches += nl;
            continue;
        }
        if (!g) {
            cudaGraph_t graph = nullptr;
            B2_CUDA(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal));
            int nl = 0;
            const int rc = enqueue_step(s, with_head, &nl, nullptr);
            cudaError_t ce = cudaStreamEndCapture(stream, &graph);
            if (rc != B2_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
            if (ce == cudaSuccess) ce = cudaGraphInstantiate(&g, graph, 0);
            if (graph) cudaGraphDestroy(graph);
            if (ce != cudaSuccess) {
                g = nullptr;
                cudaGetLastError();
                if (use_pdl) { use_pdl = false; i--; continue; }     // retry the capture without PDL edges
                B2_FAIL(B2_ERR_CUDA, "graph capture failed: %s", cudaGetErrorString(ce));
            }
            gl = nl;
        }
        B2_CUDA(cudaGraphLaunch(g, stream));
        launches += gl; last_launches += gl;
    }
    steps += n; last_steps += n;
    return B2_OK;
}

int Model::begin_call() {
    last_steps = 0; last_launches = 0;
    B2_CUDA(cudaEventRecord(ev0, stream));
    return B2_OK;
}
int Model::end_call(Slot* s) {
    B2_CUDA(cudaEventRecord(ev1, stream));
    if (s) B2_CUDA(cudaMemcpyAsync(&s->host, s->d_state, sizeof(SeqState), cudaMemcpyDeviceToHost, stream));
    B2_CUDA(cudaStreamSynchronize(stream));
    float ms = 0;
    B2_CUDA(cudaEventElapsedTime(&ms, ev0, ev1));
    last_ms = ms;
    return B2_OK;
}

int Model::prefill(int seq, const int32_t* toks, int n, float* logits) {
    B2_TRY(check_seq(seq));
    if (!toks || n <= 0) B2_FAIL(B2_ERR_ARG, "prefill: need at least one token");
    for (int i = 0; i < n; i++)
        if (toks[i] < 0 || toks[i] >= hp.n_vocab) B2_FAIL(B2_ERR_ARG, "prefill: token %d at %d outside the vocabulary [0,%d)", toks[i], i, hp.n_vocab);
    Slot& s = slots[seq];
    B2_TRY(ensure_pages(s, s.len + n));
    B2_TRY(begin_call());
    if (bt.ok && n >= batch_min_seqs) {
        // chunks of up to bt.cap prompt tokens per pass over the weights (512 on the tensor-core path, else 32); only the
        // last token needs the head
        bt.next_region = 0;
        std::vector<int> col_slot(bt.cap), col_pos(bt.cap);
        int last_b = 0;
        for (int t0 = 0; t0 < n; t0 += bt.cap) {
            const int B = std::min(bt.cap, n - t0);
            for (int c = 0; c < B; c++) { col_slot[c] = seq; col_pos[c] = s.len + t0 + c; }
            B2_TRY(batch_step(B, col_slot.data(), col_pos.data(), toks + t0, t0 + B == n ? B - 1 : B));
            last_b = B;
        }
        if (logits) B2_CUDA(cudaMemcpyAsync(logits, bt.logits, (size_t)hp.n_vocab * 4, cudaMemcpyDeviceToHost, stream));
        B2_CUDA(cudaMemcpyAsync(bt.h_next, bt.d_next, (size_t)last_b * 4, cudaMemcpyDeviceToHost, stream));
        B2_TRY(end_call(nullptr));
        s.host = SeqState{};
        s.host.cur_tok = bt.h_next[last_b - 1];
        s.host.pos = s.len + n;
        s.len += n;
        return B2_OK;
    }
    // Short prompts (and models without batched kernels) walk the prompt with the decode kernels,
    // one token per step; prompt tokens are forced from the device-side list.
    SeqState st{};
    st.pos = s.len; st.cur_tok = toks[0]; st.n_forced = n - 1; st.forced_idx = 0; st.n_out = 0;
    B2_CUDA(cudaMemcpyAsync(s.d_state, &st, sizeof(st), cudaMemcpyHostToDevice, stream));
    if (n > 1) B2_CUDA(cudaMemcpyAsync(s.d_forced, toks + 1, (size_t)(n - 1) * 4, cudaMemcpyHostToDevice, stream));
    if (n > 1) B2_TRY(run_steps(s, false, n - 1));
    B2_TRY(run_steps(s, true, 1));
    if (logits) B2_CUDA(cudaMemcpyAsync(logits, s.logits, (size_t)hp.n_vocab * 4, cudaMemcpyDeviceToHost, stream));
    B2_TRY(end_call(&s));
    s.len += n;
    return B2_OK;
}

int Model::decode(int n_seq, const int* seqs, const int32_t* last_tok, int32_t* next_tok, float* logits) {
    if (n_seq <= 0 || !seqs || !last_tok) B2_FAIL(B2_ERR_ARG, "decode: bad arguments");
    for (int i = 0; i < n_seq; i++) {
        B2_TRY(check_seq(seqs[i]));
        if (last_tok[i] < 0 || last_tok[i] >= hp.n_vocab) B2_FAIL(B2_ERR_ARG, "decode: token %d outside the vocabulary", last_tok[i]);
        for (int j = 0; j < i; j++) if (seqs[j] == seqs[i]) B2_FAIL(B2_ERR_ARG, "decode: sequence %d listed twice", seqs[i]);
        B2_TRY(ensure_pages(slots[seqs[i]], slots[seqs[i]].len + 1));
    }
    B2_TRY(begin_call());
    if (bt.ok && n_seq >= batch_min_seqs) {
        bt.next_region = 0;
        int col_pos[BATCH_MAX];
        for (int c0 = 0; c0 < n_seq; c0 += BATCH_MAX) {
            const int B = std::min(BATCH_MAX, n_seq - c0);
            for (int c = 0; c < B; c++) col_pos[c] = slots[seqs[c0 + c]].len;
            B2_TRY(batch_step(B, seqs + c0, col_pos, last_tok + c0, 0));
            if (logits) B2_CUDA(cudaMemcpyAsync(logits + (size_t)c0 * hp.n_vocab, bt.logits, (size_t)B * hp.n_vocab * 4, cudaMemcpyDeviceToHost, stream));
            B2_CUDA(cudaMemcpyAsync(bt.h_next, bt.d_next, BATCH_MAX * 4, cudaMemcpyDeviceToHost, stream));
            {
                B2_CUDA(cudaStreamSynchronize(stream));        // h_next is reused by the next group
                for (int c = 0; c < B; c++) {
                    Slot& s = slots[seqs[c0 + c]];
                    s.host = SeqState{};
                    s.host.cur_tok = bt.h_next[c];
                    s.host.pos = s.len + 1;
                    if (next_tok) next_tok[c0 + c] = bt.h_next[c];
                }
            }
        }
        B2_TRY(end_call(nullptr));
        for (int i = 0; i < n_seq; i++) slots[seqs[i]].len += 1;
        return B2_OK;
    }
    // One sequence, no logits (what a server does per generated token): the copy of the token in, the step and the copy of
    // the state out are ONE graph, so the host enqueues a single launch -- every separately enqueued operation is GPU idle
    // time here, because the GPU waits for the host between tokens (tp: the shards run this path in lock step, graphs are
    // per shard, nothing changes).
    if (n_seq == 1 && !logits && use_graph && slots[seqs[0]].warm_full && !tp_skip_reduce) {
        Slot& s = slots[seqs[0]];
        SeqState& st = pin_state[2 * seqs[0]];
        st = SeqState{};
        st.pos = s.len; st.cur_tok = last_tok[0];
        if (!s.g_decode) {
            cudaGraph_t graph = nullptr;
            B2_CUDA(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal));
            int nl = 0;
            // (no copy nodes: step_begin reads the state from the pinned buffer, step_end writes the new one there -- two
            //  copy-engine hops less per token)
            s.zc_in = &st; s.zc_out = &pin_state[2 * seqs[0] + 1];
            const int rc = enqueue_step(s, true, &nl, nullptr);
            s.zc_in = s.zc_out = nullptr;
            cudaError_t ce = cudaStreamEndCapture(stream, &graph);
            if (rc == B2_OK && ce == cudaSuccess) ce = cudaGraphInstantiate(&s.g_decode, graph, 0);
            if (graph) cudaGraphDestroy(graph);
            if (rc != B2_OK || ce != cudaSuccess) { s.g_decode = nullptr; cudaGetLastError(); }
            s.gl_full = nl > 0 ? nl : s.gl_full;
        }
        if (s.g_decode) {
            B2_CUDA(cudaGraphLaunch(s.g_decode, stream));
            launches += s.gl_full; last_launches += s.gl_full; steps += 1; last_steps += 1;
            B2_TRY(end_call(nullptr));
            s.host = pin_state[2 * seqs[0] + 1];
            s.len += 1;
            if (next_tok) next_tok[0] = s.host.cur_tok;
            return B2_OK;
        }
    }
    // the sequence state crosses the bus through pinned staging (pageable copies would each synchronise):
    // everything of this call is queued, then ONE synchronisation
    for (int i = 0; i < n_seq; i++) {
        Slot& s = slots[seqs[i]];
        SeqState& st = pin_state[2 * seqs[i]];
        st = SeqState{};
        st.pos = s.len; st.cur_tok = last_tok[i];
        B2_CUDA(cudaMemcpyAsync(s.d_state, &st, sizeof(st), cudaMemcpyHostToDevice, stream));
        B2_TRY(run_steps(s, true, 1));
        if (logits) B2_CUDA(cudaMemcpyAsync(logits + (size_t)i * hp.n_vocab, s.logits, (size_t)hp.n_vocab * 4, cudaMemcpyDeviceToHost, stream));
        B2_CUDA(cudaMemcpyAsync(&pin_state[2 * seqs[i] + 1], s.d_state, sizeof(SeqState), cudaMemcpyDeviceToHost, stream));
    }
    B2_TRY(end_call(nullptr));
    for (int i = 0; i < n_seq; i++) {
        Slot& s = slots[seqs[i]];
        s.host = pin_state[2 * seqs[i] + 1];
        s.len += 1;
        if (next_tok) next_tok[i] = s.host.cur_tok;
    }
    return B2_OK;
}

int Model::decode_greedy(int seq, int32_t first_tok, int n_steps, int32_t* out_toks) {
    B2_TRY(check_seq(seq));
    if (n_steps <= 0 || n_steps > out_cap) B2_FAIL(B2_ERR_ARG, "decode_greedy: n_steps=%d outside [1,%d]", n_steps, out_cap);
    Slot& s = slots[seq];
    if (first_tok >= hp.n_vocab) B2_FAIL(B2_ERR_ARG, "decode_greedy: token %d outside the vocabulary", first_tok);
    if (first_tok < 0 && s.len == 0) B2_FAIL(B2_ERR_STATE, "decode_greedy: nothing to continue from");
    B2_TRY(ensure_pages(s, s.len + n_steps));
    B2_TRY(begin_call());
    SeqState st{};
    st.pos = s.len; st.cur_tok = first_tok >= 0 ? first_tok : s.host.cur_tok;
    B2_CUDA(cudaMemcpyAsync(s.d_state, &st, sizeof(st), cudaMemcpyHostToDevice, stream));
    B2_TRY(run_steps(s, true, n_steps));
    if (out_toks) B2_CUDA(cudaMemcpyAsync(out_toks, s.d_out, (size_t)n_steps * 4, cudaMemcpyDeviceToHost, stream));
    B2_TRY(end_call(&s));
    s.len += n_steps;
    return B2_OK;
}

// One eager decode step with a CUDA event after every launch: per-kernel-class device time
// inside a real step (no PDL overlap, so the classes add up to a slightly longer step).
int Model::profile_step(int seq, int32_t tok, float* class_ms, int* class_launches) {
    B2_TRY(check_seq(seq));
    if (tok < 0 || tok >= hp.n_vocab || !class_ms || !class_launches) B2_FAIL(B2_ERR_ARG, "profile_step: bad arguments");
    Slot& s = slots[seq];
    B2_TRY(ensure_pages(s, s.len + 1));
    SeqState st{};
    st.pos = s.len; st.cur_tok = tok;
    B2_CUDA(cudaMemcpyAsync(s.d_state, &st, sizeof(st), cudaMemcpyHostToDevice, stream));
    StepProfile prof;
    int nl = 0;
    const bool saved = use_pdl;
    use_pdl = false;
    B2_TRY(delay_launch(stream, 300000));      // the GPU idles 0.3 ms while the host enqueues: no launch gaps between the timed kernels
    const int rc = enqueue_step(s, true, &nl, &prof);
    use_pdl = saved;
    cudaStreamSynchronize(stream);
    for (int i = 0; i < 9; i++) { class_ms[i] = 0.f; class_launches[i] = 0; }
    if (rc == B2_OK)
        for (size_t i = 1; i < prof.ev.size(); i++) {
            float ms = 0.f;
            cudaEventElapsedTime(&ms, prof.ev[i - 1], prof.ev[i]);
            class_ms[prof.cls[i]] += ms;
            class_launches[prof.cls[i]]++;
        }
    for (cudaEvent_t e : prof.ev) cudaEventDestroy(e);
    B2_TRY(rc);
    B2_CUDA(cudaMemcpy(&s.host, s.d_state, sizeof(SeqState), cudaMemcpyDeviceToHost));
    s.len += 1; steps += 1; launches += nl;
    return B2_OK;
}

// One eager decode step through the persistent kernel with per-CTA, per-stage time stamps.
int Model::mega_trace(int seq, int32_t tok, unsigned long long* out, int cap, int* n_cta, int* n_stage) {
    B2_TRY(check_seq(seq));
    Slot& s = slots[seq];
    if (!s.mega_ok) B2_FAIL(B2_ERR_STATE, "mega_trace: the persistent kernel is not in use for this model");
    const int nst = s.mega.s_end - s.mega.s_begin, ncta = s.mega.grid;
    *n_cta = ncta; *n_stage = nst;
    if (!out || cap < ncta * nst * 16) B2_FAIL(B2_ERR_ARG, "mega_trace: need room for %d values", ncta * nst * 16);
    B2_TRY(ensure_pages(s, s.len + 1));
    unsigned long long* d_tr = nullptr;
    B2_CUDA(cudaMalloc(&d_tr, (size_t)ncta * nst * 16 * 8));
    B2_CUDA(cudaMemset(d_tr, 0, (size_t)ncta * nst * 16 * 8));
    SeqState st{};
    st.pos = s.len; st.cur_tok = tok;
    B2_CUDA(cudaMemcpyAsync(s.d_state, &st, sizeof(st), cudaMemcpyHostToDevice, stream));
    s.mega.trace = d_tr;
    int nl = 0;
    const int rc = enqueue_step(s, true, &nl, nullptr);
    s.mega.trace = nullptr;
    cudaStreamSynchronize(stream);
    if (rc == B2_OK) cudaMemcpy(out, d_tr, (size_t)ncta * nst * 16 * 8, cudaMemcpyDeviceToHost);
    cudaFree(d_tr);
    B2_TRY(rc);
    B2_CUDA(cudaMemcpy(&s.host, s.d_state, sizeof(SeqState), cudaMemcpyDeviceToHost));
    s.len += 1; steps += 1; launches += nl;
    return B2_OK;
}

int Model::kv_read(int seq, int layer, uint16_t* k_out, uint16_t* v_out, int n_tok) {
    B2_TRY(check_seq(seq));
    Slot& s = slots[seq];
    if (layer < 0 || layer >= hp.n_layer || n_tok < 0 || n_tok > s.len) B2_FAIL(B2_ERR_ARG, "kv_read: layer %d / n_tok %d out of range", layer, n_tok);
    const size_t nkv = (size_t)n_head_kv_l * hp.head_dim;      // this shard's KV heads
    for (int t0 = 0; t0 < n_tok; t0 += PAGE_TOKENS) {
        const int cnt = std::min(PAGE_TOKENS, n_tok - t0);
        const size_t src = ((size_t)s.pages[t0 / PAGE_TOKENS] * PAGE_TOKENS) * nkv + layer_kv_elems * layer;
        if (k_out) B2_CUDA(cudaMemcpy(k_out + (size_t)t0 * nkv, kv_k + src, cnt * nkv * 2, cudaMemcpyDeviceToHost));
        if (v_out) B2_CUDA(cudaMemcpy(v_out + (size_t)t0 * nkv, kv_v + src, cnt * nkv * 2, cudaMemcpyDeviceToHost));
    }
    return B2_OK;
}

}  // namespace b2

// api.cu -- the extern "C" surface declared in include/b2gguf.h.
#include <string.h>

#include <new>
#include <string>
#include <thread>
#include <vector>

#include "engine.h"

struct b2_ctx {
    std::vector<int> devs;
};
struct b2_model {
    b2_ctx* ctx = nullptr;
    b2::Model m;                       // the model, or shard 0 of a tensor-parallel group
    std::vector<b2::Model*> peers;     // shards 1 .. tp_size-1, one per GPU
    ~b2_model() {
        for (size_t i = 0; i < peers.size(); i++) {
            cudaSetDevice(ctx->devs[i + 1]);
            delete peers[i];
        }
        cudaSetDevice(ctx->devs[0]);
    }
};

using namespace b2;

namespace {

struct DevBuf {          // scoped device allocation for the host-buffer hooks
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    int alloc(size_t n) {
        B2_CUDA(cudaMalloc(&p, n ? n : 1));
        return B2_OK;
    }
    template <typename T> T* as() { return (T*)p; }
};

int check_device() {
    int n = 0;
    const cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {
        cudaGetLastError();
        B2_FAIL(B2_ERR_CUDA, "no CUDA device available (%s); this library has no CPU fallback", e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    }
    return B2_OK;
}

int alloc_qmat(QMat& m, DevBuf planes[4], int type, int64_t rows, int64_t k) {
    qmat_shape(m, type, rows, k);
    for (int p = 0; p < m.n_planes; p++) {
        B2_TRY(planes[p].alloc((size_t)rows * m.stride[p]));
        B2_CUDA(cudaMemset(planes[p].p, 0, (size_t)rows * m.stride[p]));
        m.plane[p] = planes[p].as<uint8_t>();
    }
    return B2_OK;
}

// Run `f(shard, rank)` on every shard of the model: shard 0 on the calling thread, the others on their
// own threads (each shard's kernels wait for its peers inside the all-reduce, so the shards must be
// driven concurrently).  Returns the first failure, with its message.
template <class F>
int tp_call(b2_model* m, F f) {
    if (!m) B2_FAIL(B2_ERR_ARG, "null model");
    const int n = 1 + (int)m->peers.size();
    if (n == 1) {
        B2_CUDA(cudaSetDevice(m->ctx->devs[0]));
        return f(m->m, 0);
    }
    std::vector<int> rc(n, 0);
    std::vector<std::string> err(n);
    std::vector<std::thread> th;
    for (int r = 1; r < n; r++)
        th.emplace_back([&, r] {
            if (cudaSetDevice(m->ctx->devs[r]) != cudaSuccess) { rc[r] = B2_ERR_CUDA; err[r] = "cudaSetDevice failed"; return; }
            rc[r] = f(*m->peers[r - 1], r);
            if (rc[r] != B2_OK) err[r] = last_error();
        });
    if (cudaSetDevice(m->ctx->devs[0]) != cudaSuccess) { rc[0] = B2_ERR_CUDA; err[0] = "cudaSetDevice failed"; }
    else {
        rc[0] = f(m->m, 0);
        if (rc[0] != B2_OK) err[0] = last_error();
    }
    for (auto& t : th) t.join();
    for (int r = 0; r < n; r++)
        if (rc[r] != B2_OK) { set_error("%s", err[r].c_str()); return rc[r]; }
    return B2_OK;
}

int check_quant_shape(int type, int64_t rows, int64_t k) {
    if (!type_is_quant(type)) B2_FAIL(B2_ERR_UNSUPPORTED, "ggml type %d is not supported (Q4_0=2, Q8_0=8, Q4_K=12, Q6_K=14)", type);
    if (rows <= 0 || k <= 0 || k % type_block_elems(type)) B2_FAIL(B2_ERR_ARG, "rows=%lld k=%lld: k must be a positive multiple of %d", (long long)rows,
                                                                    (long long)k, type_block_elems(type));
    return B2_OK;
}

}  // namespace

extern "C" {

const char* b2_strerror(void) { return last_error(); }

int b2_init(int n_dev, const int* dev_ids, b2_ctx** out) {
    if (!out || n_dev <= 0) B2_FAIL(B2_ERR_ARG, "b2_init: need n_dev >= 1 and an out pointer");
    B2_TRY(check_device());
    int have = 0;
    B2_CUDA(cudaGetDeviceCount(&have));
    b2_ctx* c = new (std::nothrow) b2_ctx;
    if (!c) B2_FAIL(B2_ERR_NOMEM, "out of host memory");
    for (int i = 0; i < n_dev; i++) {
        const int d = dev_ids ? dev_ids[i] : i;
        if (d < 0 || d >= have) { delete c; B2_FAIL(B2_ERR_ARG, "b2_init: device %d not present (%d visible)", d, have); }
        c->devs.push_back(d);
    }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, c->devs[0]) == cudaSuccess && prop.major < 9) {
        delete c;
        B2_FAIL(B2_ERR_UNSUPPORTED, "device compute capability %d.%d: this library is built for sm_100a only", prop.major, prop.minor);
    }
    *out = c;
    return B2_OK;
}

void b2_shutdown(b2_ctx* ctx) { delete ctx; }

int b2_model_load(b2_ctx* ctx, const char* gguf_path, const b2_load_opts* opts, b2_model** out) {
    if (!ctx || !gguf_path || !out) B2_FAIL(B2_ERR_ARG, "b2_model_load: null argument");
    B2_CUDA(cudaSetDevice(ctx->devs[0]));
    b2_model* m = new (std::nothrow) b2_model;
    if (!m) B2_FAIL(B2_ERR_NOMEM, "out of host memory");
    m->ctx = ctx;
    const int tp = opts && opts->tp_size > 1 ? opts->tp_size : 1;
    if (tp > TP_MAX || tp > (int)ctx->devs.size()) {
        delete m;
        B2_FAIL(B2_ERR_ARG, "b2_model_load: tp_size=%d needs that many devices in b2_init (have %d, at most %d)", tp, (int)ctx->devs.size(), TP_MAX);
    }
    for (int r = 1; r < tp; r++) m->peers.push_back(new b2::Model);
    // every shard reads its slices of the file on its own thread
    int rc = tp_call(m, [&](Model& s, int r) { return s.load(gguf_path, opts, r, tp); });
    if (rc == B2_OK && tp > 1) {
        std::vector<Model*> all{&m->m};
        for (Model* p : m->peers) all.push_back(p);
        for (int r = 0; r < tp && rc == B2_OK; r++) rc = all[r]->tp_connect(all.data());
        cudaSetDevice(ctx->devs[0]);
    }
    if (rc != B2_OK) { delete m; return rc; }
    *out = m;
    return B2_OK;
}

void b2_model_free(b2_model* m) {
    if (!m) return;
    cudaSetDevice(m->ctx->devs[0]);
    delete m;
}

int b2_model_get_info(const b2_model* m, b2_model_info* info) {
    if (!m || !info) B2_FAIL(B2_ERR_ARG, "b2_model_get_info: null argument");
    const HParams& h = m->m.hp;
    memset(info, 0, sizeof(*info));
    info->arch = h.arch; info->n_layer = h.n_layer; info->n_embd = h.n_embd; info->n_head = h.n_head; info->n_head_kv = h.n_head_kv;
    info->head_dim = h.head_dim; info->n_ff = h.n_ff; info->n_vocab = h.n_vocab; info->n_ctx = h.n_ctx; info->n_rot = h.n_rot;
    info->max_batch = m->m.max_batch; info->max_ctx = m->m.max_ctx; info->norm_eps = h.norm_eps; info->rope_freq_base = h.rope_freq_base;
    info->weight_bytes = m->m.weight_bytes; info->step_weight_bytes = m->m.step_weight_bytes;     // per GPU under tensor parallelism
    for (const Model* p : m->peers) { info->weight_bytes += p->weight_bytes; info->step_weight_bytes += p->step_weight_bytes; }
    info->kv_bytes_per_token = (uint64_t)h.n_layer * 2 * h.n_head_kv * h.head_dim * 2;
    return B2_OK;
}

#define B2_MODEL_CALL(m, expr)                                  \
    do {                                                        \
        if (!(m)) B2_FAIL(B2_ERR_ARG, "null model");            \
        B2_CUDA(cudaSetDevice((m)->ctx->devs[0]));              \
        return (expr);                                          \
    } while (0)

int b2_seq_new(b2_model* m, int* seq_out) {
    if (!seq_out) B2_FAIL(B2_ERR_ARG, "b2_seq_new: null out pointer");
    return tp_call(m, [&](Model& s, int r) { int dummy; return s.seq_new(r == 0 ? seq_out : &dummy); });      // every shard hands out the same slot
}
int b2_seq_free(b2_model* m, int seq) { return tp_call(m, [&](Model& s, int) { return s.seq_free(seq); }); }
int b2_seq_len(const b2_model* m, int seq) {
    if (!m) B2_FAIL(B2_ERR_ARG, "null model");
    B2_TRY(m->m.check_seq(seq));
    return m->m.slots[seq].len;
}
int b2_seq_truncate(b2_model* m, int seq, int n_keep) { return tp_call(m, [&](Model& s, int) { return s.seq_truncate(seq, n_keep); }); }
int b2_seq_assume(b2_model* m, int seq, int n_tokens) { return tp_call(m, [&](Model& s, int) { return s.seq_assume(seq, n_tokens); }); }
int b2_prefill(b2_model* m, int seq, const int32_t* toks, int n, float* logits) {
    return tp_call(m, [&](Model& s, int r) { return s.prefill(seq, toks, n, r == 0 ? logits : nullptr); });
}
int b2_decode(b2_model* m, int n_seq, const int* seqs, const int32_t* last_tok, int32_t* next_tok, float* logits) {
    return tp_call(m, [&](Model& s, int r) { return s.decode(n_seq, seqs, last_tok, r == 0 ? next_tok : nullptr, r == 0 ? logits : nullptr); });
}
int b2_decode_greedy(b2_model* m, int seq, int32_t first_tok, int n_steps, int32_t* out_toks) {
    return tp_call(m, [&](Model& s, int r) { return s.decode_greedy(seq, first_tok, n_steps, r == 0 ? out_toks : nullptr); });
}
int b2_kv_read(b2_model* m, int seq, int layer, uint16_t* k_out, uint16_t* v_out, int n_tok) {
    if (!m) B2_FAIL(B2_ERR_ARG, "null model");
    if (m->peers.empty()) B2_MODEL_CALL(m, m->m.kv_read(seq, layer, k_out, v_out, n_tok));
    // tensor parallel: every shard holds its KV heads; stitch the rows together
    if (n_tok < 0) B2_FAIL(B2_ERR_ARG, "kv_read: n_tok %d", n_tok);
    const int tp = 1 + (int)m->peers.size();
    const size_t w = (size_t)m->m.n_head_kv_l * m->m.hp.head_dim;
    std::vector<uint16_t> kt(w * n_tok), vt(w * n_tok);
    for (int r = 0; r < tp; r++) {
        Model& s = r == 0 ? m->m : *m->peers[r - 1];
        B2_CUDA(cudaSetDevice(m->ctx->devs[r]));
        B2_TRY(s.kv_read(seq, layer, k_out ? kt.data() : nullptr, v_out ? vt.data() : nullptr, n_tok));
        for (int t = 0; t < n_tok; t++) {
            if (k_out) memcpy(k_out + ((size_t)t * tp + r) * w, kt.data() + (size_t)t * w, w * 2);
            if (v_out) memcpy(v_out + ((size_t)t * tp + r) * w, vt.data() + (size_t)t * w, w * 2);
        }
    }
    cudaSetDevice(m->ctx->devs[0]);
    return B2_OK;
}
int b2_profile_step(b2_model* m, int seq, int32_t tok, float* class_ms, int* class_launches) {
    float ms[9]; int nl[9];
    return tp_call(m, [&](Model& s, int r) { return s.profile_step(seq, tok, r == 0 ? class_ms : ms, r == 0 ? class_launches : nl); });
}
int b2_debug_read(b2_model* m, int seq, int what, float* out, int n) {
    if (!m || !out) B2_FAIL(B2_ERR_ARG, "b2_debug_read: null argument");
    B2_TRY(m->m.check_seq(seq));
    const Slot& s = m->m.slots[seq];
    const HParams& h = m->m.hp;
    const float* src = nullptr;
    int cap = 0;
    // shard 0's buffers: under tensor parallelism q / att / h hold this shard's heads and FFN rows only
    switch (what) {
    case 0: src = s.x; cap = h.n_embd; break;
    case 1: src = s.q; cap = m->m.n_head_l * h.head_dim; break;
    case 2: src = s.att; cap = m->m.n_head_l * h.head_dim; break;
    case 3: src = s.h; cap = m->m.n_ff_l; break;
    case 4: src = s.logits; cap = h.n_vocab; break;
    case 5: src = s.att2; cap = h.n_embd; break;
    default: B2_FAIL(B2_ERR_ARG, "b2_debug_read: unknown buffer %d", what);
    }
    if (n < 0 || n > cap) B2_FAIL(B2_ERR_ARG, "b2_debug_read: n=%d outside [0,%d]", n, cap);
    B2_CUDA(cudaSetDevice(m->ctx->devs[0]));
    B2_CUDA(cudaMemcpy(out, src, (size_t)n * 4, cudaMemcpyDeviceToHost));
    return B2_OK;
}
int b2_mega_trace(b2_model* m, int seq, int32_t tok, uint64_t* out, int cap, int* n_cta, int* n_stage) {
    if (!n_cta || !n_stage) B2_FAIL(B2_ERR_ARG, "b2_mega_trace: null argument");
    // a traced step runs shard 0 alone; its in-kernel exchange stages would wait for peers that never launch
    if (m && m->m.tp_size > 1) B2_FAIL(B2_ERR_UNSUPPORTED, "b2_mega_trace: not available for tensor-parallel models");
    B2_MODEL_CALL(m, m->m.mega_trace(seq, tok, (unsigned long long*)out, cap, n_cta, n_stage));
}
int b2_get_stats(const b2_model* m, b2_stats* out) {
    if (!m || !out) B2_FAIL(B2_ERR_ARG, "b2_get_stats: null argument");
    out->steps = m->m.steps; out->kernel_launches = m->m.launches; out->gpu_ms = m->m.last_ms;
    out->last_steps = m->m.last_steps; out->last_launches = m->m.last_launches;
    return B2_OK;
}

// ---- hooks ---------------------------------------------------------------------------------------
int b2_dequantize(int ggml_type, const void* blocks, size_t n_elem, float* out) {
    if (n_elem == 0) return B2_OK;
    if (!blocks || !out) B2_FAIL(B2_ERR_ARG, "b2_dequantize: null buffer");
    const int be = type_block_elems(ggml_type), bb = type_block_bytes(ggml_type);
    if (bb == 0) B2_FAIL(B2_ERR_UNSUPPORTED, "ggml type %d is not supported", ggml_type);
    if (n_elem % be) B2_FAIL(B2_ERR_ARG, "b2_dequantize: %zu elements is not a whole number of %d-element blocks", n_elem, be);
    B2_TRY(check_device());
    const size_t nbytes = n_elem / be * bb;
    DevBuf src, dst;
    B2_TRY(src.alloc(nbytes));
    B2_TRY(dst.alloc(n_elem * 4));
    B2_CUDA(cudaMemcpy(src.p, blocks, nbytes, cudaMemcpyHostToDevice));
    B2_TRY(dequant_raw(ggml_type, src.p, (int64_t)n_elem, dst.as<float>(), 0));
    B2_CUDA(cudaMemcpy(out, dst.p, n_elem * 4, cudaMemcpyDeviceToHost));
    return B2_OK;
}

int b2_dequantize_repacked(int ggml_type, const void* blocks, int64_t rows, int64_t k, float* out) {
    if (!blocks || !out) B2_FAIL(B2_ERR_ARG, "b2_dequantize_repacked: null buffer");
    B2_TRY(check_quant_shape(ggml_type, rows, k));
    B2_TRY(check_device());
    const size_t nbytes = (size_t)rows * type_row_bytes(ggml_type, k);
    DevBuf src, dst, planes[4];
    QMat m;
    B2_TRY(src.alloc(nbytes));
    B2_TRY(dst.alloc((size_t)rows * k * 4));
    B2_TRY(alloc_qmat(m, planes, ggml_type, rows, k));
    B2_CUDA(cudaMemcpy(src.p, blocks, nbytes, cudaMemcpyHostToDevice));
    B2_TRY(repack(ggml_type, src.p, rows, k, m, 0, 1, 0));
    B2_TRY(dequant_rows(m, 0, rows, dst.as<float>(), 0));
    B2_CUDA(cudaMemcpy(out, dst.p, (size_t)rows * k * 4, cudaMemcpyDeviceToHost));
    return B2_OK;
}

int b2_gemv(int ggml_type, const void* W, int64_t rows, int64_t k, const float* x, const float* gain, float eps, float* y) {
    if (!W || !x || !y) B2_FAIL(B2_ERR_ARG, "b2_gemv: null buffer");
    B2_TRY(check_quant_shape(ggml_type, rows, k));
    if (rows & 1) B2_FAIL(B2_ERR_ARG, "b2_gemv: rows=%lld must be even", (long long)rows);
    B2_TRY(check_device());
    const size_t nbytes = (size_t)rows * type_row_bytes(ggml_type, k);
    DevBuf src, dx, dg, dy, planes[4];
    QMat m;
    B2_TRY(src.alloc(nbytes));
    B2_TRY(dx.alloc((size_t)k * 4));
    B2_TRY(dg.alloc((size_t)k * 4));
    B2_TRY(dy.alloc((size_t)rows * 4));
    B2_TRY(alloc_qmat(m, planes, ggml_type, rows, k));
    B2_CUDA(cudaMemcpy(src.p, W, nbytes, cudaMemcpyHostToDevice));
    B2_CUDA(cudaMemcpy(dx.p, x, (size_t)k * 4, cudaMemcpyHostToDevice));
    if (gain) B2_CUDA(cudaMemcpy(dg.p, gain, (size_t)k * 4, cudaMemcpyHostToDevice));
    B2_TRY(repack(ggml_type, src.p, rows, k, m, 0, 1, 0));
    GemvArgs a;
    a.W = m; a.pro = gain ? PRO_RMSNORM : PRO_PLAIN; a.epi = EPI_STORE; a.x = dx.as<float>(); a.gain = dg.as<float>(); a.eps = eps; a.y = dy.as<float>();
    B2_TRY(gemv_launch(a, 0, false));
    B2_CUDA(cudaDeviceSynchronize());
    B2_CUDA(cudaMemcpy(y, dy.p, (size_t)rows * 4, cudaMemcpyDeviceToHost));
    return B2_OK;
}

int b2_gemv_bench(int ggml_type, int64_t rows, int64_t k, int iters, const int* tune, float* ms_out) {
    if (!ms_out || iters <= 0) B2_FAIL(B2_ERR_ARG, "b2_gemv_bench: bad arguments");
    B2_TRY(check_quant_shape(ggml_type, rows, k));
    B2_TRY(check_device());
    // enough distinct matrices that consecutive launches never hit L2 (126 MB)
    QMat shape;
    qmat_shape(shape, ggml_type, rows, k);
    const size_t mat_bytes = shape.bytes();
    const int n_mats = (int)std::min<size_t>(64, std::max<size_t>(2, (size_t)(400u << 20) / std::max<size_t>(mat_bytes, 1) + 1));
    std::vector<DevBuf> bufs((size_t)n_mats * 4);
    std::vector<QMat> mats(n_mats);
    for (int i = 0; i < n_mats; i++) {
        qmat_shape(mats[i], ggml_type, rows, k);
        for (int p = 0; p < mats[i].n_planes; p++) {
            DevBuf& b = bufs[(size_t)i * 4 + p];
            B2_TRY(b.alloc((size_t)rows * mats[i].stride[p]));
            B2_CUDA(cudaMemset(b.p, 0x11 + i, (size_t)rows * mats[i].stride[p]));
            mats[i].plane[p] = b.as<uint8_t>();
        }
    }
    DevBuf dx, dy;
    B2_TRY(dx.alloc((size_t)k * 4));
    B2_TRY(dy.alloc((size_t)rows * 4));
    B2_CUDA(cudaMemset(dx.p, 0, (size_t)k * 4));
    GemvTune t;
    if (tune) { t.nb = tune[0]; t.g = tune[1]; t.rpg = tune[2]; t.ns = tune[3]; t.ctas_per_sm = tune[4]; }
    cudaStream_t st;
    B2_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    cudaEvent_t e0, e1;
    B2_CUDA(cudaEventCreate(&e0));
    B2_CUDA(cudaEventCreate(&e1));
    int rc = B2_OK;
    for (int pass = 0; pass < 2 && rc == B2_OK; pass++) {      // pass 0 = warm-up
        const int n = pass == 0 ? std::min(iters, 3) : iters;
        cudaEventRecord(e0, st);
        for (int i = 0; i < n && rc == B2_OK; i++) {
            GemvArgs a;
            a.W = mats[i % n_mats]; a.pro = PRO_PLAIN; a.epi = EPI_STORE; a.x = dx.as<float>(); a.y = dy.as<float>();
            rc = gemv_launch(a, st, true, tune ? &t : nullptr);
        }
        cudaEventRecord(e1, st);
        if (cudaStreamSynchronize(st) != cudaSuccess && rc == B2_OK) { set_error("b2_gemv_bench: %s", cudaGetErrorString(cudaGetLastError())); rc = B2_ERR_CUDA; }
        if (pass == 1 && rc == B2_OK) {
            float ms = 0;
            cudaEventElapsedTime(&ms, e0, e1);
            *ms_out = ms / iters;
        }
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaStreamDestroy(st);
    return rc;
}

int b2_attn_decode(const float* q, const uint16_t* k, const uint16_t* v, int n_ctx, int n_head, int n_head_kv, int head_dim, float scale,
                   int cluster, float* out) {
    if (!q || !k || !v || !out || n_ctx <= 0 || n_head <= 0 || n_head_kv <= 0 || head_dim <= 0) B2_FAIL(B2_ERR_ARG, "b2_attn_decode: bad arguments");
    B2_TRY(check_device());
    const size_t kvd = (size_t)n_head_kv * head_dim, nq = (size_t)n_head * head_dim;
    const int n_pages = (n_ctx + PAGE_TOKENS - 1) / PAGE_TOKENS;
    DevBuf dq, dk, dv, dout, dpt, dpos;
    B2_TRY(dq.alloc(nq * 4)); B2_TRY(dout.alloc(nq * 4));
    B2_TRY(dk.alloc((size_t)n_pages * PAGE_TOKENS * kvd * 2)); B2_TRY(dv.alloc((size_t)n_pages * PAGE_TOKENS * kvd * 2));
    B2_TRY(dpt.alloc((size_t)n_pages * 4)); B2_TRY(dpos.alloc(4));
    // scatter the rows over physical pages in reverse order so the page table is exercised
    std::vector<int> pt(n_pages);
    for (int i = 0; i < n_pages; i++) pt[i] = n_pages - 1 - i;
    for (int i = 0; i < n_pages; i++) {
        const int t0 = i * PAGE_TOKENS, cnt = std::min(PAGE_TOKENS, n_ctx - t0);
        B2_CUDA(cudaMemcpy(dk.as<__half>() + (size_t)pt[i] * PAGE_TOKENS * kvd, k + (size_t)t0 * kvd, cnt * kvd * 2, cudaMemcpyHostToDevice));
        B2_CUDA(cudaMemcpy(dv.as<__half>() + (size_t)pt[i] * PAGE_TOKENS * kvd, v + (size_t)t0 * kvd, cnt * kvd * 2, cudaMemcpyHostToDevice));
    }
    const int pos = n_ctx - 1;
    B2_CUDA(cudaMemcpy(dpt.p, pt.data(), (size_t)n_pages * 4, cudaMemcpyHostToDevice));
    B2_CUDA(cudaMemcpy(dpos.p, &pos, 4, cudaMemcpyHostToDevice));
    B2_CUDA(cudaMemcpy(dq.p, q, nq * 4, cudaMemcpyHostToDevice));
    AttnArgs a;
    a.q = dq.as<float>(); a.out = dout.as<float>(); a.k_base = dk.as<__half>(); a.v_base = dv.as<__half>(); a.page_table = dpt.as<int>();
    a.pos = dpos.as<int>(); a.n_head = n_head; a.n_head_kv = n_head_kv; a.head_dim = head_dim; a.page_tokens = PAGE_TOKENS; a.kv_dim = (int)kvd;
    a.max_ctx = n_ctx; a.scale = scale;
    B2_TRY(attn_decode_launch(a, 0, false, cluster));
    B2_CUDA(cudaDeviceSynchronize());
    B2_CUDA(cudaMemcpy(out, dout.p, nq * 4, cudaMemcpyDeviceToHost));
    return B2_OK;
}

int b2_attn_bench(int n_ctx, int n_head, int n_head_kv, int head_dim, int cluster, int n_layers, int iters, float* ms_out) {
    if (!ms_out || n_ctx <= 0 || iters <= 0 || n_layers <= 0) B2_FAIL(B2_ERR_ARG, "b2_attn_bench: bad arguments");
    B2_TRY(check_device());
    const size_t kvd = (size_t)n_head_kv * head_dim, nq = (size_t)n_head * head_dim;
    const int n_pages = (n_ctx + PAGE_TOKENS - 1) / PAGE_TOKENS;
    const size_t layer_elems = (size_t)n_pages * PAGE_TOKENS * kvd;
    DevBuf dq, dk, dv, dout, dpt, dpos;
    B2_TRY(dq.alloc(nq * 4)); B2_TRY(dout.alloc(nq * 4));
    B2_TRY(dk.alloc(layer_elems * 2 * n_layers)); B2_TRY(dv.alloc(layer_elems * 2 * n_layers));
    B2_TRY(dpt.alloc((size_t)n_pages * 4)); B2_TRY(dpos.alloc(4));
    B2_CUDA(cudaMemset(dk.p, 0, layer_elems * 2 * n_layers)); B2_CUDA(cudaMemset(dv.p, 0, layer_elems * 2 * n_layers));
    B2_CUDA(cudaMemset(dq.p, 0, nq * 4));
    std::vector<int> pt(n_pages);
    for (int i = 0; i < n_pages; i++) pt[i] = i;
    const int pos = n_ctx - 1;
    B2_CUDA(cudaMemcpy(dpt.p, pt.data(), (size_t)n_pages * 4, cudaMemcpyHostToDevice));
    B2_CUDA(cudaMemcpy(dpos.p, &pos, 4, cudaMemcpyHostToDevice));
    cudaStream_t st;
    B2_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    cudaEvent_t e0, e1;
    B2_CUDA(cudaEventCreate(&e0)); B2_CUDA(cudaEventCreate(&e1));
    int rc = B2_OK;
    for (int pass = 0; pass < 2 && rc == B2_OK; pass++) {
        const int n = pass == 0 ? std::min(iters, 3) : iters;
        cudaEventRecord(e0, st);
        for (int i = 0; i < n && rc == B2_OK; i++) {
            AttnArgs a;
            const int l = i % n_layers;        // a different layer's cache every launch, like the decode loop
            a.q = dq.as<float>(); a.out = dout.as<float>(); a.k_base = dk.as<__half>() + layer_elems * l; a.v_base = dv.as<__half>() + layer_elems * l;
            a.page_table = dpt.as<int>(); a.pos = dpos.as<int>(); a.n_head = n_head; a.n_head_kv = n_head_kv; a.head_dim = head_dim;
            a.page_tokens = PAGE_TOKENS; a.kv_dim = (int)kvd; a.max_ctx = n_ctx; a.scale = 1.0f;
            rc = attn_decode_launch(a, st, true, cluster);
        }
        cudaEventRecord(e1, st);
        if (cudaStreamSynchronize(st) != cudaSuccess && rc == B2_OK) { set_error("b2_attn_bench: %s", cudaGetErrorString(cudaGetLastError())); rc = B2_ERR_CUDA; }
        if (pass == 1 && rc == B2_OK) { float ms = 0; cudaEventElapsedTime(&ms, e0, e1); *ms_out = ms / iters; }
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaStreamDestroy(st);
    return rc;
}

int b2_gemm(int ggml_type, const void* W, int64_t rows, int64_t k, const float* x, int n_cols, const float* gain, float eps, float* y, int iters,
            float* ms_out) {
    if (!W || !x || !y || n_cols < 1 || n_cols > BATCH_MAX) B2_FAIL(B2_ERR_ARG, "b2_gemm: bad arguments (1..%d columns)", BATCH_MAX);
    B2_TRY(check_quant_shape(ggml_type, rows, k));
    B2_TRY(check_device());
    const size_t nbytes = (size_t)rows * type_row_bytes(ggml_type, k);
    DevBuf src, dx, dg, dy, planes[4], dq, dd, ds, dargs, dpos, dpt;
    QMat m;
    B2_TRY(src.alloc(nbytes)); B2_TRY(dx.alloc((size_t)n_cols * k * 4)); B2_TRY(dg.alloc((size_t)k * 4)); B2_TRY(dy.alloc((size_t)n_cols * rows * 4));
    B2_TRY(alloc_qmat(m, planes, ggml_type, rows, k));
    B2_TRY(dq.alloc(q8cols_q_bytes(k))); B2_TRY(dd.alloc(q8cols_d_bytes(k))); B2_TRY(ds.alloc(q8cols_d_bytes(k)));
    B2_CUDA(cudaMemset(dq.p, 0, q8cols_q_bytes(k))); B2_CUDA(cudaMemset(dd.p, 0, q8cols_d_bytes(k))); B2_CUDA(cudaMemset(ds.p, 0, q8cols_d_bytes(k)));
    B2_TRY(dargs.alloc(sizeof(GemvArgs) * BATCH_MAX)); B2_TRY(dpos.alloc(BATCH_MAX * 4)); B2_TRY(dpt.alloc(BATCH_MAX * 4));
    B2_CUDA(cudaMemset(dpos.p, 0, BATCH_MAX * 4)); B2_CUDA(cudaMemset(dpt.p, 0, BATCH_MAX * 4));
    B2_CUDA(cudaMemcpy(src.p, W, nbytes, cudaMemcpyHostToDevice));
    B2_CUDA(cudaMemcpy(dx.p, x, (size_t)n_cols * k * 4, cudaMemcpyHostToDevice));
    if (gain) B2_CUDA(cudaMemcpy(dg.p, gain, (size_t)k * 4, cudaMemcpyHostToDevice));
    B2_TRY(repack(ggml_type, src.p, rows, k, m, 0, 1, 0));
    GemmMaps maps;
    B2_TRY(gemm_make_maps(m, &maps));
    std::vector<GemvArgs> ca(n_cols);
    for (int c = 0; c < n_cols; c++) { ca[c].epi = EPI_STORE; ca[c].y = dy.as<float>() + (size_t)c * rows; }
    B2_CUDA(cudaMemcpy(dargs.p, ca.data(), sizeof(GemvArgs) * n_cols, cudaMemcpyHostToDevice));
    Q8Cols q8; q8.q = dq.as<uint8_t>(); q8.d = dd.as<float>(); q8.s8 = ds.as<int>(); q8.k_cap = k;
    B2_TRY(quant_cols_launch(dx.as<float>(), k, k, n_cols, gain ? PRO_RMSNORM : PRO_PLAIN, dg.as<float>(), nullptr, eps, q8, 0));
    cudaEvent_t e0, e1;
    B2_CUDA(cudaEventCreate(&e0)); B2_CUDA(cudaEventCreate(&e1));
    const int n = iters > 0 ? iters : 1;
    B2_TRY(gemm_w4a8_launch(m, maps, q8, n_cols, dargs.as<GemvArgs>(), dpos.as<int>(), dpt.as<int>(), 1, 0));
    cudaEventRecord(e0, 0);
    for (int i = 0; i < n; i++) B2_TRY(gemm_w4a8_launch(m, maps, q8, n_cols, dargs.as<GemvArgs>(), dpos.as<int>(), dpt.as<int>(), 1, 0));
    cudaEventRecord(e1, 0);
    B2_CUDA(cudaDeviceSynchronize());
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (ms_out) *ms_out = ms / n;
    B2_CUDA(cudaMemcpy(y, dy.p, (size_t)n_cols * rows * 4, cudaMemcpyDeviceToHost));
    return B2_OK;
}

int b2_gemm_tc(int ggml_type, const void* W, int64_t rows, int64_t k, const float* x, int n_cols, const float* gain, float eps, int acc32, float* y, int iters,
               float* ms_out) {
    if (!W || !x || !y || n_cols < 1 || n_cols > TC_COLS_MAX) B2_FAIL(B2_ERR_ARG, "b2_gemm_tc: bad arguments (1..%d columns)", TC_COLS_MAX);
    B2_TRY(check_quant_shape(ggml_type, rows, k));
    B2_TRY(check_device());
    const size_t nbytes = (size_t)rows * type_row_bytes(ggml_type, k);
    DevBuf src, dx, dg, dy, planes[4], dq, dd, dargs, dpos, dpt;
    QMat m;
    B2_TRY(src.alloc(nbytes)); B2_TRY(dx.alloc((size_t)n_cols * k * 4)); B2_TRY(dg.alloc((size_t)k * 4)); B2_TRY(dy.alloc((size_t)n_cols * rows * 4));
    B2_TRY(alloc_qmat(m, planes, ggml_type, rows, k));
    if (!gemm_tc_supported(m)) B2_FAIL(B2_ERR_UNSUPPORTED, "b2_gemm_tc: Q4_0 with k a multiple of 256 and an even row count only");
    B2_TRY(dq.alloc((size_t)TC_COLS_MAX * k)); B2_TRY(dd.alloc((size_t)(k / 32) * TC_COLS_MAX * 4));
    B2_CUDA(cudaMemset(dq.p, 0, (size_t)TC_COLS_MAX * k)); B2_CUDA(cudaMemset(dd.p, 0, (size_t)(k / 32) * TC_COLS_MAX * 4));
    B2_TRY(dargs.alloc(sizeof(GemvArgs) * TC_COLS_MAX)); B2_TRY(dpos.alloc(TC_COLS_MAX * 4)); B2_TRY(dpt.alloc(TC_COLS_MAX * 4));
    B2_CUDA(cudaMemset(dpos.p, 0, TC_COLS_MAX * 4)); B2_CUDA(cudaMemset(dpt.p, 0, TC_COLS_MAX * 4));
    B2_CUDA(cudaMemcpy(src.p, W, nbytes, cudaMemcpyHostToDevice));
    B2_CUDA(cudaMemcpy(dx.p, x, (size_t)n_cols * k * 4, cudaMemcpyHostToDevice));
    if (gain) B2_CUDA(cudaMemcpy(dg.p, gain, (size_t)k * 4, cudaMemcpyHostToDevice));
    B2_TRY(repack(ggml_type, src.p, rows, k, m, 0, 1, 0));
    TcMaps maps;
    B2_TRY(gemm_tc_make_maps(m, &maps));
    TcCols cols;
    B2_TRY(tc_cols_init(cols, dq.as<int8_t>(), dd.as<float>(), k, TC_COLS_MAX));
    std::vector<GemvArgs> ca(n_cols);
    for (int c = 0; c < n_cols; c++) { ca[c].epi = EPI_STORE; ca[c].y = dy.as<float>() + (size_t)c * rows; }
    B2_CUDA(cudaMemcpy(dargs.p, ca.data(), sizeof(GemvArgs) * n_cols, cudaMemcpyHostToDevice));
    B2_TRY(quant_cols_tc_launch(dx.as<float>(), k, k, n_cols, gain ? PRO_RMSNORM : PRO_PLAIN, dg.as<float>(), nullptr, eps, cols, 0));
    cudaEvent_t e0, e1;
    B2_CUDA(cudaEventCreate(&e0)); B2_CUDA(cudaEventCreate(&e1));
    const int n = iters > 0 ? iters : 1;
    B2_TRY(gemm_tc_launch(m, maps, cols, n_cols, dargs.as<GemvArgs>(), dpos.as<int>(), dpt.as<int>(), 1, 0, false, acc32 != 0));
    B2_CUDA(cudaDeviceSynchronize());
    cudaEventRecord(e0, 0);
    for (int i = 0; i < n; i++) B2_TRY(gemm_tc_launch(m, maps, cols, n_cols, dargs.as<GemvArgs>(), dpos.as<int>(), dpt.as<int>(), 1, 0, false, acc32 != 0));
    cudaEventRecord(e1, 0);
    B2_CUDA(cudaDeviceSynchronize());
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (ms_out) *ms_out = ms / n;
    B2_CUDA(cudaMemcpy(y, dy.p, (size_t)n_cols * rows * 4, cudaMemcpyDeviceToHost));
    return B2_OK;
}

}  // extern "C"

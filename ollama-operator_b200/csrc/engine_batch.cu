// engine_batch.cu -- batched steps: several sequences per b2_decode call and chunked b2_prefill.
//
// llama.cpp evaluates a batch of tokens with the same per-row arithmetic as a single token: every
// activation row is quantised on its own and every (weight row, activation row) dot product is the
// same sum of block terms [upstream ggml_compute_forward_mul_mat].  The batched kernels here keep
// that property bit for bit (tests/test_gpu_model.py compares against the one-token path), so
// batching is purely a scheduling decision: the weights are streamed once per step for up to
// BATCH_MAX columns instead of once per token.
//
// A "column" is one token being evaluated: in a decode call column c belongs to sequence c of the
// call; in a prefill chunk the columns are consecutive prompt positions of one sequence, and column
// c attends to the cache rows the columns before it appended in the same step (the QKV kernel of a
// layer finishes before its attention kernel starts).
#include <math.h>
#include <string.h>

#include <algorithm>

#include "engine.h"

namespace b2 {

static inline size_t align_up_b(size_t v, size_t a) { return (v + a - 1) / a * a; }

int Model::setup_batch() {
    bt.ok = false;
    if (slots.empty()) return B2_OK;
    const int E = hp.n_embd, hd = hp.head_dim, nq = n_head_l * hd, F = n_ff_l, V = hp.n_vocab;      // this shard's heads and FFN rows
    int64_t k_max = 0;
    for (const StepOp& op : slots[0].ops)
        if (op.kind == 0) {
            const QMat& W = op.g.W;
            if (!gemm_supported(W)) return B2_OK;     // per-token path only
            k_max = std::max(k_max, W.k);
        }
    // The tensor-core kernel takes up to TC_COLS_MAX columns per pass over the weights (prompt chunks); it serves a model
    // whose every matrix it supports, on one GPU (the tensor-parallel exchange is sized for BATCH_MAX columns).
    bt.tc_ok = tp_size == 1 && !getenv("B2_NO_TC");
    for (const StepOp& op : slots[0].ops)
        if (op.kind == 0 && !gemm_tc_supported(op.g.W)) bt.tc_ok = false;
    bt.tc_acc32 = bt.tc_ok && getenv("B2_TC_ACC32") != nullptr;
    bt.cap = bt.tc_ok ? TC_COLS_MAX : BATCH_MAX;
    const int CAP = bt.cap;
    size_t off = 0;
    auto carve = [&](size_t bytes) { const size_t o2 = align_up_b(off, 256); off = o2 + bytes; return o2; };
    const size_t o_x = carve((size_t)CAP * E * 4), o_q = carve((size_t)CAP * nq * 4), o_att = carve((size_t)CAP * nq * 4),
                 o_att2 = carve((size_t)CAP * E * 4), o_h = carve((size_t)CAP * F * 4), o_logits = carve((size_t)BATCH_MAX * V * 4),      // logits: the head columns only (at most BATCH_MAX)
                 o_q8 = carve(q8cols_q_bytes(k_max)), o_q8d = carve(q8cols_d_bytes(k_max)), o_q8s = carve(q8cols_d_bytes(k_max)), o_tok = carve((size_t)(2 * CAP + CAP * pages_per_seq) * 4),      // tok | pos | page tables: one upload
                 o_next = carve((size_t)CAP * 4), o_keys = carve((size_t)CAP * 8);
    const size_t o_tcq = bt.tc_ok ? carve((size_t)TC_COLS_MAX * k_max) : 0, o_tcd = bt.tc_ok ? carve((size_t)(k_max / 32) * TC_COLS_MAX * 4) : 0;
    const size_t n_ops = slots[0].ops.size();
    const size_t o_cargs = carve(n_ops * CAP * sizeof(GemvArgs)), o_aargs = carve(n_ops * CAP * sizeof(AttnArgs));
    B2_CUDA(cudaMalloc(&bt.mem, off));
    B2_CUDA(cudaMemset(bt.mem, 0, off));
    uint8_t* b = (uint8_t*)bt.mem;
    bt.x = (float*)(b + o_x); bt.q = (float*)(b + o_q); bt.att = (float*)(b + o_att); bt.att2 = (float*)(b + o_att2); bt.h = (float*)(b + o_h);
    bt.logits = (float*)(b + o_logits);
    bt.q8.q = b + o_q8; bt.q8.d = (float*)(b + o_q8d); bt.q8.s8 = (int*)(b + o_q8s); bt.q8.k_cap = k_max;
    bt.d_tok = (int*)(b + o_tok); bt.d_pos = bt.d_tok + CAP; bt.d_pt = bt.d_tok + 2 * CAP; bt.d_next = (int*)(b + o_next);
    bt.d_keys = (unsigned long long*)(b + o_keys);
    bt.d_colargs = (GemvArgs*)(b + o_cargs); bt.d_attargs = (AttnArgs*)(b + o_aargs);

    // the pipeline of slot 0 re-bound to the column buffers; per-column operand tables
    const Slot& s0 = slots[0];
    bt.ops = s0.ops;
    if (bt.tc_ok) B2_TRY(tc_cols_init(bt.tc, (int8_t*)(b + o_tcq), (float*)(b + o_tcd), k_max, TC_COLS_MAX));
    std::vector<GemvArgs> ca(n_ops * CAP);
    std::vector<AttnArgs> aa(n_ops * CAP);
    auto rebind = [&](const float* p, int c) -> float* {
        if (!p) return nullptr;
        if (p == s0.x) return bt.x + (size_t)c * E;
        if (p == s0.q) return bt.q + (size_t)c * nq;
        if (p == s0.att) return bt.att + (size_t)c * nq;
        if (p == s0.att2) return bt.att2 + (size_t)c * E;
        if (p == s0.h) return bt.h + (size_t)c * F;
        return const_cast<float*>(p);
    };
    for (size_t i = 0; i < n_ops; i++)
        for (int c = 0; c < CAP; c++) {
            if (s0.ops[i].kind == 0) {
                GemvArgs g = s0.ops[i].g;
                g.x = rebind(g.x, c); g.y = rebind(g.y, c); g.y2 = rebind(g.y2, c); g.y3 = rebind(g.y3, c);
                if (s0.ops[i].tp_buf >= 0) g.y64 = tpb.part[s0.ops[i].tp_buf][tp_rank] + (size_t)c * E;      // K-split projection: per-column partial sums
                g.kv.page_table = bt.d_pt + (size_t)c * pages_per_seq;
                g.kv.pos = bt.d_pos + c;
                ca[i * CAP + c] = g;
                if (c == 0) bt.ops[i].g = g;
            } else {
                AttnArgs a = s0.ops[i].a;
                a.q = rebind(a.q, c); a.out = rebind(a.out, c);
                a.page_table = bt.d_pt + (size_t)c * pages_per_seq;
                a.pos = bt.d_pos + c;
                aa[i * CAP + c] = a;
                if (c == 0) bt.ops[i].a = a;
            }
        }
    bt.maps.resize(n_ops);
    bt.tcmaps.resize(bt.tc_ok ? n_ops : 0);
    for (size_t i = 0; i < n_ops; i++)
        if (s0.ops[i].kind == 0) {
            B2_TRY(gemm_make_maps(s0.ops[i].g.W, &bt.maps[i]));
            if (bt.tc_ok) B2_TRY(gemm_tc_make_maps(s0.ops[i].g.W, &bt.tcmaps[i]));
        }
    B2_CUDA(cudaMemcpy(bt.d_colargs, ca.data(), ca.size() * sizeof(GemvArgs), cudaMemcpyHostToDevice));
    B2_CUDA(cudaMemcpy(bt.d_attargs, aa.data(), aa.size() * sizeof(AttnArgs), cudaMemcpyHostToDevice));

    bt.region_ints = 2 * CAP + CAP * pages_per_seq;
    bt.n_regions = max_ctx / CAP + 4;
    B2_CUDA(cudaMallocHost(&bt.h_stage, (size_t)bt.n_regions * bt.region_ints * 4));
    B2_CUDA(cudaMallocHost(&bt.h_next, (size_t)CAP * 4));
    if (const char* v = getenv("B2_BATCH_MIN")) batch_min_seqs = std::max(1, atoi(v));
    bt.ok = true;
    return B2_OK;
}

// One step for B columns.  col_slot[c] owns the KV pages column c writes to and attends over,
// col_pos[c] is the cache row it appends, col_tok[c] the token.  Columns c >= head_from (at most BATCH_MAX of
// them) get logits (bt.logits + (c - head_from) * n_vocab) and a greedy next token (bt.d_next[c]).
// More than BATCH_MAX columns (prompt chunks) run on the tcgen05 kernel, otherwise the mma.sync one.
int Model::batch_step(int B, const int* col_slot, const int* col_pos, const int32_t* col_tok, int head_from) {
    if (!bt.ok) B2_FAIL(B2_ERR_STATE, "batched kernels are not available for this model");
    if (B < 1 || B > bt.cap) B2_FAIL(B2_ERR_ARG, "batch_step: %d columns", B);
    if (head_from < 0 || B - head_from > BATCH_MAX) B2_FAIL(B2_ERR_ARG, "batch_step: %d head columns", B - head_from);
    const bool tc = bt.tc_ok && (B > BATCH_MAX || getenv("B2_TC_ALWAYS"));
    const int CAP = bt.cap;
    batch_pdl_enabled() = use_pdl;              // B2_LOAD_NO_PDL covers the batched launches too
    if (bt.next_region >= bt.n_regions) {          // all upload regions of this call are in flight: drain
        B2_CUDA(cudaStreamSynchronize(stream));
        bt.next_region = 0;
    }
    int* reg = bt.h_stage + (size_t)bt.next_region++ * bt.region_ints;
    int* h_tok = reg; int* h_pos = reg + CAP; int* h_pt = reg + 2 * CAP;
    for (int c = 0; c < B; c++) {
        const Slot& s = slots[col_slot[c]];
        h_tok[c] = col_tok[c];
        h_pos[c] = col_pos[c];
        memcpy(h_pt + (size_t)c * pages_per_seq, s.pages.data(), s.pages.size() * 4);
    }
    // tok, pos and the page tables are contiguous on both sides: one copy
    B2_CUDA(cudaMemcpyAsync(bt.d_tok, reg, (size_t)(2 * CAP + B * pages_per_seq) * 4, cudaMemcpyHostToDevice, stream));
    int nl = 0;
    const int E = hp.n_embd;
    B2_TRY(batch_begin_launch(w_embd, bt.x, E, bt.d_tok, bt.d_keys, B, stream)); nl++;
    for (size_t i = 0; i < bt.ops.size(); i++) {
        const StepOp& op = bt.ops[i];
        if (op.kind == 0) {
            const GemvArgs& g = op.g;
            // the same activation vector feeds every segment of a split QKV matrix: quantise once
            const bool reuse = i > 0 && bt.ops[i - 1].kind == 0 && bt.ops[i - 1].g.x == g.x && bt.ops[i - 1].g.pro == g.pro &&
                               bt.ops[i - 1].g.gain == g.gain && bt.ops[i - 1].g.W.k == g.W.k && g.epi == EPI_QKV_ROPE;
            if (!reuse) {
                const int64_t xs = g.x == bt.x || g.x == bt.att2 ? E : g.x == bt.h ? n_ff_l : n_head_l * hp.head_dim;
                if (tc) B2_TRY(quant_cols_tc_launch(g.x, xs, g.W.k, B, g.pro, g.gain, g.nbias, g.eps, bt.tc, stream));
                else B2_TRY(quant_cols_launch(g.x, xs, g.W.k, B, g.pro, g.gain, g.nbias, g.eps, bt.q8, stream));
                nl++;
            }
            if (tc) B2_TRY(gemm_tc_launch(g.W, bt.tcmaps[i], bt.tc, B, bt.d_colargs + i * CAP, bt.d_pos, bt.d_pt, pages_per_seq, stream, op.tp_buf >= 0, bt.tc_acc32));
            else B2_TRY(gemm_w4a8_launch(g.W, bt.maps[i], bt.q8, B, bt.d_colargs + i * CAP, bt.d_pos, bt.d_pt, pages_per_seq, stream, op.tp_buf >= 0));
            nl++;
            if (op.tp_buf >= 0) {      // all columns at once: x[c] += sum over the shards of their partial sums
                B2_TRY(tp_allreduce_residual_launch(tpb, op.tp_buf, bt.x, B * E, stream, true)); nl++;
            }
        } else {
            B2_TRY(attn_decode_batch_launch(bt.d_attargs + i * CAP, op.a, B, stream)); nl++;
        }
    }
    for (int c = head_from; c < B; c++) {
            GemvArgs h;
            h.W = w_head; h.pro = hp.arch == 0 ? PRO_RMSNORM : PRO_LAYERNORM; h.epi = EPI_LOGITS;
            h.x = bt.x + (size_t)c * E; h.gain = out_norm; h.nbias = out_norm_b; h.eps = hp.norm_eps; h.bias = head_b;
            h.y = bt.logits + (size_t)(c - head_from) * hp.n_vocab; h.argmax_key = bt.d_keys + c;
            B2_TRY(gemv_launch(h, stream, use_pdl)); nl++;      // consecutive head launches overlap their ring priming with the predecessor's tail
        }
    if (head_from < B) { B2_TRY(batch_end_launch(bt.d_keys, bt.d_next, B, stream)); nl++; }
    launches += nl; last_launches += nl;
    steps += 1; last_steps += 1;
    return B2_OK;
}

}  // namespace b2

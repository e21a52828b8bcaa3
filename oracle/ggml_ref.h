/*
 * oracle/ggml_ref.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the arithmetic that runs inside the `ollama/ollama`
 * container the reference deploys (reference call sites: pkg/model/pod.go:14-66
 * `NewOllamaServerContainer` -> `ollama serve`, pkg/model/model.go:85-90).  The
 * algorithm itself lives in a third-party dependency that is absent from
 * /root/reference: the container image docker.io/ollama/ollama, pinned to no
 * version (floating `latest`, pkg/model/pod.go:10-12), which embeds
 * llama.cpp/ggml.  What follows restates ggml's published block formats and
 * reference (scalar) routines: dequantize_row_{q4_0,q8_0,q4_K,q6_K},
 * quantize_row_{q4_0,q8_0,q8_K}_ref, ggml_vec_dot_{q4_0_q8_0,q8_0_q8_0,
 * q4_K_q8_K,q6_K_q8_K} (generic C paths), and the ops rms_norm, norm, rope,
 * soft_max, silu, gelu used by llama.cpp's build_llama / build_phi2 graphs.
 *
 * PARITY STATUS: the reference holds no golden vector, fixture or known-answer
 * test for this path (internal/controller/model_controller_test.go:69-82 only
 * asserts err == nil; test/e2e/e2e_test.go:84-118 only that the manager pod
 * runs) => "parity unpinned" with respect to the reference.  The block
 * (de)quantisation routines ARE pinned against the in-container `gguf` 0.19.0
 * python package (gguf/quants.py:222-251,381-401,504-522,554-572), which restates
 * the same ggml formats; tests/golden/ holds vectors generated from it.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library.  The product (libb200gguf.so) never does.
 */
#ifndef GGML_REF_H
#define GGML_REF_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ggml type ids (same numbering as GGUF files; gguf/constants.py GGMLQuantizationType) */
enum { REF_F32 = 0, REF_F16 = 1, REF_Q4_0 = 2, REF_Q8_0 = 8, REF_Q4_K = 12, REF_Q6_K = 14, REF_Q8_K = 15 };

#define REF_QK 32
#define REF_QK_K 256

/* bytes for one row of k elements of the given type; 0 if unsupported / misaligned */
size_t ref_row_size(int type, int64_t k);

/* ---- block (de)quantisation -------------------------------------------------- */
void ref_dequantize_row(int type, const void *blocks, float *y, int64_t k);
void ref_quantize_row_q4_0(const float *x, void *y, int64_t k);
void ref_quantize_row_q8_0(const float *x, void *y, int64_t k);
void ref_quantize_row_q8_K(const float *x, void *y, int64_t k);
/* Synthetic-weight makers for K-quants.  NOT restatements of ggml's
 * quantize_row_q4_K_ref / q6_K_ref (those do an rmse search); they emit valid
 * blocks of the same format with plain max-abs scaling.  Used only to build
 * random-init test models. */
void ref_make_q4_K(const float *x, void *y, int64_t k);
void ref_make_q6_K(const float *x, void *y, int64_t k);

/* ---- dot products (weight row of `wtype` . activation row already quantised) -- */
float ref_vec_dot_q4_0_q8_0(int64_t n, const void *vx, const void *vy);
float ref_vec_dot_q8_0_q8_0(int64_t n, const void *vx, const void *vy);
float ref_vec_dot_q4_K_q8_K(int64_t n, const void *vx, const void *vy);
float ref_vec_dot_q6_K_q8_K(int64_t n, const void *vx, const void *vy);

/* y[r] = W[r,:] . x   for r < rows.  `faithful` selects the arithmetic:
 *   0  "ideal": weights dequantised to f32, f64 accumulate, activations not quantised;
 *   1  ggml's generic C path: x quantised to the weight type's vec_dot type, exact integer
 *      block dots, per-block float terms added in f32, block after block;
 *   2  "canonical": the same terms, added in f64 and rounded to f32 once, exp/tanh rounded
 *      from f64 -- order-free, so the CUDA kernels can (and must) match it bit for bit.  (The
 *      attention dots of mode 2 follow ggml's 32-lane SIMD f16 dot instead; see ggml_ref.c.)
 *   3  mode 1 with the Q4_0 / Q8_0 dots in the order of ggml's AVX2 kernels (d_w*d_a first, eight
 *      fma lanes, hsum_float_8): a second legitimate ggml build, used only to measure how far two
 *      such builds drift apart end to end (bench.py `ggml_build_noise`).
 * threads<=0 -> all cores. */
void ref_mul_mat_vec(int wtype, const void *W, int64_t rows, int64_t k,
                     const float *x, float *y, int faithful, int threads);

/* ---- element ops --------------------------------------------------------------- */
void ref_rms_norm(const float *x, const float *g, float *y, int64_t n, float eps);
void ref_layer_norm(const float *x, const float *g, const float *b, float *y, int64_t n, float eps);
/* mode 0: adjacent pairs (llama "norm" rope); mode 2: NeoX halves. in place on
 * n_head vectors of head_dim; rotates the first n_rot dims. */
void ref_rope(float *x, int n_head, int head_dim, int n_rot, int pos, float freq_base, int mode);
/* the same with ggml's frequency factors and linear position scaling: theta = freq_scale * (theta_base / factor[pair]) */
void ref_rope_ext(float *x, int n_head, int head_dim, int n_rot, int pos, float freq_base, int mode, float freq_scale, const float *factors);
void ref_softmax(float *x, int64_t n);
/* canonical-mode attention of one token (see ggml_ref.c); k, v: [n][n_head_kv*hd] f16 */
void ref_attention_canon(const float *q, const uint16_t *k, const uint16_t *v, int n, int n_head, int n_head_kv, int hd, float scale, float *out);

/* ---- whole-model forward ------------------------------------------------------- */
enum { REF_ARCH_LLAMA = 0, REF_ARCH_PHI2 = 1 };

typedef struct {
    int type;          /* ggml type id */
    const void *data;  /* row-major [rows][row_size(type,k)] */
} ref_tensor;

typedef struct {
    ref_tensor attn_norm, attn_norm_b;
    ref_tensor wq, wk, wv, wo;         /* llama: separate; phi2: wq holds fused qkv */
    ref_tensor bq, bo;                 /* phi2 biases (bq = fused qkv bias) */
    ref_tensor ffn_norm;
    ref_tensor w_gate, w_up, w_down;   /* phi2: w_up / w_down only */
    ref_tensor b_up, b_down;
} ref_layer;

typedef struct {
    int arch;
    int n_layer, n_embd, n_head, n_head_kv, head_dim, n_ff, n_vocab, n_ctx, n_rot;
    float norm_eps, rope_freq_base;
    ref_tensor tok_embd, out_norm, out_norm_b, output, output_b;
    const ref_layer *layers;
    float rope_freq_scale;        /* 1 / rope.scaling.factor for linear scaling, else 1 (0 is read as 1) */
    const float *rope_factors;    /* rope_freqs.weight [n_rot/2] or NULL: theta is divided by it per pair */
} ref_model;

typedef struct ref_ctx ref_ctx;   /* KV cache (f16 when faithful, f32 when ideal) + scratch */

ref_ctx *ref_ctx_new(const ref_model *m, int faithful, int threads);
void ref_ctx_free(ref_ctx *c);
/* Overwrite KV rows [0,n_tok) of layer `il` with f16 data laid out [n_tok][n_head_kv*head_dim]. */
void ref_ctx_set_kv_f16(ref_ctx *c, int il, const uint16_t *k, const uint16_t *v, int n_tok);
/* Evaluate one token at position pos (appends to the KV cache at pos).
 * logits may be NULL (skips the head).  Returns 0 on success. */
int ref_eval(ref_ctx *c, int32_t token, int pos, float *logits);
/* hidden state after the last evaluated token's final norm (n_embd floats) */
const float *ref_last_hidden(const ref_ctx *c);

int ref_num_threads(void);
void ref_set_num_threads(int n);     /* overrides OMP_NUM_THREADS for later calls with threads == 0 */

#ifdef __cplusplus
}
#endif
#endif

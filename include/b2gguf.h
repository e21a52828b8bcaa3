/*
 * b2gguf.h -- C ABI of libb200gguf.so, the B200-native GGUF decode engine.
 *
 * What this replaces.  LittleSound/ollama-operator has no FFI for its hot path: every
 * `Model` CR becomes a pod that runs the `ollama/ollama` image with `serve`
 * (reference pkg/model/pod.go:14-66, wired in pkg/model/model.go:85-90 and
 * pkg/model/image_store.go:169), and the GGUF decode loop runs inside that container.
 * The replacement server (cmd/server in the reference's language, Go) binds the entry
 * points below through cgo -- see INTEGRATION.md for the binding -- and keeps the
 * container contract of pod.go (argv `serve`, OLLAMA_HOST, port 11434, /api/tags).
 * Each entry point names the upstream step of the container's decode loop it stands in
 * for.
 *
 * Conventions: plain C types only; every function returns 0 on success or a negative
 * b2_status, with a human-readable message available from b2_strerror() on the calling
 * thread; the caller owns every host buffer it passes; the library owns all device memory;
 * calls are blocking and synchronise their own stream before host-visible results are
 * returned; one OS thread drives a given b2_model at a time.  There is no CPU fallback:
 * without a CUDA device b2_init fails.
 */
#ifndef B2GGUF_H
#define B2GGUF_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    B2_OK = 0,
    B2_ERR_ARG = -1,          /* invalid argument */
    B2_ERR_CUDA = -2,         /* CUDA runtime error (message has file:line) */
    B2_ERR_IO = -3,           /* cannot open / map the GGUF file */
    B2_ERR_FORMAT = -4,       /* not a GGUF v2/v3 file, or missing tensors / keys */
    B2_ERR_UNSUPPORTED = -5,  /* architecture, tensor type or shape not supported */
    B2_ERR_NOMEM = -6,        /* device memory or KV pages exhausted */
    B2_ERR_STATE = -7         /* call not valid in the current state (e.g. bad sequence id) */
} b2_status;

/* ggml tensor type ids, as stored in GGUF files */
enum { B2_TYPE_F32 = 0, B2_TYPE_F16 = 1, B2_TYPE_Q4_0 = 2, B2_TYPE_Q8_0 = 8, B2_TYPE_Q4_K = 12, B2_TYPE_Q6_K = 14 };

typedef struct b2_ctx b2_ctx;       /* one per process: the devices this engine may use */
typedef struct b2_model b2_model;   /* weights resident in HBM + KV cache + sequence slots */

typedef struct {
    int32_t max_batch;   /* sequence slots (0 -> 1) */
    int32_t max_ctx;     /* tokens per sequence (0 -> the model's context length) */
    int32_t tp_size;     /* tensor-parallel degree (0/1 = one GPU); needs that many devices in b2_init with peer access;
                            llama graph only; results are bit-identical to tp_size 1 */
    int32_t flags;       /* B2_LOAD_* */
} b2_load_opts;
enum { B2_LOAD_NO_PDL = 1,      /* do not use programmatic dependent launch */
       B2_LOAD_NO_GRAPH = 2,    /* launch kernels one by one instead of replaying CUDA graphs */
       B2_LOAD_NO_MEGA = 4,     /* one kernel per matrix instead of the persistent per-step kernel */
       B2_LOAD_NO_BATCH = 8 };  /* evaluate prompts and multi-sequence decode calls one token at a time */

typedef struct {
    int32_t arch;        /* 0 llama, 1 phi2 */
    int32_t n_layer, n_embd, n_head, n_head_kv, head_dim, n_ff, n_vocab, n_ctx, n_rot;
    int32_t max_batch, max_ctx;
    float norm_eps, rope_freq_base;
    uint64_t weight_bytes;       /* quantised weight bytes resident in HBM */
    uint64_t step_weight_bytes;  /* weight bytes one decode step streams (all layers + head) */
    uint64_t kv_bytes_per_token; /* K+V bytes one cached token occupies over all layers */
} b2_model_info;

typedef struct {
    uint64_t steps;          /* decode/prefill steps executed since load */
    uint64_t kernel_launches;/* CUDA kernels launched by those steps */
    double   gpu_ms;         /* CUDA-event time of the most recent b2_prefill / b2_decode* call */
    uint64_t last_steps;     /* steps in that call */
    uint64_t last_launches;  /* kernel launches in that call */
} b2_stats;

/* Stands in for `ollama serve` start-up inside the container (pod.go:18-20): binds the
 * CUDA devices.  dev_ids may be NULL (device 0..n_dev-1). */
int b2_init(int n_dev, const int* dev_ids, b2_ctx** out);
void b2_shutdown(b2_ctx* ctx);
/* Message of the last failing call on this thread ("" if none). */
const char* b2_strerror(void);

/* Stands in for the container's lazy model load (GGUF blob under /root/.ollama, the mount of
 * pod.go:34-40): parses the file, uploads and repacks the quantised tensors, sizes the KV cache. */
int b2_model_load(b2_ctx* ctx, const char* gguf_path, const b2_load_opts* opts, b2_model** out);
void b2_model_free(b2_model* m);
int b2_model_get_info(const b2_model* m, b2_model_info* info);

/* Sequence slots (one per concurrent request the server batches). */
int b2_seq_new(b2_model* m, int* seq_out);
int b2_seq_free(b2_model* m, int seq);
int b2_seq_len(const b2_model* m, int seq);            /* tokens in the KV cache, or < 0 */
int b2_seq_truncate(b2_model* m, int seq, int n_keep); /* drop cached tokens beyond n_keep */

/* Tooling hook: declare the first n_tokens cache rows of the sequence valid without evaluating a prompt
 * (their K/V are whatever the cache holds: zeros after load).  Lets a benchmark reach a long context
 * without paying for the prefill; not for serving. */
int b2_seq_assume(b2_model* m, int seq, int n_tokens);

/* llama_decode over a prompt (upstream "prefill"): appends toks[0..n) to the sequence.
 * logits (n_vocab floats, may be NULL) receives the logits after the last token. */
int b2_prefill(b2_model* m, int seq, const int32_t* toks, int n, float* logits);

/* One decode step for n_seq sequences: evaluates last_tok[i] at the end of seqs[i] and writes
 * the greedy (arg-max) next token to next_tok[i]; logits (n_seq * n_vocab, may be NULL)
 * receives the full logits so the caller can sample otherwise.  Host buffers; the token ids
 * cross the PCIe bus inside this call. */
int b2_decode(b2_model* m, int n_seq, const int* seqs, const int32_t* last_tok, int32_t* next_tok, float* logits);

/* Device-resident greedy loop: n_steps decode steps for one sequence with the token feedback
 * kept on the GPU (CUDA-graph replay, no host round trip per token).  first_tok < 0 continues
 * from the arg-max of the previous call.  out_toks receives the n_steps generated tokens. */
int b2_decode_greedy(b2_model* m, int seq, int32_t first_tok, int n_steps, int32_t* out_toks);

int b2_get_stats(const b2_model* m, b2_stats* out);
/* One decode step through the persistent per-step kernel with time stamps: out receives
 * [n_cta][n_stage][12] GPU-global nanosecond stamps {stage entered, grid barrier passed, prologue
 * done, stage done, then eight phase stamps of the attention stage} (stages: per layer qkv,
 * attention, wo, gate|up, down).  Tuning aid. */
int b2_mega_trace(b2_model* m, int seq, int32_t tok, uint64_t* out, int cap, int* n_cta, int* n_stage);
/* One decode step for `seq` (like b2_decode with one sequence) run with a CUDA event after every
 * kernel: class_ms[9] / class_launches[9] receive device milliseconds and launch counts per
 * kernel class {0 step-begin, 1 qkv gemv, 2 attention, 3 wo gemv, 4 gate/up gemv, 5 down gemv,
 * 6 head gemv, 7 step-end, 8 persistent all-layers kernel (then classes 1-5 stay empty)}.
 * Stands in for llama.cpp's per-op timings (GGML_PERF) [upstream]. */
int b2_profile_step(b2_model* m, int seq, int32_t tok, float* class_ms, int* class_launches);

/* ---- test / tooling hooks (host buffers in, host buffers out) ------------------------------ */
/* ggml dequantize_row_<type>: n_elem values from file-format blocks. */
int b2_dequantize(int ggml_type, const void* blocks, size_t n_elem, float* out);
/* Same, but through the load-time repack into split planes (proves repack o dequant == dequant). */
int b2_dequantize_repacked(int ggml_type, const void* blocks, int64_t rows, int64_t k, float* out);
/* ggml mul_mat with one activation column: y[rows] = W[rows][k] . x[k] with ggml's W4A8
 * semantics.  gain (k floats, may be NULL) applies the fused RMSNorm prologue first:
 * x <- rms_norm(x, eps) * gain. */
int b2_gemv(int ggml_type, const void* W, int64_t rows, int64_t k, const float* x, const float* gain, float eps, float* y);
/* b2_gemv without any device allocation -- no counterpart in the reference (nor in b2_gemv's
 * upstream analogue): it exists for callers that cannot call cudaMalloc, e.g. a server whose KV
 * cache already fills the GPU, code living inside another framework's memory pool, or a path
 * that cannot afford cudaMalloc's device-wide synchronisation.  All device staging lives in a
 * static array of B2_NOALLOC_SCRATCH_BYTES inside the library (held on every GPU that loads it,
 * until exit).  Same arguments and results as b2_gemv; matrices whose staging does not fit
 * (roughly rows*k > 6.5 M weights for Q4_0) are rejected with B2_ERR_NOMEM.  The first call on a
 * device resolves the array's address and should precede any stream capture.  Calls share the
 * array: do not let two of them overlap on the GPU. */
#define B2_NOALLOC_SCRATCH_BYTES (8u << 20)
int b2_gemv_noalloc(int ggml_type, const void* W, int64_t rows, int64_t k, const float* x, const float* gain, float eps, float* y);
/* ggml mul_mat with n_cols (1..32) activation columns, x [n_cols][k], y [n_cols][rows]: the batched
 * kernel behind multi-sequence b2_decode and chunked b2_prefill (Q4_0 weights).  Every column equals
 * b2_gemv on that column bit for bit.  The kernel runs `iters` more times for *ms_out (may be NULL). */
int b2_gemm(int ggml_type, const void* W, int64_t rows, int64_t k, const float* x, int n_cols, const float* gain, float eps, float* y, int iters,
            float* ms_out);
/* The same product on the tcgen05 tensor cores (integer MMA per 32-weight block, accumulators in tensor memory),
 * n_cols 1..128: the kernel behind chunked b2_prefill and batched b2_decode for Q4_0 matrices whose k is a multiple of
 * 256.  acc32 = 0: block terms added in f64 (every column equals b2_gemv bit for bit); acc32 = 1: f32, block after
 * block, the order of ggml's generic C kernel (ggml-quants.c ggml_vec_dot_q4_0_q8_0, scalar branch [upstream]). */
int b2_gemm_tc(int ggml_type, const void* W, int64_t rows, int64_t k, const float* x, int n_cols, const float* gain, float eps, int acc32, float* y, int iters,
               float* ms_out);
/* Same kernel launched `iters` times back to back on resident data; returns the mean device
 * time per launch in *ms_out (CUDA events).  tune = {nb, g, rpg, ns, ctas_per_sm}, zeros for
 * the built-in heuristic. */
int b2_gemv_bench(int ggml_type, int64_t rows, int64_t k, int iters, const int* tune, float* ms_out);
/* Decode attention for one token: q[n_head*head_dim] f32, K/V [n_ctx][n_head_kv*head_dim]
 * f16 (row-major, unpaged on the host side; the hook pages them) -> out[n_head*head_dim]. */
int b2_attn_decode(const float* q, const uint16_t* k, const uint16_t* v, int n_ctx, int n_head, int n_head_kv,
                   int head_dim, float scale, int cluster, float* out);
/* The attention kernel launched `iters` times back to back over `n_layers` distinct caches of n_ctx
 * tokens (cluster = 0 for the built-in choice); mean device milliseconds per launch in *ms_out. */
int b2_attn_bench(int n_ctx, int n_head, int n_head_kv, int head_dim, int cluster, int n_layers, int iters, float* ms_out);
/* Read back one per-sequence activation buffer as it stands after the last step:
 * what = 0 residual stream x[n_embd], 1 rotated query q[n_head*head_dim], 2 attention output,
 * 3 FFN hidden h[n_ff], 4 logits[n_vocab], 5 phi-2 attention projection[n_embd]. */
int b2_debug_read(b2_model* m, int seq, int what, float* out, int n);
/* Copy cached K and V rows [0, n_tok) of one layer to the host as f16 [n_tok][n_head_kv*head_dim]. */
int b2_kv_read(b2_model* m, int seq, int layer, uint16_t* k_out, uint16_t* v_out, int n_tok);

/* ---- model store and tokenizer (host only, no GPU needed) ----------------------------------------------
 * What sits either side of the decode path inside the pod the operator builds: the read-only volume at
 * /root/.ollama that `ollama pull` filled (reference pkg/model/pod.go:34-40 mount, :68-83 puller,
 * pkg/model/image_store.go:172-181 store), and the tokenizer every GGUF blob carries in its metadata. */
typedef struct b2_store_entry {
    char name[256];             /* canonical [namespace/]model:tag */
    char model_path[1024];      /* blob of the application/vnd.ollama.image.model layer: the GGUF file */
    char template_path[1024];   /* prompt template layer ("" if the manifest has none) */
    char system_path[1024];     /* system prompt layer */
    char params_path[1024];     /* JSON parameters layer (stop strings, temperature, ...) */
    char digest[80];            /* sha256:<hex> of the model layer */
    uint64_t model_size;
    int64_t modified_unix;
} b2_store_entry;
/* models: the store root (a directory with manifests/ and blobs/, or an OLLAMA home whose models/ child has them);
 * NULL: $OLLAMA_MODELS, $HOME/.ollama/models, /root/.ollama/models.  name: [registry/][namespace/]model[:tag]. */
int b2_store_resolve(const char* models, const char* name, b2_store_entry* out);
/* One line per pulled model, "name\tbytes\tdigest\tmtime\n"; returns the bytes the full listing needs. */
int b2_store_list(const char* models, char* buf, int cap);

typedef struct b2_tokenizer b2_tokenizer;
typedef struct b2_tokenizer_info {
    int32_t kind;               /* 1 = SentencePiece-style "llama", 2 = byte-level BPE "gpt2" */
    int32_t n_vocab, bos, eos, add_bos;
} b2_tokenizer_info;
int b2_tokenizer_load(const char* gguf_path, b2_tokenizer** out);
void b2_tokenizer_free(b2_tokenizer* t);
int b2_tokenizer_get_info(const b2_tokenizer* t, b2_tokenizer_info* out);
/* Returns the token count; if it exceeds cap nothing is written (call again with room).  add_bos < 0: the file's
 * tokenizer.ggml.add_bos_token default.  parse_special: control tokens spelled out in the text become their ids. */
int b2_tokenize(const b2_tokenizer* t, const char* text, int add_bos, int parse_special, int32_t* out, int cap);
/* Returns the byte count (terminator not included); beyond cap - 1 nothing is written. */
int b2_detokenize(const b2_tokenizer* t, const int32_t* toks, int n, int render_special, char* out, int cap);
int b2_token_piece(const b2_tokenizer* t, int32_t tok, int render_special, char* out, int cap);

#ifdef __cplusplus
}
#endif
#endif /* B2GGUF_H */

// tokenizer_api.cpp -- C ABI over tokenizer.cpp (include/b2gguf.h: b2_tokenizer_*).  Host only: no GPU is touched, so a
// server can tokenise (and answer its probes) before, or without, a device.
#include <string.h>

#include <memory>

#include "../../include/b2gguf.h"
#include "common.cuh"
#include "gguf.h"
#include "tokenizer.h"

struct b2_tokenizer {
    b2::Tokenizer tok;
};

extern "C" {

int b2_tokenizer_load(const char* gguf_path, b2_tokenizer** out) {
    if (!gguf_path || !out) B2_FAIL(B2_ERR_ARG, "b2_tokenizer_load: null argument");
    *out = nullptr;
    b2::GgufFile f;
    B2_TRY(f.open(gguf_path));
    std::unique_ptr<b2_tokenizer> t(new b2_tokenizer());
    B2_TRY(t->tok.load(f));
    *out = t.release();
    return B2_OK;
}
void b2_tokenizer_free(b2_tokenizer* t) { delete t; }

int b2_tokenizer_get_info(const b2_tokenizer* t, b2_tokenizer_info* out) {
    if (!t || !out) B2_FAIL(B2_ERR_ARG, "b2_tokenizer_get_info: null argument");
    out->kind = (int)t->tok.kind(); out->n_vocab = t->tok.n_vocab(); out->bos = t->tok.bos(); out->eos = t->tok.eos();
    out->add_bos = t->tok.add_bos_default() ? 1 : 0;
    return B2_OK;
}

int b2_tokenize(const b2_tokenizer* t, const char* text, int add_bos, int parse_special, int32_t* out, int cap) {
    if (!t || !text || cap < 0 || (cap > 0 && !out)) B2_FAIL(B2_ERR_ARG, "b2_tokenize: bad arguments");
    const std::vector<int32_t> ids = t->tok.encode(text, add_bos < 0 ? t->tok.add_bos_default() : add_bos != 0, parse_special != 0);
    const int n = (int)ids.size();
    if (n <= cap) memcpy(out, ids.data(), (size_t)n * 4);
    return n;                   // more than cap: nothing was written, call again with room for n
}

int b2_detokenize(const b2_tokenizer* t, const int32_t* toks, int n, int render_special, char* out, int cap) {
    if (!t || n < 0 || (n > 0 && !toks) || cap < 0 || (cap > 0 && !out)) B2_FAIL(B2_ERR_ARG, "b2_detokenize: bad arguments");
    const std::string s = t->tok.decode(toks, n, render_special != 0);
    if ((int)s.size() + 1 <= cap) { memcpy(out, s.data(), s.size()); out[s.size()] = 0; }
    return (int)s.size();       // bytes (without the terminator); beyond cap - 1: nothing was written
}

int b2_token_piece(const b2_tokenizer* t, int32_t tok, int render_special, char* out, int cap) {
    if (!t || cap < 0 || (cap > 0 && !out)) B2_FAIL(B2_ERR_ARG, "b2_token_piece: bad arguments");
    const std::string s = t->tok.piece(tok, render_special != 0);
    if ((int)s.size() + 1 <= cap) { memcpy(out, s.data(), s.size()); out[s.size()] = 0; }
    return (int)s.size();
}

}  // extern "C"

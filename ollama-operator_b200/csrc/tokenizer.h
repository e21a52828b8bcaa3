// tokenizer.h -- the tokenizer a GGUF file carries in its metadata (`tokenizer.ggml.*`), as llama.cpp reads it
// [upstream llama-vocab.cpp]: "llama" = SentencePiece-style BPE over scored pieces with byte fallback (Llama-2,
// Mistral), "gpt2" = byte-level BPE over ranked merges (phi-2).  Host only; what a server needs between the text of
// /api/generate and the token ids of b2_prefill / b2_decode.
#pragma once
#include <stdint.h>

#include <string>
#include <unordered_map>
#include <vector>

#include "gguf.h"

namespace b2 {

class Tokenizer {
  public:
    enum Kind { NONE = 0, SPM = 1, BPE = 2 };
    // token types of tokenizer.ggml.token_type (gguf TokenType)
    enum Type { NORMAL = 1, UNKNOWN = 2, CONTROL = 3, USER_DEFINED = 4, UNUSED = 5, BYTE = 6 };

    int load(const GgufFile& f);          // B2_OK, or B2_ERR_FORMAT / B2_ERR_UNSUPPORTED with the message set
    Kind kind() const { return kind_; }
    int n_vocab() const { return (int)tokens_.size(); }
    int bos() const { return bos_; }
    int eos() const { return eos_; }
    bool add_bos_default() const { return add_bos_; }

    // text -> ids.  parse_special: control / user-defined tokens written out in the text ("<s>", "<|im_start|>")
    // become their ids instead of being tokenised as characters.
    std::vector<int32_t> encode(const std::string& text, bool add_bos, bool parse_special) const;
    // ids -> bytes of text (control tokens render as nothing unless render_special)
    std::string decode(const int32_t* toks, int n, bool render_special = false) const;
    std::string piece(int32_t tok, bool render_special = false) const;
    bool is_control(int32_t tok) const { return tok >= 0 && tok < n_vocab() && types_[tok] == CONTROL; }

  private:
    void encode_spm(const std::string& text, std::vector<int32_t>& out) const;
    void encode_bpe(const std::string& text, std::vector<int32_t>& out) const;
    void bpe_word(const std::string& word, std::vector<int32_t>& out) const;

    Kind kind_ = NONE;
    std::vector<std::string> tokens_;
    std::vector<float> scores_;
    std::vector<int32_t> types_;
    std::unordered_map<std::string, int32_t> index_;
    std::unordered_map<std::string, int32_t> merge_rank_;      // "left right" -> rank (BPE)
    std::vector<std::pair<std::string, int32_t>> specials_;     // control / user-defined tokens, longest first
    int32_t byte_token_[256];                                   // SPM byte fallback: <0xXX>
    int bos_ = -1, eos_ = -1, unk_ = -1;
    bool add_bos_ = true, add_space_prefix_ = true;
};

}  // namespace b2

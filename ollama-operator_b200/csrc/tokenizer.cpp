// tokenizer.cpp -- SentencePiece-style ("llama") and byte-level BPE ("gpt2") tokenizers from GGUF metadata.
// Algorithms as llama.cpp applies them to these vocabularies [upstream llama-vocab.cpp: llm_tokenizer_spm,
// llm_tokenizer_bpe]; the tests compare with `sentencepiece` and `tokenizers` on vocabularies trained in the test.
#include "tokenizer.h"

#include <string.h>

#include <algorithm>
#include <queue>

#include "common.cuh"

namespace b2 {

namespace {

inline int utf8_len(unsigned char c) { return c < 0x80 ? 1 : (c >> 5) == 6 ? 2 : (c >> 4) == 14 ? 3 : (c >> 3) == 30 ? 4 : 1; }

// GPT-2's reversible byte <-> unicode mapping: printable bytes map to themselves, the rest to U+0100...
struct ByteMap {
    std::string enc[256];                                  // byte -> UTF-8 of its stand-in code point
    std::unordered_map<std::string, unsigned char> dec;
    ByteMap() {
        int n = 0;
        for (int b = 0; b < 256; b++) {
            const bool keep = (b >= 33 && b <= 126) || (b >= 161 && b <= 172) || (b >= 174 && b <= 255);
            const int cp = keep ? b : 256 + n++;
            std::string s;
            if (cp < 0x80) s.push_back((char)cp);
            else { s.push_back((char)(0xC0 | (cp >> 6))); s.push_back((char)(0x80 | (cp & 0x3F))); }
            enc[b] = s;
            dec[s] = (unsigned char)b;
        }
    }
};
const ByteMap& byte_map() { static const ByteMap m; return m; }

inline bool is_letter(unsigned char c) { return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || c >= 0x80; }      // non-ASCII bytes count as letters
inline bool is_digit(unsigned char c) { return c >= '0' && c <= '9'; }
inline bool is_space(unsigned char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\v' || c == '\f'; }

// GPT-2's pre-tokeniser  's|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+   on bytes
void gpt2_split(const std::string& t, std::vector<std::string>& out) {
    const size_t n = t.size();
    size_t i = 0;
    while (i < n) {
        const unsigned char c = (unsigned char)t[i];
        if (c == '\'' && i + 1 < n) {
            static const char* suf[] = {"s", "t", "re", "ve", "m", "ll", "d"};
            bool hit = false;
            for (const char* s : suf) {
                const size_t l = strlen(s);
                if (t.compare(i + 1, l, s) == 0) { out.push_back(t.substr(i, l + 1)); i += l + 1; hit = true; break; }
            }
            if (hit) continue;
        }
        size_t j = i;
        const bool lead = c == ' ' && i + 1 < n && !is_space((unsigned char)t[i + 1]);
        if (lead) j++;
        const unsigned char d = (unsigned char)t[j];
        if (j < n && is_letter(d)) {
            while (j < n && is_letter((unsigned char)t[j])) j++;
        } else if (j < n && is_digit(d)) {
            while (j < n && is_digit((unsigned char)t[j])) j++;
        } else if (j < n && !is_space(d)) {
            while (j < n && !is_space((unsigned char)t[j]) && !is_letter((unsigned char)t[j]) && !is_digit((unsigned char)t[j])) j++;
        } else {
            // white space: all of it if the text ends there, otherwise all but the last character (which leads the next word)
            j = i;
            while (j < n && is_space((unsigned char)t[j])) j++;
            if (j < n && j - i > 1) j--;
        }
        out.push_back(t.substr(i, j - i));
        i = j;
    }
}

}  // namespace

int Tokenizer::load(const GgufFile& f) {
    std::string model;
    if (!f.get_str("tokenizer.ggml.model", &model)) B2_FAIL(B2_ERR_FORMAT, "tokenizer: the file carries no tokenizer.ggml.model");
    if (model == "llama") kind_ = SPM;
    else if (model == "gpt2") kind_ = BPE;
    else B2_FAIL(B2_ERR_UNSUPPORTED, "tokenizer: model '%s' is not supported (llama, gpt2)", model.c_str());
    if (!f.get_str_array("tokenizer.ggml.tokens", &tokens_) || tokens_.empty()) B2_FAIL(B2_ERR_FORMAT, "tokenizer: tokenizer.ggml.tokens missing");
    if (!f.get_f32_array("tokenizer.ggml.scores", &scores_)) scores_.assign(tokens_.size(), 0.f);
    if (!f.get_i32_array("tokenizer.ggml.token_type", &types_)) types_.assign(tokens_.size(), NORMAL);
    if (scores_.size() != tokens_.size() || types_.size() != tokens_.size()) B2_FAIL(B2_ERR_FORMAT, "tokenizer: scores / token types do not match the vocabulary size");
    index_.clear();
    index_.reserve(tokens_.size() * 2);
    for (size_t i = 0; i < tokens_.size(); i++) index_.emplace(tokens_[i], (int32_t)i);      // first id wins for duplicate pieces
    for (int b = 0; b < 256; b++) {
        char name[8];
        snprintf(name, sizeof name, "<0x%02X>", b);
        auto it = index_.find(name);
        byte_token_[b] = it == index_.end() ? -1 : it->second;
    }
    if (kind_ == BPE) {
        std::vector<std::string> merges;
        if (!f.get_str_array("tokenizer.ggml.merges", &merges)) B2_FAIL(B2_ERR_FORMAT, "tokenizer: tokenizer.ggml.merges missing (gpt2)");
        merge_rank_.reserve(merges.size() * 2);
        for (size_t i = 0; i < merges.size(); i++) merge_rank_.emplace(merges[i], (int32_t)i);
    }
    uint64_t v;
    bos_ = f.get_u64("tokenizer.ggml.bos_token_id", &v) ? (int)v : (kind_ == SPM ? 1 : -1);
    eos_ = f.get_u64("tokenizer.ggml.eos_token_id", &v) ? (int)v : (kind_ == SPM ? 2 : -1);
    unk_ = f.get_u64("tokenizer.ggml.unknown_token_id", &v) ? (int)v : (kind_ == SPM ? 0 : -1);
    add_bos_ = f.get_u64("tokenizer.ggml.add_bos_token", &v) ? v != 0 : kind_ == SPM;
    add_space_prefix_ = f.get_u64("tokenizer.ggml.add_space_prefix", &v) ? v != 0 : true;
    specials_.clear();
    for (size_t i = 0; i < tokens_.size(); i++)
        if ((types_[i] == CONTROL || types_[i] == USER_DEFINED) && !tokens_[i].empty()) specials_.emplace_back(tokens_[i], (int32_t)i);
    std::sort(specials_.begin(), specials_.end(), [](const auto& a, const auto& b) { return a.first.size() > b.first.size(); });
    return B2_OK;
}

// ---- SentencePiece-style: merge the adjacent pair whose concatenation is the best-scored piece, until none is ----
void Tokenizer::encode_spm(const std::string& raw, std::vector<int32_t>& out) const {
    std::string text;
    text.reserve(raw.size() + 8);
    for (char c : raw) {
        if (c == ' ') text += "\xE2\x96\x81";      // U+2581
        else text.push_back(c);
    }
    struct Sym { int prev, next; size_t off, len; };
    std::vector<Sym> sym;
    for (size_t off = 0; off < text.size();) {
        const size_t l = std::min((size_t)utf8_len((unsigned char)text[off]), text.size() - off);
        sym.push_back({(int)sym.size() - 1, (int)sym.size() + 1, off, l});
        off += l;
    }
    if (sym.empty()) return;
    sym.back().next = -1;
    struct Bigram { float score; int left, right; size_t len; };
    auto worse = [](const Bigram& a, const Bigram& b) { return a.score < b.score || (a.score == b.score && a.left > b.left); };
    std::priority_queue<Bigram, std::vector<Bigram>, decltype(worse)> q(worse);
    auto try_add = [&](int l, int r) {
        if (l < 0 || r < 0) return;
        const std::string piece = text.substr(sym[l].off, sym[l].len + sym[r].len);
        auto it = index_.find(piece);
        if (it == index_.end()) return;
        q.push({scores_[it->second], l, r, piece.size()});
    };
    for (int i = 1; i < (int)sym.size(); i++) try_add(i - 1, i);
    while (!q.empty()) {
        const Bigram b = q.top();
        q.pop();
        Sym &L = sym[b.left], &R = sym[b.right];
        if (L.len == 0 || R.len == 0 || L.len + R.len != b.len || L.next != b.right) continue;      // one side was merged away meanwhile
        L.len += R.len;
        R.len = 0;
        L.next = R.next;
        if (R.next >= 0) sym[R.next].prev = b.left;
        try_add(L.prev, b.left);
        try_add(b.left, L.next);
    }
    for (int i = 0; i >= 0; i = sym[i].next) {
        const std::string piece = text.substr(sym[i].off, sym[i].len);
        auto it = index_.find(piece);
        if (it != index_.end()) { out.push_back(it->second); continue; }
        for (unsigned char c : piece) {                 // byte fallback
            if (byte_token_[c] >= 0) out.push_back(byte_token_[c]);
            else if (unk_ >= 0) out.push_back(unk_);
        }
    }
}

// ---- byte-level BPE: inside each pre-token, merge the adjacent pair of lowest rank until none has a rank ----
void Tokenizer::bpe_word(const std::string& word, std::vector<int32_t>& out) const {
    const ByteMap& bm = byte_map();
    std::vector<std::string> parts;
    for (unsigned char c : word) parts.push_back(bm.enc[c]);
    while (parts.size() > 1) {
        int best = -1, best_rank = INT32_MAX;
        for (size_t i = 0; i + 1 < parts.size(); i++) {
            auto it = merge_rank_.find(parts[i] + " " + parts[i + 1]);
            if (it != merge_rank_.end() && it->second < best_rank) { best_rank = it->second; best = (int)i; }
        }
        if (best < 0) break;
        // merge every occurrence of that pair, left to right
        const std::string a = parts[best], b = parts[best + 1];
        std::vector<std::string> next;
        for (size_t i = 0; i < parts.size();) {
            if (i + 1 < parts.size() && parts[i] == a && parts[i + 1] == b) { next.push_back(a + b); i += 2; }
            else next.push_back(parts[i++]);
        }
        parts.swap(next);
    }
    for (const std::string& p : parts) {
        auto it = index_.find(p);
        if (it != index_.end()) { out.push_back(it->second); continue; }
        for (size_t off = 0; off < p.size();) {        // an unmerged stand-in character that is not a piece: its own byte pieces
            const size_t l = std::min((size_t)utf8_len((unsigned char)p[off]), p.size() - off);
            auto jt = index_.find(p.substr(off, l));
            if (jt != index_.end()) out.push_back(jt->second);
            else if (unk_ >= 0) out.push_back(unk_);
            off += l;
        }
    }
}
void Tokenizer::encode_bpe(const std::string& text, std::vector<int32_t>& out) const {
    std::vector<std::string> words;
    gpt2_split(text, words);
    for (const std::string& w : words) bpe_word(w, out);
}

std::vector<int32_t> Tokenizer::encode(const std::string& text, bool add_bos, bool parse_special) const {
    std::vector<int32_t> out;
    if (add_bos && bos_ >= 0) out.push_back(bos_);
    // split the text at special tokens written out literally, tokenise the stretches between them
    size_t pos = 0;
    bool first = true;
    auto plain = [&](const std::string& s) {
        if (s.empty()) return;
        if (kind_ == SPM) encode_spm((first && add_space_prefix_ ? " " : "") + s, out);
        else encode_bpe(s, out);
        first = false;
    };
    if (!parse_special || specials_.empty()) { plain(text); return out; }
    while (pos < text.size()) {
        size_t hit = std::string::npos;
        int32_t id = -1;
        size_t hit_len = 0;
        for (const auto& sp : specials_) {
            const size_t at = text.find(sp.first, pos);
            if (at != std::string::npos && (at < hit || (at == hit && sp.first.size() > hit_len))) { hit = at; id = sp.second; hit_len = sp.first.size(); }
        }
        if (hit == std::string::npos) { plain(text.substr(pos)); break; }
        plain(text.substr(pos, hit - pos));
        out.push_back(id);
        first = false;
        pos = hit + hit_len;
    }
    return out;
}

std::string Tokenizer::piece(int32_t tok, bool render_special) const {
    if (tok < 0 || tok >= n_vocab()) return std::string();
    const int ty = types_[tok];
    if (ty == CONTROL || ty == UNKNOWN || ty == UNUSED) return render_special ? tokens_[tok] : std::string();
    const std::string& t = tokens_[tok];
    if (kind_ == SPM) {
        if (ty == BYTE && t.size() == 6) return std::string(1, (char)strtol(t.c_str() + 3, nullptr, 16));
        std::string s;
        for (size_t i = 0; i < t.size();) {
            if (t.compare(i, 3, "\xE2\x96\x81") == 0) { s.push_back(' '); i += 3; }
            else s.push_back(t[i++]);
        }
        return s;
    }
    if (ty == USER_DEFINED) return t;
    const ByteMap& bm = byte_map();
    std::string s;
    for (size_t off = 0; off < t.size();) {
        const size_t l = std::min((size_t)utf8_len((unsigned char)t[off]), t.size() - off);
        auto it = bm.dec.find(t.substr(off, l));
        if (it != bm.dec.end()) s.push_back((char)it->second);
        else s += t.substr(off, l);
        off += l;
    }
    return s;
}

std::string Tokenizer::decode(const int32_t* toks, int n, bool render_special) const {
    std::string s;
    for (int i = 0; i < n; i++) s += piece(toks[i], render_special);
    // the space that encode() put in front of the text (SentencePiece's dummy prefix) is not part of it
    if (kind_ == SPM && add_space_prefix_ && !s.empty() && s[0] == ' ') s.erase(0, 1);
    return s;
}

}  // namespace b2

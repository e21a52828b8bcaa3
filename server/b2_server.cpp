// b2_server.cpp -- the process the operator's pod template starts: `serve`, listening on $OLLAMA_HOST (default port
// 11434), answering the Ollama HTTP API from the read-only model store with the B200 engine behind the C ABI.
//
// Container contract (reference pkg/model/pod.go:14-66): argv `serve`; env OLLAMA_HOST=0.0.0.0; port 11434 named "ollama";
// volume /root/.ollama (read-only in model pods); readiness AND liveness probes are `GET /api/tags`, which therefore never
// touches the GPU or a lock a generation holds.  HTTP surface [upstream Ollama API; reference
// docs/pages/en/guide/getting-started.md:129-149 shows `ollama run` and the OpenAI-compatible call]:
//   GET|HEAD /            "Ollama is running"
//   GET  /api/tags        pulled models (b2_store_list)
//   GET  /api/version
//   POST /api/show        {name|model}
//   POST /api/generate    {model, prompt, raw, stream, context[], options{num_predict, temperature, top_k, top_p, seed, stop}} -> NDJSON
//   POST /api/chat        {model, messages[], stream, options}                                                          -> NDJSON
//   POST /v1/chat/completions  (OpenAI-compatible; stream = SSE)
//
// Threads: one per connection (parse, tokenise, stream), plus ONE engine thread that owns the b2_model (the C ABI allows
// one OS thread per model): it admits waiting requests into free sequence slots (b2_prefill), then advances every running
// sequence by one token per b2_decode call -- the batched step -- and retires sequences on EOS, a stop string or
// num_predict.  This file uses nothing of the library but include/b2gguf.h.
//
// The reference's host language is Go; no Go toolchain exists in this image, so this stand-in is C++.  The Go `cmd/server`
// a maintainer would ship binds the same entry points through cgo (server/go/).
#include <arpa/inet.h>
#include <errno.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <signal.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <fstream>
#include <map>
#include <memory>
#include <mutex>
#include <random>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include "../include/b2gguf.h"
#include "../ollama-operator_b200/csrc/json_min.h"

using b2::Json;
using b2::json_escape;
using Clock = std::chrono::steady_clock;

static const char* kVersion = "0.1.32-b200";

static void logf(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    fprintf(stderr, "[b2-server] ");
    vfprintf(stderr, fmt, ap);
    fputc('\n', stderr);
    va_end(ap);
}
static std::string now_rfc3339() {
    char buf[64];
    timespec ts;
    clock_gettime(CLOCK_REALTIME, &ts);
    tm t;
    gmtime_r(&ts.tv_sec, &t);
    const size_t n = strftime(buf, sizeof buf, "%Y-%m-%dT%H:%M:%S", &t);
    snprintf(buf + n, sizeof buf - n, ".%06ldZ", ts.tv_nsec / 1000);
    return buf;
}
static std::string read_text(const char* path) {
    if (!path || !*path) return std::string();
    std::ifstream f(path, std::ios::binary);
    std::ostringstream ss;
    ss << f.rdbuf();
    return ss.str();
}

// Byte-fallback tokens can split a UTF-8 character over several tokens (and a random-init model emits arbitrary bytes):
// take the longest prefix of `pending` that is complete, valid UTF-8, replacing invalid bytes with U+FFFD; an incomplete
// sequence at the end stays pending unless `flush`.
static std::string take_valid_utf8(std::string& pending, bool flush) {
    std::string out;
    size_t i = 0;
    const size_t n = pending.size();
    while (i < n) {
        const unsigned char c = (unsigned char)pending[i];
        const int len = c < 0x80 ? 1 : (c >> 5) == 6 && c >= 0xC2 ? 2 : (c >> 4) == 14 ? 3 : (c >> 3) == 30 && c <= 0xF4 ? 4 : 0;
        if (len == 0) { out += "\xEF\xBF\xBD"; i++; continue; }
        if (i + (size_t)len > n) {
            bool plausible = true;
            for (size_t k = i + 1; k < n; k++) plausible &= ((unsigned char)pending[k] & 0xC0) == 0x80;
            if (plausible && !flush) break;                  // wait for the rest of the character
            out += "\xEF\xBF\xBD"; i++; continue;
        }
        bool ok = true;
        for (int k = 1; k < len; k++) ok &= ((unsigned char)pending[i + k] & 0xC0) == 0x80;
        if (ok && len == 3) { const unsigned cp = ((c & 0x0F) << 12) | (((unsigned char)pending[i + 1] & 0x3F) << 6); ok = cp >= 0x800 && (cp < 0xD800 || cp > 0xDFFF); }
        if (ok && len == 4) { const unsigned cp = ((c & 0x07) << 18) | (((unsigned char)pending[i + 1] & 0x3F) << 12); ok = cp >= 0x10000 && cp <= 0x10FFFF; }
        if (!ok) { out += "\xEF\xBF\xBD"; i++; continue; }
        out.append(pending, i, (size_t)len);
        i += (size_t)len;
    }
    pending.erase(0, i);
    return out;
}

// ---- one generation request, shared between its connection thread and the engine thread -------------------------
struct Options {
    int num_predict = 128;
    float temperature = 0.8f, top_p = 0.9f;
    int top_k = 40;
    uint32_t seed = 0;
    bool seeded = false;
    std::vector<std::string> stop;
};
struct Request {
    std::vector<int32_t> prompt;           // token ids (the connection thread tokenised)
    Options opt;
    // engine -> connection
    std::mutex mu;
    std::condition_variable cv;
    std::deque<std::string> pieces;        // detokenised text, in order
    std::vector<int32_t> generated;
    bool done = false;
    std::string error, done_reason = "stop";
    double load_s = 0, prompt_eval_s = 0, eval_s = 0;
    std::atomic<bool> cancelled{false};    // the client went away
};

// ---- the engine: model residency + the admit / step / retire loop ----------------------------------------------
class Engine {
  public:
    Engine(const std::string& store, int max_batch, int max_ctx) : store_(store), max_batch_(max_batch), max_ctx_(max_ctx) { thread_ = std::thread([this] { run(); }); }
    ~Engine() {
        { std::lock_guard<std::mutex> l(mu_); stop_ = true; }
        cv_.notify_all();
        thread_.join();
    }
    // Make `name` the resident model (loads on the engine thread; one model at a time) and return its tokenizer.
    // Blocks until loaded.  Thread-safe.
    int ensure_model(const std::string& name, b2_tokenizer** tok, b2_store_entry* entry, std::string* err) {
        std::unique_lock<std::mutex> l(mu_);
        load_want_ = name;
        load_done_ = false;
        cv_.notify_all();
        cv_load_.wait(l, [&] { return load_done_ && load_name_ == name; });
        if (!load_err_.empty()) { *err = load_err_; return -1; }
        *tok = tok_;
        *entry = entry_;
        return 0;
    }
    void submit(const std::shared_ptr<Request>& r) {
        { std::lock_guard<std::mutex> l(mu_); waiting_.push_back(r); }
        cv_.notify_all();
    }
    double last_load_s() const { return last_load_s_; }

  private:
    struct Running { std::shared_ptr<Request> r; int seq; int32_t last; int n_gen; std::mt19937 rng; std::string tail; Clock::time_point t_eval; std::string utf8; };

    void fail(const std::shared_ptr<Request>& r, const std::string& e) {
        std::lock_guard<std::mutex> l(r->mu);
        r->error = e; r->done = true;
        r->cv.notify_all();
    }
    void unload() {
        if (model_) { b2_model_free(model_); model_ = nullptr; }
        if (tok_) { b2_tokenizer_free(tok_); tok_ = nullptr; }
    }
    void load(const std::string& name) {
        const auto t0 = Clock::now();
        load_err_.clear();
        if (name == load_name_ && model_) return;
        unload();
        load_name_ = name;
        if (b2_store_resolve(store_.c_str(), name.c_str(), &entry_) != B2_OK) { load_err_ = std::string("model '") + name + "' not found: " + b2_strerror(); return; }
        const int tp = std::max(1, std::min(8, getenv("B2_TP_SIZE") ? atoi(getenv("B2_TP_SIZE")) : 1));       // GPUs of this replica (spec.b200.tensorParallel)
        if (!ctx_) {
            int devs[8];
            for (int i = 0; i < tp; i++) devs[i] = i;
            if (b2_init(tp, devs, &ctx_) != B2_OK) { load_err_ = b2_strerror(); return; }
        }
        b2_load_opts o;
        memset(&o, 0, sizeof o);
        o.max_batch = max_batch_; o.max_ctx = max_ctx_; o.tp_size = tp;
        if (b2_model_load(ctx_, entry_.model_path, &o, &model_) != B2_OK) { load_err_ = b2_strerror(); model_ = nullptr; return; }
        if (b2_tokenizer_load(entry_.model_path, &tok_) != B2_OK) { load_err_ = b2_strerror(); unload(); return; }
        b2_model_get_info(model_, &info_);
        b2_tokenizer_get_info(tok_, &tinfo_);
        last_load_s_ = std::chrono::duration<double>(Clock::now() - t0).count();
        logf("loaded %s (%s, %d layers, ctx %d, %d slots) in %.2f s", entry_.name, entry_.model_path, info_.n_layer, info_.max_ctx, info_.max_batch, last_load_s_);
    }

    int32_t sample(Running& g, const float* logits) {
        const Options& o = g.r->opt;
        const int V = info_.n_vocab;
        if (o.temperature <= 0.f) return (int32_t)(std::max_element(logits, logits + V) - logits);
        std::vector<std::pair<float, int>> c(V);
        for (int i = 0; i < V; i++) c[i] = {logits[i], i};
        const int k = o.top_k > 0 && o.top_k < V ? o.top_k : V;
        std::partial_sort(c.begin(), c.begin() + k, c.end(), [](const auto& a, const auto& b) { return a.first > b.first; });
        c.resize(k);
        double sum = 0;
        std::vector<double> p(k);
        for (int i = 0; i < k; i++) { p[i] = exp((double)(c[i].first - c[0].first) / o.temperature); sum += p[i]; }
        int keep = k;
        if (o.top_p < 1.f) {
            double acc = 0;
            for (int i = 0; i < k; i++) { acc += p[i] / sum; if (acc >= o.top_p) { keep = i + 1; break; } }
        }
        double s2 = 0;
        for (int i = 0; i < keep; i++) s2 += p[i];
        double u = std::uniform_real_distribution<double>(0.0, s2)(g.rng);
        for (int i = 0; i < keep; i++) { u -= p[i]; if (u <= 0) return c[i].second; }
        return c[keep - 1].second;
    }

    // hand one token's text to the connection; returns true when the sequence must stop
    bool emit(Running& g, int32_t tok) {
        Request& r = *g.r;
        char buf[256];
        const int n = b2_token_piece(tok_, tok, 0, buf, sizeof buf);
        std::string piece = n > 0 && n < (int)sizeof buf ? std::string(buf, (size_t)n) : std::string();
        bool stop = tok == tinfo_.eos;
        if (!stop && !r.opt.stop.empty()) {          // stop strings may span tokens: hold back a tail as long as the longest of them
            g.tail += piece;
            for (const std::string& s : r.opt.stop) {
                const size_t at = g.tail.find(s);
                if (at != std::string::npos) { piece = g.tail.substr(0, at); g.tail.clear(); stop = true; goto flush; }
            }
            size_t hold = 0;
            for (const std::string& s : r.opt.stop) hold = std::max(hold, s.size() - 1);
            if (g.tail.size() > hold) { piece = g.tail.substr(0, g.tail.size() - hold); g.tail.erase(0, g.tail.size() - hold); }
            else piece.clear();
        }
    flush:
        if (tok == tinfo_.eos) { piece = g.tail; g.tail.clear(); }
        g.utf8 += piece;
        piece = take_valid_utf8(g.utf8, stop);
        {
            std::lock_guard<std::mutex> l(r.mu);
            if (tok != tinfo_.eos) r.generated.push_back(tok);
            if (!piece.empty()) r.pieces.push_back(piece);
        }
        r.cv.notify_all();
        return stop;
    }
    void finish(Running& g, const char* reason) {
        Request& r = *g.r;
        b2_seq_free(model_, g.seq);
        std::lock_guard<std::mutex> l(r.mu);
        g.utf8 += g.tail;
        const std::string rest = take_valid_utf8(g.utf8, true);
        if (!rest.empty()) r.pieces.push_back(rest);
        r.eval_s = std::chrono::duration<double>(Clock::now() - g.t_eval).count();
        r.done_reason = reason;
        r.done = true;
        r.cv.notify_all();
    }

    void run() {
        std::vector<Running> running;
        std::vector<float> logits;
        for (;;) {
            std::shared_ptr<Request> admit;
            {
                std::unique_lock<std::mutex> l(mu_);
                cv_.wait(l, [&] { return stop_ || !load_want_.empty() || !waiting_.empty() || !running.empty(); });
                if (stop_) break;
                if (!load_want_.empty() && running.empty()) {         // switch / load models only between generations
                    const std::string want = load_want_;
                    load_want_.clear();
                    l.unlock();
                    load(want);
                    l.lock();
                    load_done_ = true;
                    cv_load_.notify_all();
                    continue;
                }
                if (!waiting_.empty() && model_ && (int)running.size() < info_.max_batch && load_want_.empty()) { admit = waiting_.front(); waiting_.pop_front(); }
            }
            if (admit) {                       // ---- admit: prefill into a free slot; its last logits give the first token ----
                Running g{admit, -1, 0, 0, std::mt19937(admit->opt.seeded ? admit->opt.seed : std::random_device{}()), std::string(), Clock::now()};
                const int n = (int)admit->prompt.size();
                if (n == 0) { fail(admit, "empty prompt"); continue; }
                if (n + 1 > info_.max_ctx) { fail(admit, "prompt of " + std::to_string(n) + " tokens exceeds the context of " + std::to_string(info_.max_ctx)); continue; }
                if (b2_seq_new(model_, &g.seq) != B2_OK) { fail(admit, b2_strerror()); continue; }
                logits.resize((size_t)info_.n_vocab * info_.max_batch);
                const auto t0 = Clock::now();
                if (b2_prefill(model_, g.seq, admit->prompt.data(), n, logits.data()) != B2_OK) { b2_seq_free(model_, g.seq); fail(admit, b2_strerror()); continue; }
                admit->prompt_eval_s = std::chrono::duration<double>(Clock::now() - t0).count();
                admit->load_s = last_load_s_;
                g.t_eval = Clock::now();
                g.last = sample(g, logits.data());
                g.n_gen = 1;
                if (emit(g, g.last) || g.n_gen >= admit->opt.num_predict) { finish(g, g.last == tinfo_.eos ? "stop" : g.n_gen >= admit->opt.num_predict ? "length" : "stop"); continue; }
                running.push_back(std::move(g));
                continue;                      // look for more waiting requests before the next step
            }
            if (running.empty()) continue;
            // ---- one batched decode step for every running sequence ----
            const int B = (int)running.size();
            std::vector<int> seqs(B);
            std::vector<int32_t> last(B), next(B);
            bool need_logits = false;
            for (int i = 0; i < B; i++) { seqs[i] = running[i].seq; last[i] = running[i].last; need_logits |= running[i].r->opt.temperature > 0.f; }
            logits.resize((size_t)info_.n_vocab * info_.max_batch);
            if (b2_decode(model_, B, seqs.data(), last.data(), next.data(), need_logits ? logits.data() : nullptr) != B2_OK) {
                const std::string e = b2_strerror();
                for (Running& g : running) { b2_seq_free(model_, g.seq); fail(g.r, e); }
                running.clear();
                continue;
            }
            for (int i = B - 1; i >= 0; i--) {
                Running& g = running[i];
                g.last = g.r->opt.temperature > 0.f ? sample(g, logits.data() + (size_t)i * info_.n_vocab) : next[i];
                g.n_gen++;
                const bool ctx_full = b2_seq_len(model_, g.seq) + 1 >= info_.max_ctx;
                const bool stop = g.r->cancelled.load() || emit(g, g.last);
                if (stop || g.n_gen >= g.r->opt.num_predict || ctx_full) {
                    finish(g, stop ? "stop" : "length");
                    running.erase(running.begin() + i);
                }
            }
        }
        for (Running& g : running) { b2_seq_free(model_, g.seq); fail(g.r, "server shutting down"); }
        unload();
        if (ctx_) b2_shutdown(ctx_);
    }

    std::string store_;
    int max_batch_, max_ctx_;
    std::thread thread_;
    std::mutex mu_;
    std::condition_variable cv_, cv_load_;
    bool stop_ = false, load_done_ = false;
    std::string load_want_, load_name_, load_err_;
    std::deque<std::shared_ptr<Request>> waiting_;
    b2_ctx* ctx_ = nullptr;
    b2_model* model_ = nullptr;
    b2_tokenizer* tok_ = nullptr;
    b2_store_entry entry_;
    b2_model_info info_;
    b2_tokenizer_info tinfo_;
    double last_load_s_ = 0;
};

// ---- HTTP plumbing ----------------------------------------------------------------------------------------------
struct HttpReq { std::string method, path, body; std::map<std::string, std::string> hdr; };

static bool send_all(int fd, const char* p, size_t n) {
    while (n) {
        const ssize_t k = send(fd, p, n, MSG_NOSIGNAL);
        if (k <= 0) { if (errno == EINTR) continue; return false; }
        p += k; n -= (size_t)k;
    }
    return true;
}
static bool send_str(int fd, const std::string& s) { return send_all(fd, s.data(), s.size()); }
static bool read_request(int fd, HttpReq* r) {
    std::string buf;
    char tmp[8192];
    size_t hdr_end;
    while ((hdr_end = buf.find("\r\n\r\n")) == std::string::npos) {
        const ssize_t k = recv(fd, tmp, sizeof tmp, 0);
        if (k <= 0) return false;
        buf.append(tmp, (size_t)k);
        if (buf.size() > (1u << 20)) return false;
    }
    std::istringstream hs(buf.substr(0, hdr_end));
    std::string line;
    std::getline(hs, line);
    std::istringstream rl(line);
    std::string ver;
    rl >> r->method >> r->path >> ver;
    while (std::getline(hs, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        const size_t c = line.find(':');
        if (c == std::string::npos) continue;
        std::string k = line.substr(0, c), v = line.substr(c + 1);
        std::transform(k.begin(), k.end(), k.begin(), ::tolower);
        while (!v.empty() && v[0] == ' ') v.erase(0, 1);
        r->hdr[k] = v;
    }
    size_t want = 0;
    if (r->hdr.count("content-length")) want = (size_t)strtoull(r->hdr["content-length"].c_str(), nullptr, 10);
    if (want > (64u << 20)) return false;
    r->body = buf.substr(hdr_end + 4);
    while (r->body.size() < want) {
        const ssize_t k = recv(fd, tmp, sizeof tmp, 0);
        if (k <= 0) return false;
        r->body.append(tmp, (size_t)k);
    }
    const size_t q = r->path.find('?');
    if (q != std::string::npos) r->path.erase(q);
    return true;
}
static void respond(int fd, int code, const char* ctype, const std::string& body, bool head_only = false) {
    const char* reason = code == 200 ? "OK" : code == 400 ? "Bad Request" : code == 404 ? "Not Found" : code == 405 ? "Method Not Allowed" : "Internal Server Error";
    std::string h = "HTTP/1.1 " + std::to_string(code) + " " + reason + "\r\nContent-Type: " + ctype + "\r\nContent-Length: " + std::to_string(body.size()) +
                    "\r\nConnection: close\r\n\r\n";
    send_str(fd, h);
    if (!head_only) send_str(fd, body);
}
static void respond_error(int fd, int code, const std::string& msg) { respond(fd, code, "application/json", "{\"error\":\"" + json_escape(msg) + "\"}"); }
struct Chunked {      // a streamed response body (HTTP/1.1 chunked transfer coding)
    int fd; bool ok = true;
    Chunked(int f, const char* ctype) : fd(f) { ok = send_str(fd, std::string("HTTP/1.1 200 OK\r\nContent-Type: ") + ctype + "\r\nTransfer-Encoding: chunked\r\nConnection: close\r\n\r\n"); }
    void write(const std::string& s) {
        if (!ok || s.empty()) return;
        char len[32];
        snprintf(len, sizeof len, "%zx\r\n", s.size());
        ok = send_str(fd, len) && send_str(fd, s) && send_str(fd, "\r\n");
    }
    void end() { if (ok) send_str(fd, "0\r\n\r\n"); }
};

// ---- request handling ---------------------------------------------------------------------------------------------
static std::string g_store;
static Engine* g_engine = nullptr;

static std::string jstr(const Json& j, const char* k, const std::string& d = "") { const Json* v = j.get(k); return v && v->t == Json::STR ? v->str : d; }
static bool jbool(const Json& j, const char* k, bool d) { const Json* v = j.get(k); return v && v->t == Json::BOOL ? v->b : d; }

static Options parse_options(const Json& body, const std::string& params_json) {
    Options o;
    auto apply = [&](const Json& j) {
        if (j.t != Json::OBJ) return;
        if (const Json* v = j.get("num_predict")) if (v->t == Json::NUM) o.num_predict = (int)v->num;
        if (const Json* v = j.get("max_tokens")) if (v->t == Json::NUM) o.num_predict = (int)v->num;
        if (const Json* v = j.get("temperature")) if (v->t == Json::NUM) o.temperature = (float)v->num;
        if (const Json* v = j.get("top_k")) if (v->t == Json::NUM) o.top_k = (int)v->num;
        if (const Json* v = j.get("top_p")) if (v->t == Json::NUM) o.top_p = (float)v->num;
        if (const Json* v = j.get("seed")) if (v->t == Json::NUM) { o.seed = (uint32_t)v->num; o.seeded = true; }
        if (const Json* v = j.get("stop")) {
            if (v->t == Json::STR) o.stop = {v->str};
            else if (v->t == Json::ARR) { o.stop.clear(); for (const Json& s : v->arr) if (s.t == Json::STR && !s.str.empty()) o.stop.push_back(s.str); }
        }
    };
    bool ok = false;
    if (!params_json.empty()) apply(b2::json_parse(params_json, &ok));      // the model's params layer first, the request overrides it
    if (const Json* opt = body.get("options")) apply(*opt);
    apply(body);                                                             // OpenAI-style top-level fields
    if (o.num_predict < 0) o.num_predict = 1 << 30;
    if (o.num_predict == 0) o.num_predict = 1;
    return o;
}

// Ollama's Go templates reduced to what the stock model templates use: {{ .System }}, {{ .Prompt }}, {{ .Response }} and
// the {{ if .System }} ... {{ end }} guard.  A model without a template layer gets the bare prompt.
static std::string render_template(std::string t, const std::string& system, const std::string& prompt) {
    if (t.empty()) return (system.empty() ? "" : system + "\n\n") + prompt;
    for (;;) {                                   // {{ if .System }} body {{ end }}
        const size_t a = t.find("{{ if .System }}"), a2 = t.find("{{- if .System }}");
        const size_t s = std::min(a, a2);
        if (s == std::string::npos) break;
        const size_t open_end = t.find("}}", s) + 2;
        const size_t e = t.find("{{ end }}", open_end), e2 = t.find("{{- end }}", open_end);
        const size_t en = std::min(e, e2);
        if (en == std::string::npos) break;
        const size_t close_end = t.find("}}", en) + 2;
        const std::string body = system.empty() ? "" : t.substr(open_end, en - open_end);
        t = t.substr(0, s) + body + t.substr(close_end);
    }
    auto sub = [&](const char* key, const std::string& val) {
        for (size_t at; (at = t.find(key)) != std::string::npos;) t.replace(at, strlen(key), val);
    };
    sub("{{ .System }}", system); sub("{{.System}}", system);
    sub("{{ .Prompt }}", prompt); sub("{{.Prompt}}", prompt);
    const size_t r = std::min(t.find("{{ .Response }}"), t.find("{{.Response}}"));
    if (r != std::string::npos) t.erase(r);       // generation continues where the response would go
    return t;
}

static std::vector<int32_t> tokenize(b2_tokenizer* tok, const std::string& text, bool add_bos, bool special) {
    std::vector<int32_t> ids(text.size() + 16);
    int n = b2_tokenize(tok, text.c_str(), add_bos ? -1 : 0, special ? 1 : 0, ids.data(), (int)ids.size());
    if (n > (int)ids.size()) { ids.resize((size_t)n); n = b2_tokenize(tok, text.c_str(), add_bos ? -1 : 0, special ? 1 : 0, ids.data(), n); }
    ids.resize((size_t)std::max(0, n));
    return ids;
}

enum class Api { GENERATE, CHAT, OPENAI };

static void handle_generation(int fd, const HttpReq& hr, Api api) {
    bool ok = false;
    const Json body = b2::json_parse(hr.body, &ok);
    if (!ok || body.t != Json::OBJ) return respond_error(fd, 400, "invalid JSON body");
    const std::string model = jstr(body, "model", jstr(body, "name"));
    if (model.empty()) return respond_error(fd, 400, "model is required");
    b2_tokenizer* tok = nullptr;
    b2_store_entry entry;
    std::string err;
    if (g_engine->ensure_model(model, &tok, &entry, &err) != 0) return respond_error(fd, 404, err);
    const std::string tmpl = read_text(entry.template_path), sys_default = read_text(entry.system_path), params = read_text(entry.params_path);

    auto req = std::make_shared<Request>();
    req->opt = parse_options(body, params);
    std::string prompt_text;
    if (api == Api::GENERATE) {
        const std::string system = jstr(body, "system", sys_default);
        prompt_text = jbool(body, "raw", false) ? jstr(body, "prompt") : render_template(tmpl, system, jstr(body, "prompt"));
        if (const Json* c = body.get("context")) if (c->t == Json::ARR) for (const Json& t : c->arr) if (t.t == Json::NUM) req->prompt.push_back((int32_t)t.num);
    } else {
        const Json* msgs = body.get("messages");
        if (!msgs || msgs->t != Json::ARR || msgs->arr.empty()) return respond_error(fd, 400, "messages is required");
        // each user turn goes through the template; earlier assistant turns are appended as plain text after their prompt
        std::string system = sys_default, acc;
        for (const Json& m : msgs->arr) {
            const std::string role = jstr(m, "role"), content = jstr(m, "content");
            if (role == "system") system = content;
            else if (role == "user") { acc += render_template(tmpl, system, content); system.clear(); }
            else if (role == "assistant") acc += content;
        }
        prompt_text = acc;
    }
    {
        const std::vector<int32_t> ids = tokenize(tok, prompt_text, req->prompt.empty(), true);
        req->prompt.insert(req->prompt.end(), ids.begin(), ids.end());
    }
    const bool stream = jbool(body, "stream", api != Api::OPENAI);
    const auto t_start = Clock::now();
    const std::string created = now_rfc3339();
    const long created_unix = (long)time(nullptr);
    g_engine->submit(req);

    auto frame = [&](const std::string& piece, bool done) -> std::string {          // one streamed object
        if (api == Api::GENERATE) return "{\"model\":\"" + json_escape(model) + "\",\"created_at\":\"" + created + "\",\"response\":\"" + json_escape(piece) + "\",\"done\":" + (done ? "true" : "false");
        if (api == Api::CHAT) return "{\"model\":\"" + json_escape(model) + "\",\"created_at\":\"" + created + "\",\"message\":{\"role\":\"assistant\",\"content\":\"" + json_escape(piece) + "\"},\"done\":" + (done ? "true" : "false");
        return "{\"id\":\"chatcmpl-b2\",\"object\":\"chat.completion.chunk\",\"created\":" + std::to_string(created_unix) + ",\"model\":\"" + json_escape(model) +
               "\",\"choices\":[{\"index\":0,\"delta\":{\"role\":\"assistant\",\"content\":\"" + json_escape(piece) + "\"},\"finish_reason\":" + (done ? "\"stop\"" : "null") + "}]";
    };
    auto stats = [&]() -> std::string {
        std::ostringstream s;
        const double total = std::chrono::duration<double>(Clock::now() - t_start).count();
        s << ",\"done_reason\":\"" << req->done_reason << "\",\"total_duration\":" << (long long)(total * 1e9) << ",\"load_duration\":" << (long long)(req->load_s * 1e9)
          << ",\"prompt_eval_count\":" << req->prompt.size() << ",\"prompt_eval_duration\":" << (long long)(req->prompt_eval_s * 1e9) << ",\"eval_count\":" << req->generated.size()
          << ",\"eval_duration\":" << (long long)(req->eval_s * 1e9);
        if (api == Api::GENERATE) {
            s << ",\"context\":[";
            bool first = true;
            for (int32_t t : req->prompt) { s << (first ? "" : ",") << t; first = false; }
            for (int32_t t : req->generated) { s << (first ? "" : ",") << t; first = false; }
            s << "]";
        }
        return s.str();
    };

    std::unique_ptr<Chunked> out;
    if (stream) out.reset(new Chunked(fd, api == Api::OPENAI ? "text/event-stream" : "application/x-ndjson"));
    std::string whole;
    for (;;) {
        std::deque<std::string> pieces;
        bool done;
        {
            std::unique_lock<std::mutex> l(req->mu);
            req->cv.wait(l, [&] { return req->done || !req->pieces.empty(); });
            pieces.swap(req->pieces);
            done = req->done;
        }
        for (const std::string& p : pieces) {
            whole += p;
            if (stream) {
                out->write(api == Api::OPENAI ? "data: " + frame(p, false) + "}\n\n" : frame(p, false) + "}\n");
                if (!out->ok) req->cancelled = true;                 // the client went away: the engine retires the sequence at its next step
            }
        }
        if (done) break;
    }
    if (!req->error.empty()) {
        if (stream) { out->write("{\"error\":\"" + json_escape(req->error) + "\"}\n"); out->end(); }
        else respond_error(fd, 500, req->error);
        return;
    }
    if (stream) {
        if (api == Api::OPENAI) { out->write("data: " + frame("", true) + "}\n\n"); out->write("data: [DONE]\n\n"); }
        else out->write(frame("", true) + stats() + "}\n");
        out->end();
    } else if (api == Api::OPENAI) {
        respond(fd, 200, "application/json",
                "{\"id\":\"chatcmpl-b2\",\"object\":\"chat.completion\",\"created\":" + std::to_string(created_unix) + ",\"model\":\"" + json_escape(model) +
                    "\",\"choices\":[{\"index\":0,\"message\":{\"role\":\"assistant\",\"content\":\"" + json_escape(whole) + "\"},\"finish_reason\":\"" +
                    (req->done_reason == "length" ? "length" : "stop") + "\"}],\"usage\":{\"prompt_tokens\":" + std::to_string(req->prompt.size()) + ",\"completion_tokens\":" +
                    std::to_string(req->generated.size()) + ",\"total_tokens\":" + std::to_string(req->prompt.size() + req->generated.size()) + "}}");
    } else {
        respond(fd, 200, "application/json", frame(whole, true) + stats() + "}");
    }
}

static std::string tags_json() {
    std::string buf(1 << 16, '\0');
    int n = b2_store_list(g_store.c_str(), &buf[0], (int)buf.size());
    if (n >= (int)buf.size()) { buf.assign((size_t)n + 1, '\0'); n = b2_store_list(g_store.c_str(), &buf[0], (int)buf.size()); }
    buf.resize((size_t)std::max(0, n));
    std::istringstream ls(buf);
    std::string line, out = "{\"models\":[";
    bool first = true;
    while (std::getline(ls, line)) {
        std::istringstream f(line);
        std::string name, size, digest, mtime;
        std::getline(f, name, '\t'); std::getline(f, size, '\t'); std::getline(f, digest, '\t'); std::getline(f, mtime, '\t');
        const time_t mt = (time_t)strtoll(mtime.c_str(), nullptr, 10);
        tm t; gmtime_r(&mt, &t);
        char ts[40]; strftime(ts, sizeof ts, "%Y-%m-%dT%H:%M:%SZ", &t);
        out += std::string(first ? "" : ",") + "{\"name\":\"" + json_escape(name) + "\",\"model\":\"" + json_escape(name) + "\",\"modified_at\":\"" + ts + "\",\"size\":" +
               (size.empty() ? "0" : size) + ",\"digest\":\"" + json_escape(digest.size() > 7 ? digest.substr(7) : digest) + "\",\"details\":{\"format\":\"gguf\"}}";
        first = false;
    }
    return out + "]}";
}

static void handle_show(int fd, const HttpReq& hr) {
    bool ok = false;
    const Json body = b2::json_parse(hr.body, &ok);
    const std::string model = ok ? jstr(body, "name", jstr(body, "model")) : "";
    if (model.empty()) return respond_error(fd, 400, "name is required");
    b2_store_entry e;
    if (b2_store_resolve(g_store.c_str(), model.c_str(), &e) != B2_OK) return respond_error(fd, 404, std::string("model '") + model + "' not found");
    const std::string tmpl = read_text(e.template_path), params = read_text(e.params_path), sys = read_text(e.system_path);
    respond(fd, 200, "application/json",
            "{\"modelfile\":\"FROM " + json_escape(e.model_path) + "\",\"parameters\":\"" + json_escape(params) + "\",\"template\":\"" + json_escape(tmpl) + "\",\"system\":\"" +
                json_escape(sys) + "\",\"details\":{\"format\":\"gguf\",\"digest\":\"" + json_escape(e.digest) + "\",\"size\":" + std::to_string(e.model_size) + "}}");
}

static void serve_connection(int fd) {
    HttpReq r;
    if (read_request(fd, &r)) {
        const bool get = r.method == "GET", head = r.method == "HEAD", post = r.method == "POST";
        if ((get || head) && r.path == "/") respond(fd, 200, "text/plain; charset=utf-8", "Ollama is running", head);
        else if ((get || head) && r.path == "/api/tags") respond(fd, 200, "application/json", tags_json(), head);          // both probes: no GPU, no engine lock
        else if (get && r.path == "/api/version") respond(fd, 200, "application/json", std::string("{\"version\":\"") + kVersion + "\"}");
        else if (post && r.path == "/api/show") handle_show(fd, r);
        else if (post && r.path == "/api/generate") handle_generation(fd, r, Api::GENERATE);
        else if (post && r.path == "/api/chat") handle_generation(fd, r, Api::CHAT);
        else if (post && r.path == "/v1/chat/completions") handle_generation(fd, r, Api::OPENAI);
        else if (r.path == "/api/pull" || r.path == "/api/push" || r.path == "/api/create" || r.path == "/api/delete" || r.path == "/api/copy")
            respond_error(fd, 405, "this server serves a read-only model store; pulling is the puller container's job");
        else respond_error(fd, 404, "not found");
    }
    shutdown(fd, SHUT_RDWR);
    close(fd);
}

int main(int argc, char** argv) {
    if (argc > 1 && strcmp(argv[1], "serve") != 0) {
        fprintf(stderr, "usage: %s serve     (OLLAMA_HOST=host[:port], OLLAMA_MODELS=<store>, B2_MAX_BATCH, B2_MAX_CTX)\n", argv[0]);
        return 2;
    }
    signal(SIGPIPE, SIG_IGN);
    std::string host = getenv("OLLAMA_HOST") ? getenv("OLLAMA_HOST") : "127.0.0.1";
    int port = 11434;
    const size_t scheme = host.find("://");
    if (scheme != std::string::npos) host = host.substr(scheme + 3);
    const size_t colon = host.rfind(':');
    if (colon != std::string::npos) { port = atoi(host.c_str() + colon + 1); host = host.substr(0, colon); }
    if (host.empty()) host = "0.0.0.0";
    g_store = getenv("OLLAMA_MODELS") ? getenv("OLLAMA_MODELS") : "/root/.ollama";
    Engine engine(g_store, getenv("B2_MAX_BATCH") ? atoi(getenv("B2_MAX_BATCH")) : 8, getenv("B2_MAX_CTX") ? atoi(getenv("B2_MAX_CTX")) : 4096);
    g_engine = &engine;

    const int ls = socket(AF_INET, SOCK_STREAM, 0);
    int one = 1;
    setsockopt(ls, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
    sockaddr_in addr;
    memset(&addr, 0, sizeof addr);
    addr.sin_family = AF_INET;
    addr.sin_port = htons((uint16_t)port);
    if (inet_pton(AF_INET, host.c_str(), &addr.sin_addr) != 1) addr.sin_addr.s_addr = htonl(host == "localhost" ? INADDR_LOOPBACK : INADDR_ANY);
    if (bind(ls, (sockaddr*)&addr, sizeof addr) != 0 || listen(ls, 128) != 0) { logf("cannot listen on %s:%d: %s", host.c_str(), port, strerror(errno)); return 1; }
    logf("listening on %s:%d, model store %s, version %s", host.c_str(), port, g_store.c_str(), kVersion);
    for (;;) {
        const int fd = accept(ls, nullptr, nullptr);
        if (fd < 0) { if (errno == EINTR) continue; break; }
        setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
        std::thread(serve_connection, fd).detach();
    }
    return 0;
}

// store.cpp -- reader for the model store `ollama pull` fills (the read-only volume the operator mounts at
// /root/.ollama in every model pod: reference pkg/model/pod.go:34-40, filled through the puller of
// pkg/model/pod.go:68-83 into the store of pkg/model/image_store.go:172-181).
//
// Layout [upstream ollama server/modelpath.go, server/manifest.go]:
//   <models>/manifests/<registry>/<namespace>/<model>/<tag>      OCI-style JSON manifest
//   <models>/blobs/sha256-<hex>                                   layer blobs (older stores: sha256:<hex>)
// A name is [registry/][namespace/]model[:tag]; defaults registry.ollama.ai, library, latest.  The layer of media type
// application/vnd.ollama.image.model is the GGUF file; template / system / params layers are small text or JSON blobs.
#include <dirent.h>
#include <string.h>
#include <sys/stat.h>

#include <algorithm>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/b2gguf.h"
#include "common.cuh"
#include "json_min.h"

namespace b2 {
namespace {

bool read_file(const std::string& path, std::string* out, size_t limit = 16u << 20) {
    std::ifstream f(path, std::ios::binary);
    if (!f) return false;
    std::ostringstream ss;
    ss << f.rdbuf();
    *out = ss.str();
    return out->size() <= limit;
}
bool exists(const std::string& p) {
    struct stat sb;
    return stat(p.c_str(), &sb) == 0;
}
bool is_dir(const std::string& p) {
    struct stat sb;
    return stat(p.c_str(), &sb) == 0 && S_ISDIR(sb.st_mode);
}

// the directory that holds manifests/ and blobs/: the argument itself, its models/ child (an OLLAMA home such as
// /root/.ollama), $OLLAMA_MODELS, or $HOME/.ollama/models
std::string models_dir(const char* arg) {
    std::vector<std::string> cand;
    if (arg && *arg) { cand.push_back(arg); cand.push_back(std::string(arg) + "/models"); }
    if (const char* e = getenv("OLLAMA_MODELS")) cand.push_back(e);
    if (const char* h = getenv("HOME")) cand.push_back(std::string(h) + "/.ollama/models");
    cand.push_back("/root/.ollama/models");
    for (const std::string& c : cand)
        if (is_dir(c + "/manifests")) return c;
    return std::string();
}

struct Name { std::string registry = "registry.ollama.ai", ns = "library", model, tag = "latest"; };
bool parse_name(const std::string& s, Name* n) {
    std::string rest = s;
    const size_t scheme = rest.find("://");
    if (scheme != std::string::npos) rest = rest.substr(scheme + 3);
    const size_t slash_last = rest.rfind('/');
    const size_t colon = rest.rfind(':');
    if (colon != std::string::npos && (slash_last == std::string::npos || colon > slash_last)) { n->tag = rest.substr(colon + 1); rest = rest.substr(0, colon); }
    std::vector<std::string> parts;
    for (size_t at = 0;;) {
        const size_t sl = rest.find('/', at);
        parts.push_back(rest.substr(at, sl == std::string::npos ? std::string::npos : sl - at));
        if (sl == std::string::npos) break;
        at = sl + 1;
    }
    if (parts.empty() || parts.size() > 3) return false;
    n->model = parts.back();
    if (parts.size() >= 2) n->ns = parts[parts.size() - 2];
    if (parts.size() == 3) n->registry = parts[0];
    for (const std::string* p : {&n->registry, &n->ns, &n->model, &n->tag})
        if (p->empty() || p->find("..") != std::string::npos || p->find('/') != std::string::npos) return false;      // no path tricks
    return true;
}

std::string blob_path(const std::string& root, const std::string& digest) {
    std::string d = digest;
    std::replace(d.begin(), d.end(), ':', '-');
    const std::string p = root + "/blobs/" + d;
    if (exists(p)) return p;
    const std::string old = root + "/blobs/" + digest;          // stores written before the '-' convention
    return exists(old) ? old : p;
}

void put(char* dst, size_t cap, const std::string& s) {
    const size_t n = std::min(cap - 1, s.size());
    memcpy(dst, s.data(), n);
    dst[n] = 0;
}

}  // namespace
}  // namespace b2

using namespace b2;

extern "C" int b2_store_resolve(const char* models, const char* name, b2_store_entry* out) {
    if (!name || !out) B2_FAIL(B2_ERR_ARG, "b2_store_resolve: null argument");
    memset(out, 0, sizeof *out);
    const std::string root = models_dir(models);
    if (root.empty()) B2_FAIL(B2_ERR_IO, "model store: no manifests/ directory under %s (nor $OLLAMA_MODELS, ~/.ollama/models)", models && *models ? models : "(default)");
    Name n;
    if (!parse_name(name, &n)) B2_FAIL(B2_ERR_ARG, "model store: '%s' is not a [registry/][namespace/]model[:tag] name", name);
    const std::string mpath = root + "/manifests/" + n.registry + "/" + n.ns + "/" + n.model + "/" + n.tag;
    std::string text;
    if (!read_file(mpath, &text)) B2_FAIL(B2_ERR_IO, "model store: model '%s' is not pulled (no manifest %s)", name, mpath.c_str());
    JsonParser jp{text.data(), text.data() + text.size()};
    const Json man = jp.value();
    const Json* layers = man.get("layers");
    if (!jp.ok || man.t != Json::OBJ || !layers || layers->t != Json::ARR) B2_FAIL(B2_ERR_FORMAT, "model store: %s is not a manifest", mpath.c_str());
    for (const Json& l : layers->arr) {
        const Json *mt = l.get("mediaType"), *dg = l.get("digest"), *sz = l.get("size");
        if (!mt || !dg || mt->t != Json::STR || dg->t != Json::STR) continue;
        if (dg->str.compare(0, 7, "sha256:") != 0 || dg->str.find('/') != std::string::npos || dg->str.find("..") != std::string::npos)
            B2_FAIL(B2_ERR_FORMAT, "model store: bad digest '%s' in %s", dg->str.c_str(), mpath.c_str());
        const std::string bp = blob_path(root, dg->str);
        if (mt->str == "application/vnd.ollama.image.model") {
            put(out->model_path, sizeof out->model_path, bp);
            put(out->digest, sizeof out->digest, dg->str);
            out->model_size = sz && sz->t == Json::NUM ? (uint64_t)sz->num : 0;
        } else if (mt->str == "application/vnd.ollama.image.template") put(out->template_path, sizeof out->template_path, bp);
        else if (mt->str == "application/vnd.ollama.image.system") put(out->system_path, sizeof out->system_path, bp);
        else if (mt->str == "application/vnd.ollama.image.params") put(out->params_path, sizeof out->params_path, bp);
    }
    if (!out->model_path[0]) B2_FAIL(B2_ERR_FORMAT, "model store: manifest %s has no application/vnd.ollama.image.model layer", mpath.c_str());
    if (!exists(out->model_path)) B2_FAIL(B2_ERR_IO, "model store: blob %s of '%s' is missing", out->model_path, name);
    put(out->name, sizeof out->name, (n.ns == "library" ? "" : n.ns + "/") + n.model + ":" + n.tag);
    struct stat sb;
    if (stat(mpath.c_str(), &sb) == 0) out->modified_unix = (int64_t)sb.st_mtime;
    return B2_OK;
}

// Every pulled model, one line each: "<name>:<tag>\t<bytes of the GGUF layer>\t<digest>\t<mtime>\n" (what GET /api/tags lists).
// Returns the number of bytes the full listing needs (excluding the terminator); fills at most cap - 1 of them.
extern "C" int b2_store_list(const char* models, char* buf, int cap) {
    const std::string root = models_dir(models);
    std::string listing;
    if (!root.empty()) {
        auto subdirs = [](const std::string& d, bool dirs) {
            std::vector<std::string> v;
            if (DIR* h = opendir(d.c_str())) {
                while (dirent* e = readdir(h)) {
                    const std::string n = e->d_name;
                    if (n == "." || n == "..") continue;
                    if (is_dir(d + "/" + n) == dirs) v.push_back(n);
                }
                closedir(h);
            }
            std::sort(v.begin(), v.end());
            return v;
        };
        const std::string mroot = root + "/manifests";
        for (const std::string& reg : subdirs(mroot, true))
            for (const std::string& ns : subdirs(mroot + "/" + reg, true))
                for (const std::string& model : subdirs(mroot + "/" + reg + "/" + ns, true))
                    for (const std::string& tag : subdirs(mroot + "/" + reg + "/" + ns + "/" + model, false)) {
                        b2_store_entry e;
                        const std::string full = reg + "/" + ns + "/" + model + ":" + tag;
                        if (b2_store_resolve(root.c_str(), full.c_str(), &e) != B2_OK) continue;
                        listing += std::string(e.name) + "\t" + std::to_string(e.model_size) + "\t" + e.digest + "\t" + std::to_string(e.modified_unix) + "\n";
                    }
    }
    if (buf && cap > 0) put(buf, (size_t)cap, listing);
    return (int)listing.size();
}

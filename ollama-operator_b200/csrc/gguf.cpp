// gguf.cpp -- GGUF v2/v3 parser over an mmap of the file.
#include "gguf.h"

#include <errno.h>
#include <fcntl.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "common.cuh"
#include "formats.cuh"

namespace b2 {

namespace {

struct Cursor {
    const uint8_t* p;
    const uint8_t* end;
    bool ok = true;
    template <typename T> T get() {
        T v{};
        if (!ok || (size_t)(end - p) < sizeof(T)) { ok = false; return v; }
        memcpy(&v, p, sizeof(T));
        p += sizeof(T);
        return v;
    }
    std::string str() {
        const uint64_t n = get<uint64_t>();
        if (!ok || (uint64_t)(end - p) < n) { ok = false; return std::string(); }
        std::string s((const char*)p, (size_t)n);
        p += n;
        return s;
    }
    void skip(uint64_t n) {
        if (!ok || (uint64_t)(end - p) < n) { ok = false; return; }
        p += n;
    }
};

// sizes of the scalar GGUF value types; 0 = not a fixed-size scalar
size_t scalar_size(uint32_t t) {
    switch (t) {
    case 0: case 1: case 7: return 1;
    case 2: case 3: return 2;
    case 4: case 5: case 6: return 4;
    case 10: case 11: case 12: return 8;
    }
    return 0;
}

bool read_value(Cursor& c, uint32_t t, GgufValue& v, int depth = 0) {
    v.type = (int)t;
    switch (t) {
    case 0: v.u = c.get<uint8_t>(); break;
    case 1: v.u = (uint64_t)(int64_t)c.get<int8_t>(); break;
    case 2: v.u = c.get<uint16_t>(); break;
    case 3: v.u = (uint64_t)(int64_t)c.get<int16_t>(); break;
    case 4: v.u = c.get<uint32_t>(); break;
    case 5: v.u = (uint64_t)(int64_t)c.get<int32_t>(); break;
    case 6: v.f = c.get<float>(); break;
    case 7: v.u = c.get<uint8_t>() != 0; break;
    case 8: v.s = c.str(); break;
    case 9: {
        const uint32_t et = c.get<uint32_t>();
        const uint64_t n = c.get<uint64_t>();
        v.arr_len = n;
        v.arr_type = (int)et;
        v.arr_data = c.p;
        if (const size_t es = scalar_size(et)) {
            uint64_t nbytes = 0;
            if (__builtin_mul_overflow(n, (uint64_t)es, &nbytes)) { c.ok = false; return false; }
            c.skip(nbytes);
        }
        else if (et == 8) { for (uint64_t i = 0; i < n && c.ok; i++) c.str(); }
        else if (et == 9 && depth < 4) { GgufValue tmp; for (uint64_t i = 0; i < n && c.ok; i++) read_value(c, 9, tmp, depth + 1); }
        else return false;
    } break;
    case 10: v.u = c.get<uint64_t>(); break;
    case 11: v.u = (uint64_t)c.get<int64_t>(); break;
    case 12: v.f = c.get<double>(); break;
    default: return false;
    }
    return c.ok;
}

}  // namespace

GgufFile::~GgufFile() {
    if (map_) munmap((void*)map_, size_);
    if (fd_ >= 0) close(fd_);
}

int GgufFile::open(const char* path) {
    if (!path) B2_FAIL(B2_ERR_ARG, "gguf: null path");
    fd_ = ::open(path, O_RDONLY);
    if (fd_ < 0) B2_FAIL(B2_ERR_IO, "gguf: cannot open %s: %s", path, strerror(errno));
    struct stat sb;
    if (fstat(fd_, &sb) != 0 || sb.st_size < 24) B2_FAIL(B2_ERR_FORMAT, "gguf: %s is too small to be a GGUF file", path);
    size_ = (size_t)sb.st_size;
    void* m = mmap(nullptr, size_, PROT_READ, MAP_PRIVATE, fd_, 0);
    if (m == MAP_FAILED) { map_ = nullptr; B2_FAIL(B2_ERR_IO, "gguf: mmap of %s failed: %s", path, strerror(errno)); }
    map_ = (const uint8_t*)m;

    Cursor c{map_, map_ + size_};
    if (c.get<uint32_t>() != 0x46554747u) B2_FAIL(B2_ERR_FORMAT, "gguf: %s: bad magic", path);
    version_ = c.get<uint32_t>();
    if (version_ != 2 && version_ != 3) B2_FAIL(B2_ERR_FORMAT, "gguf: %s: version %u not supported (2 or 3)", path, version_);
    const uint64_t n_tensors = c.get<uint64_t>(), n_kv = c.get<uint64_t>();
    if (n_tensors > (1u << 20) || n_kv > (1u << 20)) B2_FAIL(B2_ERR_FORMAT, "gguf: %s: implausible counts", path);
    for (uint64_t i = 0; i < n_kv; i++) {
        std::string key = c.str();
        const uint32_t t = c.get<uint32_t>();
        GgufValue v;
        if (!c.ok || !read_value(c, t, v)) B2_FAIL(B2_ERR_FORMAT, "gguf: %s: truncated or malformed metadata at key %llu", path, (unsigned long long)i);
        kv_[key] = v;
    }
    tensors_.resize((size_t)n_tensors);
    for (auto& t : tensors_) {
        t.name = c.str();
        const uint32_t nd = c.get<uint32_t>();
        if (!c.ok || nd > 4) B2_FAIL(B2_ERR_FORMAT, "gguf: %s: bad tensor header", path);
        t.ne.resize(nd);
        for (uint32_t d = 0; d < nd; d++) t.ne[d] = (int64_t)c.get<uint64_t>();
        t.type = (int)c.get<uint32_t>();
        t.offset = c.get<uint64_t>();
    }
    if (!c.ok) B2_FAIL(B2_ERR_FORMAT, "gguf: %s: truncated tensor table", path);
    uint64_t align = 32;
    get_u64("general.alignment", &align);
    if (align == 0 || (align & (align - 1))) B2_FAIL(B2_ERR_FORMAT, "gguf: %s: bad alignment %llu", path, (unsigned long long)align);
    const size_t data_start = ((size_t)(c.p - map_) + align - 1) / align * align;
    if (data_start > size_) B2_FAIL(B2_ERR_FORMAT, "gguf: %s: no tensor data section", path);
    const uint64_t data_size = size_ - data_start;
    for (size_t i = 0; i < tensors_.size(); i++) {
        auto& t = tensors_[i];
        // element count with overflow checks: a zero or negative dimension, or a product that wraps, must not
        // yield a small `bytes` that passes the range test below
        uint64_t n = 1;
        for (int64_t d : t.ne) {
            if (d <= 0) B2_FAIL(B2_ERR_FORMAT, "gguf: tensor %s: dimension %lld is not positive", t.name.c_str(), (long long)d);
            if (__builtin_mul_overflow(n, (uint64_t)d, &n)) B2_FAIL(B2_ERR_FORMAT, "gguf: tensor %s: element count overflows", t.name.c_str());
        }
        const int be = type_block_elems(t.type), bb = type_block_bytes(t.type);
        if (bb == 0) { t.bytes = 0; t.data = nullptr; index_[t.name] = i; continue; }   // unknown type: only an error if used
        if (t.ne.empty() || t.ne[0] % be) B2_FAIL(B2_ERR_FORMAT, "gguf: tensor %s: row length not a multiple of the block size", t.name.c_str());
        uint64_t bytes = 0;
        if (__builtin_mul_overflow(n / (uint64_t)be, (uint64_t)bb, &bytes)) B2_FAIL(B2_ERR_FORMAT, "gguf: tensor %s: byte size overflows", t.name.c_str());
        if (t.offset % align) B2_FAIL(B2_ERR_FORMAT, "gguf: tensor %s: offset %llu is not aligned to %llu", t.name.c_str(), (unsigned long long)t.offset, (unsigned long long)align);
        if (t.offset > data_size || bytes > data_size - t.offset) B2_FAIL(B2_ERR_FORMAT, "gguf: tensor %s runs past the end of the file", t.name.c_str());
        t.bytes = (size_t)bytes;
        t.data = map_ + data_start + t.offset;
        index_[t.name] = i;
    }
    return B2_OK;
}

const GgufTensor* GgufFile::find(const std::string& name) const {
    auto it = index_.find(name);
    return it == index_.end() ? nullptr : &tensors_[it->second];
}
bool GgufFile::get_u64(const std::string& key, uint64_t* v) const {
    auto it = kv_.find(key);
    if (it == kv_.end()) return false;
    const int t = it->second.type;
    if (t == 6 || t == 12 || t == 8 || t == 9) return false;
    *v = it->second.u;
    return true;
}
bool GgufFile::get_f64(const std::string& key, double* v) const {
    auto it = kv_.find(key);
    if (it == kv_.end()) return false;
    const int t = it->second.type;
    if (t == 6 || t == 12) { *v = it->second.f; return true; }
    if (t == 8 || t == 9) return false;
    *v = (double)(int64_t)it->second.u;
    return true;
}
bool GgufFile::get_str_array(const std::string& key, std::vector<std::string>* v) const {
    auto it = kv_.find(key);
    if (it == kv_.end() || it->second.type != 9 || it->second.arr_type != 8) return false;
    Cursor c{it->second.arr_data, map_ + size_};
    v->clear();
    v->reserve((size_t)it->second.arr_len);
    for (uint64_t i = 0; i < it->second.arr_len && c.ok; i++) v->push_back(c.str());
    return c.ok;
}
bool GgufFile::get_f32_array(const std::string& key, std::vector<float>* v) const {
    auto it = kv_.find(key);
    if (it == kv_.end() || it->second.type != 9 || it->second.arr_type != 6) return false;
    v->resize((size_t)it->second.arr_len);
    memcpy(v->data(), it->second.arr_data, v->size() * 4);       // bounds were checked when the array was skipped over in open()
    return true;
}
bool GgufFile::get_i32_array(const std::string& key, std::vector<int32_t>* v) const {
    auto it = kv_.find(key);
    if (it == kv_.end() || it->second.type != 9 || (it->second.arr_type != 5 && it->second.arr_type != 4)) return false;
    v->resize((size_t)it->second.arr_len);
    memcpy(v->data(), it->second.arr_data, v->size() * 4);
    return true;
}
bool GgufFile::get_str(const std::string& key, std::string* v) const {
    auto it = kv_.find(key);
    if (it == kv_.end() || it->second.type != 8) return false;
    *v = it->second.s;
    return true;
}

}  // namespace b2

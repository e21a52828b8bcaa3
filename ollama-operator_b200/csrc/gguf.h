// gguf.h -- minimal reader for GGUF v2/v3 files (the blobs `ollama pull` leaves under
// /root/.ollama/models/blobs, reference pkg/model/pod.go:34-40,68-83).  Host only.
#pragma once
#include <stdint.h>

#include <map>
#include <string>
#include <vector>

namespace b2 {

struct GgufTensor {
    std::string name;
    int type = 0;
    std::vector<int64_t> ne;     // ne[0] is the contiguous (column) dimension
    uint64_t offset = 0;         // relative to the data section
    const uint8_t* data = nullptr;
    size_t bytes = 0;
};

struct GgufValue {
    int type = -1;               // GGUF value type id
    uint64_t u = 0;              // integers / bool
    double f = 0;                // floats
    std::string s;               // strings
    uint64_t arr_len = 0;        // arrays: element count, element type and where the elements sit in the mapping
    int arr_type = -1;           // (scalar elements: packed; strings: {u64 length, bytes} one after the other)
    const uint8_t* arr_data = nullptr;
};

class GgufFile {
  public:
    ~GgufFile();
    int open(const char* path);                      // 0 or a b2 error code (message set)
    const GgufTensor* find(const std::string& name) const;
    bool get_u64(const std::string& key, uint64_t* v) const;
    bool get_f64(const std::string& key, double* v) const;
    bool get_str(const std::string& key, std::string* v) const;
    // arrays (tokenizer.ggml.tokens / scores / token_type / merges): copies out of the mapping
    bool get_str_array(const std::string& key, std::vector<std::string>* v) const;
    bool get_f32_array(const std::string& key, std::vector<float>* v) const;
    bool get_i32_array(const std::string& key, std::vector<int32_t>* v) const;
    const std::vector<GgufTensor>& tensors() const { return tensors_; }
    uint32_t version() const { return version_; }

  private:
    int fd_ = -1;
    const uint8_t* map_ = nullptr;
    size_t size_ = 0;
    uint32_t version_ = 0;
    std::map<std::string, GgufValue> kv_;
    std::vector<GgufTensor> tensors_;
    std::map<std::string, size_t> index_;
};

}  // namespace b2

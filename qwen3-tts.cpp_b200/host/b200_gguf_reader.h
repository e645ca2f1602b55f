// Minimal GGUF v2/v3 reader (metadata + tensor directory) with an mmap of the data section.
// Written from the file-format description in ggml/include/gguf.h:1-60 and the reader
// ggml/src/gguf.cpp:319+ (layout only — no ggml code is linked or needed at run time).
#pragma once

#include <cstddef>
#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace qwen3_tts {
namespace b200 {

enum gguf_value_type : uint32_t {
    GGUF_T_U8 = 0, GGUF_T_I8, GGUF_T_U16, GGUF_T_I16, GGUF_T_U32, GGUF_T_I32, GGUF_T_F32, GGUF_T_BOOL,
    GGUF_T_STRING, GGUF_T_ARRAY, GGUF_T_U64, GGUF_T_I64, GGUF_T_F64,
};

struct gguf_kv {
    uint32_t type = 0;
    uint64_t u = 0;          // integer / bool payload
    double   f = 0.0;        // float payload
    std::string s;           // string payload
    uint32_t arr_type = 0;   // arrays: element type and count (elements are skipped, strings kept)
    uint64_t arr_n = 0;
    std::vector<std::string> arr_s;
};

struct gguf_tensor_info {
    std::string name;
    uint32_t n_dims = 0;
    int64_t  ne[4] = {1, 1, 1, 1};
    uint32_t type = 0;       // ggml type id
    uint64_t offset = 0;     // relative to the data section
    size_t   nbytes = 0;
};

class GGUFFile {
public:
    ~GGUFFile();
    bool open(const std::string & path, std::string & err, bool keep_string_arrays = false);
    void close();

    const std::vector<gguf_tensor_info> & tensors() const { return tensors_; }
    const gguf_kv * find(const std::string & key) const;
    int64_t find_tensor(const std::string & name) const;
    size_t data_offset() const { return data_offset_; }
    uint32_t version() const { return version_; }
    uint32_t alignment() const { return alignment_; }
    // pointer into the read-only mapping of the whole file
    const uint8_t * tensor_data(size_t idx) const { return map_ + data_offset_ + tensors_[idx].offset; }
    size_t file_size() const { return size_; }

    static size_t type_nbytes(uint32_t type, const int64_t ne[4]);   // 0 = unknown type

private:
    int fd_ = -1;
    const uint8_t * map_ = nullptr;
    size_t size_ = 0;
    uint32_t version_ = 0, alignment_ = 32;
    size_t data_offset_ = 0;
    std::map<std::string, gguf_kv> kv_;
    std::vector<gguf_tensor_info> tensors_;
    std::map<std::string, size_t> by_name_;
};

} // namespace b200
} // namespace qwen3_tts

// GGUF loader and engine-context helpers of the B200 host classes: the role of the reference's
// GGUFLoader / init_preferred_backend / load_tensor_data_from_file (src/gguf_loader.h:16-81), same method
// names and meaning, on top of our own GGUF reader and the q3t engine instead of ggml.
//
// They live in namespace qwen3_tts::b200 (and in files of their own name) on purpose: the reference's
// pipeline keeps using ITS src/gguf_loader.h for the text tokenizer and the speaker encoder
// (src/qwen3_tts.cpp:138-151 needs loader.get_ctx(), a ggml handle), so both loaders are linked into one
// program and must not collide.
#pragma once

#include "../../include/q3t.h"
#include "b200_gguf_reader.h"

#include <cstdint>
#include <string>

namespace qwen3_tts {
namespace b200 {

class Q3T_API GGUFLoader {
public:
    GGUFLoader();
    ~GGUFLoader();

    bool open(const std::string & path);                       // src/gguf_loader.cpp:84-100
    void close();
    const std::string & get_error() const { return error_msg_; }

    int64_t get_n_tensors() const;
    const char * get_tensor_name(int64_t idx) const;
    int32_t get_tensor_type(int64_t idx) const;                // ggml type id (the reference returns enum ggml_type)
    size_t get_tensor_offset(int64_t idx) const;
    size_t get_tensor_size(int64_t idx) const;
    bool   get_tensor_ne(int64_t idx, int64_t ne_out[4]) const;
    int32_t get_u32(const char * key, int32_t default_val = 0) const;
    float get_f32(const char * key, float default_val = 0.0f) const;
    size_t get_data_offset() const;

    const GGUFFile & file() const { return file_; }
    const uint8_t * get_tensor_data(int64_t idx) const;        // host pointer into the mapping

protected:
    GGUFFile file_;
    bool is_open_ = false;
    std::string error_msg_;
    std::string file_path_;
};

// Upload every tensor whose name starts with one of `prefixes` to the engine
// (role of load_tensor_data_from_file, src/gguf_loader.cpp:154-227).
Q3T_API bool load_tensor_data_from_file(const GGUFLoader & loader, q3t_ctx * engine,
                                const char * const * prefixes, int n_prefixes,
                                std::string & error_msg, int64_t * n_uploaded = nullptr);

// Process-wide, ref-counted engine contexts = the reference's shared backend singleton
// (src/gguf_loader.cpp:10-82); one context serves one transformer and one vocoder, further
// instances of the same component get their own context.  The CUDA device is taken from QWEN3_TTS_DEVICE (default 0);
// there is no IGPU/ACCEL/CPU fallback chain: without a B200 this returns nullptr.
Q3T_API q3t_ctx * init_preferred_backend(const char * component_name, std::string * error_msg);
Q3T_API void release_preferred_backend(q3t_ctx * backend, const char * component_name = nullptr);

} // namespace b200
} // namespace qwen3_tts

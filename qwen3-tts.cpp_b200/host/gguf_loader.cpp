#include "gguf_loader.h"

#include <cstdlib>
#include <cstring>

namespace qwen3_tts {

namespace {
struct shared_backend_state {
    q3t_ctx * backend = nullptr;
    int32_t ref_count = 0;
};
shared_backend_state & get_shared_backend_state() {
    static shared_backend_state state;
    return state;
}
}

GGUFLoader::GGUFLoader() = default;
GGUFLoader::~GGUFLoader() { close(); }

q3t_ctx * init_preferred_backend(const char * component_name, std::string * error_msg) {
    if (error_msg) error_msg->clear();
    auto & shared = get_shared_backend_state();
    if (shared.backend) { shared.ref_count++; return shared.backend; }
    int device = 0;
    if (const char * e = std::getenv("QWEN3_TTS_DEVICE")) device = std::atoi(e);
    q3t_ctx * ctx = nullptr;
    if (q3t_ctx_create(device, &ctx) != Q3T_OK || !ctx) {
        if (error_msg) {
            const char * name = component_name ? component_name : "component";
            *error_msg = "Failed to initialize backend (CUDA sm_100a) for " + std::string(name) + ": " + q3t_last_error(nullptr);
        }
        return nullptr;
    }
    shared.backend = ctx;
    shared.ref_count = 1;
    return ctx;
}

void release_preferred_backend(q3t_ctx * backend) {
    if (!backend) return;
    auto & shared = get_shared_backend_state();
    if (shared.backend == backend) {
        if (--shared.ref_count <= 0) {
            q3t_ctx_destroy(shared.backend);
            shared.backend = nullptr;
            shared.ref_count = 0;
        }
        return;
    }
    q3t_ctx_destroy(backend);
}

bool GGUFLoader::open(const std::string & path) {
    close();
    file_path_ = path;
    if (!file_.open(path, error_msg_)) return false;
    is_open_ = true;
    return true;
}

void GGUFLoader::close() {
    file_.close();
    is_open_ = false;
    file_path_.clear();
}

int64_t GGUFLoader::get_n_tensors() const { return is_open_ ? (int64_t) file_.tensors().size() : 0; }

const char * GGUFLoader::get_tensor_name(int64_t idx) const {
    if (!is_open_ || idx < 0 || idx >= get_n_tensors()) return nullptr;
    return file_.tensors()[(size_t) idx].name.c_str();
}
int32_t GGUFLoader::get_tensor_type(int64_t idx) const {
    if (!is_open_ || idx < 0 || idx >= get_n_tensors()) return 0;     // GGML_TYPE_F32, as the reference
    return (int32_t) file_.tensors()[(size_t) idx].type;
}
size_t GGUFLoader::get_tensor_offset(int64_t idx) const {
    if (!is_open_ || idx < 0 || idx >= get_n_tensors()) return 0;
    return (size_t) file_.tensors()[(size_t) idx].offset;
}
size_t GGUFLoader::get_tensor_size(int64_t idx) const {
    if (!is_open_ || idx < 0 || idx >= get_n_tensors()) return 0;
    return file_.tensors()[(size_t) idx].nbytes;
}
bool GGUFLoader::get_tensor_ne(int64_t idx, int64_t ne_out[4]) const {
    if (!is_open_ || idx < 0 || idx >= get_n_tensors()) return false;
    for (int i = 0; i < 4; ++i) ne_out[i] = file_.tensors()[(size_t) idx].ne[i];
    return true;
}
const uint8_t * GGUFLoader::get_tensor_data(int64_t idx) const {
    if (!is_open_ || idx < 0 || idx >= get_n_tensors()) return nullptr;
    return file_.tensor_data((size_t) idx);
}
int32_t GGUFLoader::get_u32(const char * key, int32_t default_val) const {
    if (!is_open_) return default_val;
    const gguf_kv * kv = file_.find(key);
    if (!kv || kv->type == GGUF_T_STRING || kv->type == GGUF_T_ARRAY) return default_val;
    if (kv->type == GGUF_T_F32 || kv->type == GGUF_T_F64) return (int32_t) kv->f;
    return (int32_t) kv->u;
}
float GGUFLoader::get_f32(const char * key, float default_val) const {
    if (!is_open_) return default_val;
    const gguf_kv * kv = file_.find(key);
    if (!kv || kv->type == GGUF_T_STRING || kv->type == GGUF_T_ARRAY) return default_val;
    return (float) kv->f;
}
size_t GGUFLoader::get_data_offset() const { return is_open_ ? file_.data_offset() : 0; }

bool load_tensor_data_from_file(const GGUFLoader & loader, q3t_ctx * engine, const char * const * prefixes,
                                int n_prefixes, std::string & error_msg, int64_t * n_uploaded) {
    int64_t count = 0;
    const int64_t n = loader.get_n_tensors();
    for (int64_t i = 0; i < n; ++i) {
        const char * name = loader.get_tensor_name(i);
        bool take = false;
        for (int p = 0; p < n_prefixes && !take; ++p) take = strncmp(name, prefixes[p], strlen(prefixes[p])) == 0;
        if (!take) continue;
        int64_t ne[4];
        loader.get_tensor_ne(i, ne);
        const size_t nbytes = loader.get_tensor_size(i);
        if (nbytes == 0) { error_msg = "Failed to read tensor data: " + std::string(name) + " (unsupported dtype)"; return false; }
        const int rc = q3t_upload_tensor(engine, name, loader.get_tensor_type(i), ne, loader.get_tensor_data(i), nbytes);
        if (rc < 0) { error_msg = "Failed to read tensor data: " + std::string(name) + ": " + q3t_last_error(engine); return false; }
        if (rc == 0) ++count;
    }
    if (n_uploaded) *n_uploaded = count;
    return true;
}

} // namespace qwen3_tts

#include "b200_gguf_loader.h"

#include <cstdlib>
#include <cstring>
#include <vector>

namespace qwen3_tts {
namespace b200 {

namespace {
// The reference shares one ggml backend between all components of a process
// (src/gguf_loader.cpp:10-82).  Here an engine context also owns the model state of at most one
// transformer and one vocoder (so generate -> decode can stay on the device), hence a small pool:
// a component takes the first context on its device whose slot of its kind is free.
enum slot_kind { SLOT_TTS = 0, SLOT_DEC = 1, SLOT_OTHER = 2 };
struct engine_entry {
    q3t_ctx * ctx = nullptr;
    int device = 0;
    int32_t refs[3] = {0, 0, 0};
};
std::vector<engine_entry> & engine_pool() {
    static std::vector<engine_entry> pool;
    return pool;
}
slot_kind kind_of(const char * component_name) {
    if (component_name && strstr(component_name, "TTSTransformer")) return SLOT_TTS;
    if (component_name && strstr(component_name, "AudioTokenizerDecoder")) return SLOT_DEC;
    return SLOT_OTHER;
}
}

GGUFLoader::GGUFLoader() = default;
GGUFLoader::~GGUFLoader() { close(); }

q3t_ctx * init_preferred_backend(const char * component_name, std::string * error_msg) {
    if (error_msg) error_msg->clear();
    int device = 0;
    if (const char * e = std::getenv("QWEN3_TTS_DEVICE")) device = std::atoi(e);
    const slot_kind kind = kind_of(component_name);
    auto & pool = engine_pool();
    for (auto & e : pool) {
        if (e.device != device) continue;
        if (kind == SLOT_OTHER || e.refs[kind] == 0) { e.refs[kind]++; return e.ctx; }
    }
    q3t_ctx * ctx = nullptr;
    if (q3t_ctx_create(device, &ctx) != Q3T_OK || !ctx) {
        if (error_msg) {
            const char * name = component_name ? component_name : "component";
            *error_msg = "Failed to initialize backend (CUDA sm_100a) for " + std::string(name) + ": " + q3t_last_error(nullptr);
        }
        return nullptr;
    }
    engine_entry e;
    e.ctx = ctx; e.device = device; e.refs[kind] = 1;
    pool.push_back(e);
    return ctx;
}

void release_preferred_backend(q3t_ctx * backend, const char * component_name) {
    if (!backend) return;
    const slot_kind kind = kind_of(component_name);
    auto & pool = engine_pool();
    for (size_t i = 0; i < pool.size(); ++i) {
        if (pool[i].ctx != backend) continue;
        if (pool[i].refs[kind] > 0) pool[i].refs[kind]--;
        if (pool[i].refs[0] + pool[i].refs[1] + pool[i].refs[2] <= 0) {
            q3t_ctx_destroy(backend);
            pool.erase(pool.begin() + (long) i);
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
    const int64_t v = (int64_t) kv->u;                       // 64-bit keys that do not fit are treated as absent,
    if (v < INT32_MIN || v > (int64_t) UINT32_MAX) return default_val;   // not silently truncated
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

} // namespace b200
} // namespace qwen3_tts

#include "b200_gguf_reader.h"

#include <cstring>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace qwen3_tts {
namespace b200 {

namespace {

struct Cursor {
    const uint8_t * p;
    const uint8_t * end;
    bool ok = true;
    template <typename T> T get() {
        T v{};
        if ((size_t) (end - p) < sizeof(T)) { ok = false; return v; }
        memcpy(&v, p, sizeof(T)); p += sizeof(T);
        return v;
    }
    std::string str() {
        const uint64_t n = get<uint64_t>();
        if (!ok || (uint64_t) (end - p) < n) { ok = false; return {}; }
        std::string s((const char *) p, (size_t) n); p += n;
        return s;
    }
    void skip(uint64_t n) { if ((uint64_t) (end - p) < n) ok = false; else p += n; }
};

size_t scalar_size(uint32_t t) {
    switch (t) {
        case GGUF_T_U8: case GGUF_T_I8: case GGUF_T_BOOL: return 1;
        case GGUF_T_U16: case GGUF_T_I16: return 2;
        case GGUF_T_U32: case GGUF_T_I32: case GGUF_T_F32: return 4;
        case GGUF_T_U64: case GGUF_T_I64: case GGUF_T_F64: return 8;
        default: return 0;
    }
}

bool read_scalar(Cursor & c, uint32_t t, gguf_kv & kv) {
    switch (t) {
        case GGUF_T_U8:   kv.u = c.get<uint8_t>(); break;
        case GGUF_T_I8:   kv.u = (uint64_t) (int64_t) c.get<int8_t>(); break;
        case GGUF_T_U16:  kv.u = c.get<uint16_t>(); break;
        case GGUF_T_I16:  kv.u = (uint64_t) (int64_t) c.get<int16_t>(); break;
        case GGUF_T_U32:  kv.u = c.get<uint32_t>(); break;
        case GGUF_T_I32:  kv.u = (uint64_t) (int64_t) c.get<int32_t>(); break;
        case GGUF_T_F32:  kv.f = c.get<float>(); break;
        case GGUF_T_BOOL: kv.u = c.get<uint8_t>() != 0; break;
        case GGUF_T_U64:  kv.u = c.get<uint64_t>(); break;
        case GGUF_T_I64:  kv.u = (uint64_t) c.get<int64_t>(); break;
        case GGUF_T_F64:  kv.f = c.get<double>(); break;
        default: return false;
    }
    if (t != GGUF_T_F32 && t != GGUF_T_F64) kv.f = (double) (int64_t) kv.u;
    return c.ok;
}

} // namespace

// block size / bytes per block for the ggml types that can appear in these files
// (ggml/include/ggml.h:390-430, ggml/src/ggml-common.h)
size_t GGUFFile::type_nbytes(uint32_t type, const int64_t ne[4]) {
    int64_t blck = 1; size_t bsz = 0;
    switch (type) {
        case 0:  bsz = 4; break;              // F32
        case 1:  bsz = 2; break;              // F16
        case 2:  blck = 32; bsz = 18; break;  // Q4_0
        case 3:  blck = 32; bsz = 20; break;  // Q4_1
        case 6:  blck = 32; bsz = 22; break;  // Q5_0
        case 7:  blck = 32; bsz = 24; break;  // Q5_1
        case 8:  blck = 32; bsz = 34; break;  // Q8_0
        case 9:  blck = 32; bsz = 36; break;  // Q8_1
        case 10: blck = 256; bsz = 84; break;   // Q2_K
        case 11: blck = 256; bsz = 110; break;  // Q3_K
        case 12: blck = 256; bsz = 144; break;  // Q4_K
        case 13: blck = 256; bsz = 176; break;  // Q5_K
        case 14: blck = 256; bsz = 210; break;  // Q6_K
        case 15: blck = 256; bsz = 292; break;  // Q8_K
        case 24: bsz = 1; break;              // I8
        case 25: bsz = 2; break;              // I16
        case 26: bsz = 4; break;              // I32
        case 27: bsz = 8; break;              // I64
        case 28: bsz = 8; break;              // F64
        case 30: bsz = 2; break;              // BF16
        default: return 0;
    }
    // every extent must be positive and the product must fit: a hostile header must not be able to wrap the
    // byte count past the bounds check in open() (0 = "unknown / invalid", which open() rejects)
    for (int d = 0; d < 4; ++d) if (ne[d] <= 0) return 0;
    if (ne[0] % blck != 0) return 0;
    unsigned __int128 n = (unsigned __int128) (uint64_t) (ne[0] / blck) * bsz;
    for (int d = 1; d < 4; ++d) {
        n *= (uint64_t) ne[d];
        if (n > (unsigned __int128) 1 << 62) return 0;
    }
    return (size_t) n;
}

GGUFFile::~GGUFFile() { close(); }

void GGUFFile::close() {
    if (map_) { munmap((void *) map_, size_); map_ = nullptr; }
    if (fd_ >= 0) { ::close(fd_); fd_ = -1; }
    size_ = 0; kv_.clear(); tensors_.clear(); by_name_.clear(); data_offset_ = 0;
    alignment_ = 32; version_ = 0;
}

bool GGUFFile::open(const std::string & path, std::string & err, bool keep_string_arrays) {
    close();
    fd_ = ::open(path.c_str(), O_RDONLY);
    if (fd_ < 0) { err = "Failed to open GGUF file: " + path; return false; }
    struct stat st;
    if (fstat(fd_, &st) != 0 || st.st_size < 24) { err = "Failed to open GGUF file: " + path; close(); return false; }
    size_ = (size_t) st.st_size;
    void * m = mmap(nullptr, size_, PROT_READ, MAP_PRIVATE, fd_, 0);
    if (m == MAP_FAILED) { map_ = nullptr; err = "Failed to mmap GGUF file: " + path; close(); return false; }
    map_ = (const uint8_t *) m;
    madvise(m, size_, MADV_SEQUENTIAL);

    Cursor c{map_, map_ + size_};
    if (memcmp(c.p, "GGUF", 4) != 0) { err = "Failed to open GGUF file: " + path + " (bad magic)"; close(); return false; }
    c.skip(4);
    version_ = c.get<uint32_t>();
    if (version_ < 2 || version_ > 3) { err = "Failed to open GGUF file: " + path + " (unsupported version)"; close(); return false; }
    const uint64_t n_tensors = c.get<uint64_t>();
    const uint64_t n_kv = c.get<uint64_t>();
    if (!c.ok || n_tensors > (1u << 24) || n_kv > (1u << 24)) { err = "Failed to open GGUF file: " + path + " (corrupt header)"; close(); return false; }

    for (uint64_t i = 0; i < n_kv && c.ok; ++i) {
        std::string key = c.str();
        gguf_kv kv;
        kv.type = c.get<uint32_t>();
        if (kv.type == GGUF_T_STRING) {
            kv.s = c.str();
        } else if (kv.type == GGUF_T_ARRAY) {
            kv.arr_type = c.get<uint32_t>();
            kv.arr_n = c.get<uint64_t>();
            if (kv.arr_type == GGUF_T_STRING) {
                for (uint64_t j = 0; j < kv.arr_n && c.ok; ++j) {
                    if (keep_string_arrays) kv.arr_s.push_back(c.str());
                    else { const uint64_t n = c.get<uint64_t>(); c.skip(n); }
                }
            } else {
                const size_t es = scalar_size(kv.arr_type);
                if (es == 0 || kv.arr_n > (uint64_t) (c.end - c.p) / es) { c.ok = false; break; }
                c.skip(es * kv.arr_n);
            }
        } else if (!read_scalar(c, kv.type, kv)) {
            c.ok = false; break;
        }
        kv_[key] = std::move(kv);
    }
    if (!c.ok) { err = "Failed to open GGUF file: " + path + " (corrupt metadata)"; close(); return false; }
    if (const gguf_kv * a = find("general.alignment")) {
        // ggml accepts only powers of two here; anything else (or > 1 MiB) is a corrupt file
        if (a->u == 0 || a->u > (1u << 20) || (a->u & (a->u - 1)) != 0) {
            err = "Failed to open GGUF file: " + path + " (bad general.alignment)"; close(); return false;
        }
        alignment_ = (uint32_t) a->u;
    }

    tensors_.reserve((size_t) n_tensors);
    for (uint64_t i = 0; i < n_tensors && c.ok; ++i) {
        gguf_tensor_info ti;
        ti.name = c.str();
        ti.n_dims = c.get<uint32_t>();
        if (ti.n_dims > 4) { c.ok = false; break; }
        for (uint32_t d = 0; d < ti.n_dims; ++d) ti.ne[d] = (int64_t) c.get<uint64_t>();
        ti.type = c.get<uint32_t>();
        ti.offset = c.get<uint64_t>();
        ti.nbytes = type_nbytes(ti.type, ti.ne);
        if (c.ok && ti.nbytes == 0) {
            err = "Failed to open GGUF file: " + path + " (bad type or shape: " + ti.name + ")"; close(); return false;
        }
        by_name_[ti.name] = tensors_.size();
        tensors_.push_back(std::move(ti));
    }
    if (!c.ok) { err = "Failed to open GGUF file: " + path + " (corrupt tensor directory)"; close(); return false; }
    const size_t pos = (size_t) (c.p - map_);
    data_offset_ = (pos + alignment_ - 1) / alignment_ * alignment_;
    if (data_offset_ > size_) { err = "Failed to open GGUF file: " + path + " (truncated before tensor data)"; close(); return false; }
    const size_t data_size = size_ - data_offset_;
    for (const auto & t : tensors_) {
        // subtraction form: no sum of attacker-controlled 64-bit values that could wrap
        if (t.offset > data_size || t.nbytes > data_size - t.offset) { err = "Failed to open GGUF file: " + path + " (tensor data out of bounds: " + t.name + ")"; close(); return false; }
    }
    return true;
}

const gguf_kv * GGUFFile::find(const std::string & key) const {
    auto it = kv_.find(key);
    return it == kv_.end() ? nullptr : &it->second;
}

int64_t GGUFFile::find_tensor(const std::string & name) const {
    auto it = by_name_.find(name);
    return it == by_name_.end() ? -1 : (int64_t) it->second;
}

} // namespace b200
} // namespace qwen3_tts

// gguf.cpp -- GGUF v2/v3 container parser over an mmap'd file.
//
// Replaces the GGUF-loading stage of the llama-server the reference chart launches (SURVEY 8a row a8;
// launch site ramalama-models/helm-chart/templates/model-deployments.yaml:26-35).  The on-disk layout
// followed here is the one llama.cpp's own gguf-py reads: gguf_reader.py:132-185 (header, KV section,
// tensor infos, padding to general.alignment), :217-255 (strings, arrays, nesting), :259-287 (tensor
// info records); magic/version/default alignment constants.py:10-13.  Every length read from the file
// is bounds-checked against the mapping, so a truncated or hostile file fails with B2_EFORMAT instead
// of faulting.
#include "gguf.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cerrno>
#include <cstring>

namespace b2 {

static thread_local std::string g_err;
void set_error(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
}
const char* get_error() { return g_err.c_str(); }

static const TypeTraits k_f32{1, 4, "F32"}, k_f16{1, 2, "F16"}, k_bf16{1, 2, "BF16"}, k_q8_0{32, 34, "Q8_0"},
    k_q4_K{256, 144, "Q4_K"}, k_q5_K{256, 176, "Q5_K"}, k_q6_K{256, 210, "Q6_K"};
const TypeTraits* type_traits(int type) {
    switch (type) {
        case B2_TYPE_F32: return &k_f32;
        case B2_TYPE_F16: return &k_f16;
        case B2_TYPE_BF16: return &k_bf16;
        case B2_TYPE_Q8_0: return &k_q8_0;
        case B2_TYPE_Q4_K: return &k_q4_K;
        case B2_TYPE_Q5_K: return &k_q5_K;
        case B2_TYPE_Q6_K: return &k_q6_K;
    }
    return nullptr;
}

// sizes of the scalar KV types; 0 for string/array
static size_t scalar_size(uint32_t t) {
    switch (t) {
        case B2_KV_U8: case B2_KV_I8: case B2_KV_BOOL: return 1;
        case B2_KV_U16: case B2_KV_I16: return 2;
        case B2_KV_U32: case B2_KV_I32: case B2_KV_F32: return 4;
        case B2_KV_U64: case B2_KV_I64: case B2_KV_F64: return 8;
    }
    return 0;
}

// generic block sizes for tensor types we do not compute with but must still skip over correctly
// (gguf-py constants.py:4231-4267): {type id, block elems, block bytes}
static const int k_all_types[][3] = {
    {0, 1, 4}, {1, 1, 2}, {2, 32, 18}, {3, 32, 20}, {6, 32, 22}, {7, 32, 24}, {8, 32, 34}, {9, 32, 40},
    {10, 256, 84}, {11, 256, 110}, {12, 256, 144}, {13, 256, 176}, {14, 256, 210}, {15, 256, 292},
    {16, 256, 66}, {17, 256, 74}, {18, 256, 98}, {19, 256, 50}, {20, 32, 18}, {21, 256, 110},
    {22, 256, 82}, {23, 256, 136}, {24, 1, 1}, {25, 1, 2}, {26, 1, 4}, {27, 1, 8}, {28, 1, 8},
    {29, 256, 56}, {30, 1, 2}, {34, 256, 54}, {35, 256, 66}, {39, 32, 17}};

static bool generic_block(uint32_t type, uint64_t* be, uint64_t* bb) {
    for (auto& t : k_all_types)
        if ((uint32_t)t[0] == type) {
            *be = t[1];
            *bb = t[2];
            return true;
        }
    return false;
}

struct Cursor {
    const uint8_t* base;
    uint64_t size, pos;
    bool ok = true;
    bool need(uint64_t n) {
        if (!ok || n > size || pos > size - n) {
            ok = false;
            return false;
        }
        return true;
    }
    template <typename T>
    T get() {
        T v{};
        if (need(sizeof(T))) {
            memcpy(&v, base + pos, sizeof(T));
            pos += sizeof(T);
        }
        return v;
    }
    bool str(const char** s, uint64_t* len) {
        uint64_t n = get<uint64_t>();
        if (!need(n)) return false;
        *s = (const char*)(base + pos);
        *len = n;
        pos += n;
        return true;
    }
};

// advance over one value of type t (recursing into arrays); records nothing
static bool skip_value(Cursor& c, uint32_t t, int depth) {
    if (depth > 8) return c.ok = false;
    if (size_t s = scalar_size(t)) {
        if (!c.need(s)) return false;
        c.pos += s;
        return true;
    }
    if (t == B2_KV_STRING) {
        const char* s;
        uint64_t n;
        return c.str(&s, &n);
    }
    if (t == B2_KV_ARRAY) {
        uint32_t et = c.get<uint32_t>();
        uint64_t n = c.get<uint64_t>();
        if (!c.ok) return false;
        if (size_t s = scalar_size(et)) {
            if (n > c.size / s || !c.need(n * s)) return c.ok = false;
            c.pos += n * s;
            return true;
        }
        for (uint64_t i = 0; i < n; i++)
            if (!skip_value(c, et, depth + 1)) return false;
        return true;
    }
    return c.ok = false;
}

Gguf::~Gguf() {
    if (map && map != MAP_FAILED) munmap((void*)map, map_size);
    if (fd >= 0) close(fd);
}

int Gguf::open(const char* path, Gguf** out) {
    *out = nullptr;
    std::unique_ptr<Gguf> g(new Gguf());
    g->fd = ::open(path, O_RDONLY);
    if (g->fd < 0) {
        set_error("cannot open '%s': %s", path, strerror(errno));
        return B2_EIO;
    }
    struct stat st;
    if (fstat(g->fd, &st) != 0 || st.st_size < 24) {
        set_error("'%s': not a GGUF file (too small)", path);
        return B2_EFORMAT;
    }
    g->map_size = (uint64_t)st.st_size;
    g->map = (const uint8_t*)mmap(nullptr, g->map_size, PROT_READ, MAP_PRIVATE, g->fd, 0);
    if (g->map == MAP_FAILED) {
        g->map = nullptr;
        set_error("mmap('%s') failed: %s", path, strerror(errno));
        return B2_EIO;
    }
    g->path = path;
    Cursor c{g->map, g->map_size, 0};
    if (c.get<uint32_t>() != 0x46554747u) {
        set_error("'%s': bad magic, not a GGUF file", path);
        return B2_EFORMAT;
    }
    g->version = c.get<uint32_t>();
    if (g->version != 2 && g->version != 3) {
        // a byte-swapped version word means a big-endian file (gguf_reader.py:141-148)
        set_error("'%s': unsupported GGUF version %u (only little-endian v2/v3)", path, g->version);
        return B2_EUNSUP;
    }
    uint64_t n_tensors = c.get<uint64_t>(), n_kv = c.get<uint64_t>();
    if (!c.ok || n_tensors > (1u << 24) || n_kv > (1u << 24)) {
        set_error("'%s': implausible header counts", path);
        return B2_EFORMAT;
    }
    g->kvs.reserve(n_kv);
    for (uint64_t i = 0; i < n_kv; i++) {
        KV kv;
        const char* ks;
        uint64_t kl;
        if (!c.str(&ks, &kl)) break;
        kv.key.assign(ks, kl);
        kv.type = c.get<uint32_t>();
        kv.value_pos = c.pos;
        if (kv.type == B2_KV_ARRAY) {
            Cursor p = c;
            kv.arr_type = p.get<uint32_t>();
            kv.arr_len = p.get<uint64_t>();
            kv.value_pos = p.pos;  // first element
        }
        if (!skip_value(c, kv.type, 0)) break;
        // index variable-size array elements lazily (strings): remember nothing here
        g->kv_index[kv.key] = (int64_t)g->kvs.size();
        g->kvs.push_back(std::move(kv));
    }
    if (!c.ok || g->kvs.size() != n_kv) {
        set_error("'%s': truncated or malformed key-value section", path);
        return B2_EFORMAT;
    }
    g->alignment = 32;
    auto it = g->kv_index.find("general.alignment");
    if (it != g->kv_index.end()) {
        uint64_t a = 0;
        if (g->kv_u64(it->second, &a) == 0 && a && (a & (a - 1)) == 0 && a <= (1u << 20)) g->alignment = (uint32_t)a;
        else {
            set_error("'%s': general.alignment is not a power of two", path);
            return B2_EFORMAT;
        }
    }
    g->tensors.reserve(n_tensors);
    for (uint64_t i = 0; i < n_tensors; i++) {
        Tensor t;
        const char* ns;
        uint64_t nl;
        if (!c.str(&ns, &nl)) break;
        t.name.assign(ns, nl);
        t.n_dims = c.get<uint32_t>();
        if (t.n_dims > 4) {
            c.ok = false;
            break;
        }
        uint64_t n_el = 1;
        for (uint32_t d = 0; d < 4; d++) t.dims[d] = 1;
        for (uint32_t d = 0; d < t.n_dims; d++) {
            t.dims[d] = c.get<uint64_t>();
            if (t.dims[d] == 0 || n_el > UINT64_MAX / t.dims[d]) c.ok = false;
            else n_el *= t.dims[d];
        }
        t.type = c.get<uint32_t>();
        t.offset = c.get<uint64_t>();
        if (!c.ok) break;
        uint64_t be, bb;
        if (!generic_block(t.type, &be, &bb)) {
            set_error("'%s': tensor '%s' has unknown ggml type %u", path, t.name.c_str(), t.type);
            return B2_EUNSUP;
        }
        if (t.dims[0] % be) {
            set_error("'%s': tensor '%s' row length %llu is not a multiple of its block size %llu", path,
                      t.name.c_str(), (unsigned long long)t.dims[0], (unsigned long long)be);
            return B2_EFORMAT;
        }
        t.n_bytes = n_el / be * bb;
        g->tensor_index[t.name] = (int64_t)g->tensors.size();
        g->tensors.push_back(std::move(t));
    }
    if (!c.ok || g->tensors.size() != n_tensors) {
        set_error("'%s': truncated or malformed tensor-info section", path);
        return B2_EFORMAT;
    }
    const uint64_t a = g->alignment;
    g->data_offset = (c.pos + a - 1) / a * a;
    for (auto& t : g->tensors) {
        if (t.offset % a) {
            set_error("'%s': tensor '%s' offset is not aligned to %u", path, t.name.c_str(), g->alignment);
            return B2_EFORMAT;
        }
        if (g->data_offset > g->map_size || t.offset > g->map_size - g->data_offset ||
            t.n_bytes > g->map_size - g->data_offset - t.offset) {
            set_error("'%s': tensor '%s' extends past the end of the file", path, t.name.c_str());
            return B2_EFORMAT;
        }
    }
    g->data_size = g->map_size - g->data_offset;
    *out = g.release();
    return 0;
}

int Gguf::kv_u64(int64_t i, uint64_t* out) const {
    int64_t v;
    int r = kv_i64(i, &v);
    *out = (uint64_t)v;
    return r;
}

static int scalar_to_i64(const uint8_t* p, uint32_t type, int64_t* out) {
    switch (type) {
        case B2_KV_U8: *out = *(const uint8_t*)p; return 0;
        case B2_KV_I8: *out = *(const int8_t*)p; return 0;
        case B2_KV_BOOL: *out = *(const uint8_t*)p != 0; return 0;
        case B2_KV_U16: { uint16_t v; memcpy(&v, p, 2); *out = v; return 0; }
        case B2_KV_I16: { int16_t v; memcpy(&v, p, 2); *out = v; return 0; }
        case B2_KV_U32: { uint32_t v; memcpy(&v, p, 4); *out = v; return 0; }
        case B2_KV_I32: { int32_t v; memcpy(&v, p, 4); *out = v; return 0; }
        case B2_KV_U64: { uint64_t v; memcpy(&v, p, 8); *out = (int64_t)v; return 0; }
        case B2_KV_I64: { int64_t v; memcpy(&v, p, 8); *out = v; return 0; }
        case B2_KV_F32: { float v; memcpy(&v, p, 4); *out = (int64_t)v; return 0; }
        case B2_KV_F64: { double v; memcpy(&v, p, 8); *out = (int64_t)v; return 0; }
    }
    return B2_EINVAL;
}
static int scalar_to_f64(const uint8_t* p, uint32_t type, double* out) {
    if (type == B2_KV_F32) { float v; memcpy(&v, p, 4); *out = v; return 0; }
    if (type == B2_KV_F64) { memcpy(out, p, 8); return 0; }
    int64_t v;
    int r = scalar_to_i64(p, type, &v);
    if (type == B2_KV_U64) { uint64_t u; memcpy(&u, p, 8); *out = (double)u; return 0; }
    *out = (double)v;
    return r;
}

int Gguf::kv_i64(int64_t i, int64_t* out) const {
    if (i < 0 || i >= (int64_t)kvs.size()) return B2_EINVAL;
    return scalar_to_i64(map + kvs[i].value_pos, kvs[i].type, out);
}
int Gguf::kv_f64(int64_t i, double* out) const {
    if (i < 0 || i >= (int64_t)kvs.size()) return B2_EINVAL;
    return scalar_to_f64(map + kvs[i].value_pos, kvs[i].type, out);
}
int Gguf::kv_str(int64_t i, const char** s, uint64_t* len) const {
    if (i < 0 || i >= (int64_t)kvs.size() || kvs[i].type != B2_KV_STRING) return B2_EINVAL;
    Cursor c{map, map_size, kvs[i].value_pos};
    return c.str(s, len) ? 0 : B2_EFORMAT;
}

// offsets of string-array elements are built on first use
const std::vector<uint64_t>& Gguf::str_offsets(int64_t i) const {
    std::lock_guard<std::mutex> lk(mu);
    auto& v = str_off_cache[i];
    if (v.empty() && kvs[i].arr_len) {
        v.reserve(kvs[i].arr_len);
        Cursor c{map, map_size, kvs[i].value_pos};
        for (uint64_t j = 0; j < kvs[i].arr_len; j++) {
            v.push_back(c.pos);
            const char* s;
            uint64_t n;
            if (!c.str(&s, &n)) break;
        }
    }
    return v;
}

int Gguf::arr_str(int64_t i, int64_t j, const char** s, uint64_t* len) const {
    if (i < 0 || i >= (int64_t)kvs.size() || kvs[i].type != B2_KV_ARRAY || kvs[i].arr_type != B2_KV_STRING) return B2_EINVAL;
    const auto& off = str_offsets(i);
    if (j < 0 || j >= (int64_t)off.size()) return B2_ERANGE;
    Cursor c{map, map_size, off[j]};
    return c.str(s, len) ? 0 : B2_EFORMAT;
}
int Gguf::arr_f64(int64_t i, int64_t j, double* out) const {
    if (i < 0 || i >= (int64_t)kvs.size() || kvs[i].type != B2_KV_ARRAY) return B2_EINVAL;
    size_t s = scalar_size(kvs[i].arr_type);
    if (!s) return B2_EINVAL;
    if (j < 0 || (uint64_t)j >= kvs[i].arr_len) return B2_ERANGE;
    return scalar_to_f64(map + kvs[i].value_pos + (uint64_t)j * s, kvs[i].arr_type, out);
}
int Gguf::arr_i64(int64_t i, int64_t j, int64_t* out) const {
    if (i < 0 || i >= (int64_t)kvs.size() || kvs[i].type != B2_KV_ARRAY) return B2_EINVAL;
    size_t s = scalar_size(kvs[i].arr_type);
    if (!s) return B2_EINVAL;
    if (j < 0 || (uint64_t)j >= kvs[i].arr_len) return B2_ERANGE;
    return scalar_to_i64(map + kvs[i].value_pos + (uint64_t)j * s, kvs[i].arr_type, out);
}

int64_t Gguf::find_kv(const std::string& key) const {
    auto it = kv_index.find(key);
    return it == kv_index.end() ? -1 : it->second;
}
int64_t Gguf::find_tensor(const std::string& name) const {
    auto it = tensor_index.find(name);
    return it == tensor_index.end() ? -1 : it->second;
}

bool Gguf::get_u32(const std::string& key, uint32_t* out) const {
    int64_t i = find_kv(key);
    int64_t v;
    if (i < 0 || kv_i64(i, &v) != 0) return false;
    *out = (uint32_t)v;
    return true;
}
bool Gguf::get_f32(const std::string& key, float* out) const {
    int64_t i = find_kv(key);
    double v;
    if (i < 0 || kv_f64(i, &v) != 0) return false;
    *out = (float)v;
    return true;
}
bool Gguf::get_str(const std::string& key, std::string* out) const {
    int64_t i = find_kv(key);
    const char* s;
    uint64_t n;
    if (i < 0 || kv_str(i, &s, &n) != 0) return false;
    out->assign(s, n);
    return true;
}
bool Gguf::get_bool(const std::string& key, bool* out) const {
    int64_t i = find_kv(key);
    int64_t v;
    if (i < 0 || kv_i64(i, &v) != 0) return false;
    *out = v != 0;
    return true;
}

}  // namespace b2

// gguf.h -- parsed GGUF container (see gguf.cpp)
#pragma once
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "common.h"

namespace b2 {

struct Gguf {
    struct KV {
        std::string key;
        uint32_t type = 0;
        uint32_t arr_type = 0;
        uint64_t arr_len = 0;
        uint64_t value_pos = 0;  // file offset of the value (first element for arrays)
    };
    struct Tensor {
        std::string name;
        uint32_t n_dims = 0;
        uint64_t dims[4] = {1, 1, 1, 1};
        uint32_t type = 0;
        uint64_t offset = 0, n_bytes = 0;
    };

    std::string path;
    int fd = -1;
    const uint8_t* map = nullptr;
    uint64_t map_size = 0;
    uint32_t version = 0, alignment = 32;
    uint64_t data_offset = 0, data_size = 0;
    std::vector<KV> kvs;
    std::vector<Tensor> tensors;
    std::unordered_map<std::string, int64_t> kv_index, tensor_index;

    ~Gguf();
    static int open(const char* path, Gguf** out);

    int64_t find_kv(const std::string& key) const;
    int64_t find_tensor(const std::string& name) const;
    int kv_u64(int64_t i, uint64_t* out) const;
    int kv_i64(int64_t i, int64_t* out) const;
    int kv_f64(int64_t i, double* out) const;
    int kv_str(int64_t i, const char** s, uint64_t* len) const;
    int arr_str(int64_t i, int64_t j, const char** s, uint64_t* len) const;
    int arr_f64(int64_t i, int64_t j, double* out) const;
    int arr_i64(int64_t i, int64_t j, int64_t* out) const;
    // convenience getters; false when the key is absent or of the wrong kind
    bool get_u32(const std::string& key, uint32_t* out) const;
    bool get_f32(const std::string& key, float* out) const;
    bool get_str(const std::string& key, std::string* out) const;
    bool get_bool(const std::string& key, bool* out) const;
    const uint8_t* tensor_data(int64_t i) const { return map + data_offset + tensors[i].offset; }

   private:
    const std::vector<uint64_t>& str_offsets(int64_t i) const;
    mutable std::mutex mu;
    mutable std::map<int64_t, std::vector<uint64_t>> str_off_cache;
};

}  // namespace b2

struct b2_gguf {
    b2::Gguf* g = nullptr;
    bool owned = false;   // false for the view a b2_model hands out
};

// common.h -- error slot, type table and CUDA error macro shared by the host-side sources.
#pragma once
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <mutex>
#include <string>

#include "../../include/b2.h"

namespace b2 {

// thread-local last-error message behind b2_last_error()
void set_error(const char* fmt, ...) __attribute__((format(printf, 1, 2)));
const char* get_error();

struct TypeTraits {
    int block_elems;   // elements per block
    int block_bytes;   // bytes per block
    const char* name;
};
// nullptr for types this build does not handle
const TypeTraits* type_traits(int type);

inline int64_t row_bytes(int type, int64_t k) {
    const TypeTraits* t = type_traits(type);
    if (!t || k % t->block_elems) return -1;
    return k / t->block_elems * t->block_bytes;
}

// One process may drive several GPUs (b2_model_load takes a device ordinal), and what is set up once -- a kernel's
// opt-in dynamic shared-memory size (cudaFuncSetAttribute), the SM count -- is per DEVICE: this keeps the largest
// value configured so far for each device ordinal.
constexpr int MAX_DEVICES = 64;
struct PerDeviceSize {
    std::mutex mu;
    size_t v[MAX_DEVICES] = {};
    // true if device `dev` is not yet configured for `want` (an ordinal outside the table is configured every time)
    bool below(int dev, size_t want) {
        std::lock_guard<std::mutex> l(mu);
        return dev < 0 || dev >= MAX_DEVICES || v[dev] < want;
    }
    void set(int dev, size_t want) {
        std::lock_guard<std::mutex> l(mu);
        if (dev >= 0 && dev < MAX_DEVICES && v[dev] < want) v[dev] = want;
    }
};
// SMs of the current device (looked up once per device); 0 and the error slot set on failure
int device_sm_count();
int current_device();   // ordinal of the current device, -1 on failure

}  // namespace b2

#define B2_CUDA_CHECK(expr)                                                                      \
    do {                                                                                         \
        cudaError_t _e = (expr);                                                                 \
        if (_e != cudaSuccess) {                                                                 \
            b2::set_error("CUDA error %s at %s:%d: %s", cudaGetErrorName(_e), __FILE__, __LINE__, \
                          cudaGetErrorString(_e));                                               \
            return B2_ECUDA;                                                                     \
        }                                                                                        \
    } while (0)

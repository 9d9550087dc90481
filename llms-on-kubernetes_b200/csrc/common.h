// common.h -- error slot, type table and CUDA error macro shared by the host-side sources.
#pragma once
#include <cstdarg>
#include <cstdint>
#include <cstdio>
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

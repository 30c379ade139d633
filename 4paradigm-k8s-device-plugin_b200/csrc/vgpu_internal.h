// Shared between the translation units of libvgpu_dp.so (not part of the C-ABI).
#pragma once
#include <cstdarg>
#include <cstdint>
#include <cstdio>

#include "vgpu.h"

namespace vgpu {

// Thread-local message behind vgpu_last_error().
char *err_buf();
constexpr size_t kErrLen = 512;

inline int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err_buf(), kErrLen, fmt, ap);
    va_end(ap);
    return code;
}

}  // namespace vgpu

#define VGPU_CUDA(call)                                                                  \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess)                                                           \
            return vgpu::fail(VGPU_E_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #call,  \
                              cudaGetErrorString(e_));                                   \
    } while (0)

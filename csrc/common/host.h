// b200ml — host-side helpers shared by every launcher: error reporting through the C ABI,
// cuTensorMapEncodeTiled resolved at run time (no link-time libcuda dependency), SM count cache.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

namespace b200 {

// Last error text, readable through b200ml_last_error().
char* last_error_buf();
int set_error(const char* fmt, ...);

#define B200_CHECK_CUDA(expr)                                                                    \
    do {                                                                                         \
        cudaError_t _e = (expr);                                                                 \
        if (_e != cudaSuccess)                                                                   \
            return ::b200::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
    } while (0)

#define B200_REQUIRE(cond, ...)                       \
    do {                                              \
        if (!(cond)) return ::b200::set_error(__VA_ARGS__); \
    } while (0)

int sm_count();  // of the current device (cached per device)

// 2-D / 3-D bf16 tensor map with a 128B-swizzled box whose inner extent is 64 elements (128 bytes).
// dims/strides are innermost-first; strides[0] is implied (2 bytes).  Returns 0 on success.
int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                   const uint32_t* box);

}  // namespace b200

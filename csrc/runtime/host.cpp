// b200ml — host runtime glue: error buffer, driver entry-point lookup for TMA descriptors, device info.
#include "../common/host.h"

#include <stdarg.h>

#include <mutex>

#include "../../include/b200ml.h"

namespace b200 {

static thread_local char g_err[1024] = {0};
char* last_error_buf() { return g_err; }

int set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return -1;
}

int sm_count() {
    static int cache[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cache[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cache[dev] = n;
    }
    return cache[dev];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                   const uint32_t* box) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return set_error("cuTensorMapEncodeTiled not available from the driver");
    // Driver-API entry points need a current context on the calling thread.  Runtime calls bind the primary context
    // implicitly, driver calls do not — and autograd worker threads may not have issued a runtime call yet.
    static thread_local bool ctx_bound = false;
    if (!ctx_bound) {
        cudaFree(nullptr);
        ctx_bound = true;
    }
    if ((reinterpret_cast<uintptr_t>(base) & 15) != 0) return set_error("TMA base %p not 16-byte aligned", base);
    cuuint64_t gdim[5];
    cuuint64_t gstr[5];
    cuuint32_t gbox[5];
    cuuint32_t estr[5];
    for (int i = 0; i < rank; ++i) {
        gdim[i] = dims[i];
        gbox[i] = box[i];
        estr[i] = 1;
        if (i > 0) {
            gstr[i - 1] = strides_bytes[i];
            if (strides_bytes[i] % 16 != 0)
                return set_error("TMA stride[%d]=%llu bytes is not a multiple of 16", i,
                                 (unsigned long long)strides_bytes[i]);
        }
    }
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, gbox,
                    estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return 0;
}

}  // namespace b200

extern "C" {

const char* b200ml_last_error(void) { return b200::last_error_buf(); }

int b200ml_version(void) { return 100; }

int b200ml_device_info(int* sm_count, int* cc_major, int* cc_minor, size_t* smem_optin) {
    int dev = 0;
    B200_CHECK_CUDA(cudaGetDevice(&dev));
    cudaDeviceProp p;
    B200_CHECK_CUDA(cudaGetDeviceProperties(&p, dev));
    if (sm_count) *sm_count = p.multiProcessorCount;
    if (cc_major) *cc_major = p.major;
    if (cc_minor) *cc_minor = p.minor;
    if (smem_optin) *smem_optin = p.sharedMemPerBlockOptin;
    return 0;
}

}  // extern "C"

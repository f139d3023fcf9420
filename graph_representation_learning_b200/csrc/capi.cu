// Error plumbing and library-level entry points of the C ABI (include/grl_b200.h).
#include "common.cuh"

namespace grl {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
    set_error("%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
    return static_cast<int>(e);
}

static const uint64_t* g_rng_offset = nullptr;
const uint64_t* rng_offset_ptr() { return g_rng_offset; }

}  // namespace grl

extern "C" int grl_rng_set_offset_ptr(const void* dev_u64) {
    grl::g_rng_offset = static_cast<const uint64_t*>(dev_u64);
    return 0;
}

extern "C" int grl_abi_version(void) { return GRL_ABI_VERSION; }

extern "C" const char* grl_last_error(void) { return grl::g_err; }

extern "C" int grl_device_ok(void) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        grl::cuda_fail(e, "cudaGetDevice");
        return -static_cast<int>(e);
    }
    int major = 0;
    e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    if (e != cudaSuccess) {
        grl::cuda_fail(e, "cudaDeviceGetAttribute");
        return -static_cast<int>(e);
    }
    return major == 10 ? 1 : 0;
}

// Library-wide plumbing: thread-local error text, launch counter, device query.
#include "common.cuh"
#include <stdarg.h>
#include <string.h>

namespace penroz {

static thread_local char tl_error[512] = "";
std::atomic<unsigned long long> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(tl_error, sizeof(tl_error), fmt, ap);
    va_end(ap);
}

int check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        return PENROZ_ERR_CUDA;
    }
    return PENROZ_OK;
}

int num_sms() {
    static std::atomic<int> cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    int v = cached[dev].load(std::memory_order_relaxed);
    if (v > 0) return v;
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev].store(n, std::memory_order_relaxed);
    return n;
}

}  // namespace penroz

using namespace penroz;

extern "C" const char* penroz_last_error(void) { return tl_error; }

extern "C" unsigned long long penroz_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

extern "C" int penroz_abi_version(void) { return 4; }

// Refuses anything but a compute-capability 10.x device: the kernels are sm_100a-only SASS.
extern "C" int penroz_query(int* cc_major, int* cc_minor, int* sm_count) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { set_error("penroz_query: %s", cudaGetErrorString(e)); return PENROZ_ERR_CUDA; }
    int maj = 0, min = 0, sms = 0;
    cudaDeviceGetAttribute(&maj, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(&min, cudaDevAttrComputeCapabilityMinor, dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cc_major) *cc_major = maj;
    if (cc_minor) *cc_minor = min;
    if (sm_count) *sm_count = sms;
    if (maj != 10) {
        set_error("penroz_b200 needs an sm_100a (B200) device, found compute capability %d.%d", maj, min);
        return PENROZ_ERR_ARCH;
    }
    return PENROZ_OK;
}

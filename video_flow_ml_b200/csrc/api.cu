// Library-level entry points of libvfml: version, per-thread last error, launch counter.
#include <stdarg.h>

#include "common.cuh"

namespace vfml {

static thread_local char g_err[512] = "";
std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int device_index() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return dev;
}

int device_sms() {
    static std::atomic<int> cache[64] = {};
    const int dev = device_index() & 63;
    int n = cache[dev].load(std::memory_order_relaxed);
    if (n <= 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) {
            cudaGetLastError();
            n = 148;          // B200
        }
        cache[dev].store(n, std::memory_order_relaxed);
    }
    return n;
}

}  // namespace vfml

extern "C" int vfml_version(void) { return VFML_VERSION; }
extern "C" const char* vfml_last_error(void) { return vfml::g_err; }
extern "C" uint64_t vfml_launch_count(void) { return vfml::g_launches.load(); }
extern "C" int vfml_device_sms(void) { return vfml::device_sms(); }

namespace vfml {
bool gemm_tc_supported(const vfml_gemm_desc& d);
int validate_gemm(const vfml_gemm_desc* d);
}  // namespace vfml

extern "C" int vfml_gemm_uses_tensor_cores(const vfml_gemm_desc* d) {
    if (vfml::validate_gemm(d)) return 0;
    return vfml::gemm_tc_supported(*d) ? 1 : 0;
}

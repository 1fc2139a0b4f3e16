// Shared helpers for libvfml (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>

#include "../../include/vfml.h"

namespace vfml {

void set_error(const char* fmt, ...);
extern std::atomic<uint64_t> g_launches;

inline cudaStream_t to_stream(vfml_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

#define VFML_REQUIRE(cond, ...)               \
    do {                                      \
        if (!(cond)) {                        \
            vfml::set_error(__VA_ARGS__);     \
            return VFML_ERR_ARG;              \
        }                                     \
    } while (0)

// call after every kernel launch
#define VFML_LAUNCH_CHECK(name)                                                      \
    do {                                                                             \
        vfml::g_launches.fetch_add(1, std::memory_order_relaxed);                    \
        cudaError_t e__ = cudaPeekAtLastError();                                     \
        if (e__ != cudaSuccess) {                                                    \
            vfml::set_error("%s: launch failed: %s", name, cudaGetErrorString(e__)); \
            cudaGetLastError();                                                      \
            return VFML_ERR_CUDA;                                                    \
        }                                                                            \
    } while (0)

#define VFML_CUDA(call)                                                          \
    do {                                                                         \
        cudaError_t e__ = (call);                                                \
        if (e__ != cudaSuccess) {                                                \
            vfml::set_error("%s failed: %s", #call, cudaGetErrorString(e__));    \
            return VFML_ERR_CUDA;                                                \
        }                                                                        \
    } while (0)

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }

template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }

// exact (erf) GELU, as torch.nn.functional.gelu default
__device__ __forceinline__ float gelu_erf(float x) {
    return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}

// SM count of the calling thread's current device (cached per device; 148 on B200) and the device's index
int device_sms();
int device_index();
// Per-device one-time setup (cudaFuncSetAttribute is per device): a static `PerDevice<bool> done;` reads as false the
// first time each device gets here.
template <typename T> struct PerDevice {
    T v[64] = {};
    T& get() { return v[device_index() & 63]; }
};

inline int cdiv(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

inline bool valid_dtype(int d) { return d == VFML_F32 || d == VFML_F16; }
inline size_t dtype_size(int d) { return d == VFML_F16 ? 2 : 4; }

}  // namespace vfml

// Micro-benchmark: DRAM fetch granularity of random 32-byte-sector gathers on B200, per load flavour and
// per cudaLimitMaxL2FetchGranularity.  Each lane pair reads one random, distinct 32-byte sector (16 B per lane)
// of a 2 GiB buffer; useful GB/s = sectors * 32 B / time.  If the memory system fetches 64 or 128 bytes per
// miss, the useful rate is 1/2 or 1/4 of what a sector-granular fetch reaches.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/sector_gather scripts/micro/sector_gather.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>
__device__ __forceinline__ uint4 ld16(const void* p) {
    uint4 v;
    if (MODE == 0) asm volatile("ld.global.ca.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    if (MODE == 1) asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    if (MODE == 2) asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    if (MODE == 3) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    if (MODE == 4) asm volatile("ld.global.cs.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    if (MODE == 5) asm volatile("ld.global.cv.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

// unit = contiguous bytes per gather target (32, 64, 128): lanes of a group of unit/16 read one random unit
template <int MODE>
__global__ void gather(const uint8_t* buf, uint32_t n_units_log2, int unit, uint32_t* sink, int iters) {
    const uint32_t lanes_per_unit = unit / 16;
    const uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    const uint64_t nthreads = (uint64_t)gridDim.x * blockDim.x;
    uint32_t acc = 0;
    const uint32_t mask = (1u << n_units_log2) - 1;
    for (int it = 0; it < iters; ++it) {
        const uint64_t g = (t + it * nthreads) / lanes_per_unit;              // unit ordinal
        const uint32_t u = (uint32_t)(g * 2654435761ull + 12345u) & mask;     // bijective scatter (odd multiplier)
        const uint4 v = ld16<MODE>(buf + (uint64_t)u * unit + (t % lanes_per_unit) * 16);
        acc ^= v.x ^ v.y ^ v.z ^ v.w;
    }
    if (acc == 0x12345678u) sink[0] = acc;
}

template <int MODE>
float run(const uint8_t* buf, size_t bytes, int unit, uint32_t* sink) {
    uint32_t lg = 0;
    while (((size_t)unit << (lg + 1)) <= bytes) ++lg;
    const int blocks = 148 * 16, threads = 256, iters = 16;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    gather<MODE><<<blocks, threads>>>(buf, lg, unit, sink, iters);
    cudaEventRecord(e0);
    gather<MODE><<<blocks, threads>>>(buf, lg, unit, sink, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double useful = (double)blocks * threads * iters * 16.0;
    return (float)(useful / (ms * 1e-3) / 1e9);
}

int main() {
    const size_t bytes = 2ull << 30;
    uint8_t* buf; uint32_t* sink;
    cudaMalloc(&buf, bytes); cudaMalloc(&sink, 4);
    cudaMemset(buf, 1, bytes);
    const char* names[] = {"ld.global.ca", "ld.global.cg", "ld.global.nc", "ld.global.nc.L1::no_allocate", "ld.global.cs", "ld.global.cv"};
    for (int gran : {0, 32, 64, 128}) {
        if (gran) {
            cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, gran);
            size_t got = 0; cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity);
            printf("cudaLimitMaxL2FetchGranularity := %d (%s), now %zu\n", gran, cudaGetErrorString(e), got);
        } else {
            size_t got = 0; cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity);
            printf("default cudaLimitMaxL2FetchGranularity = %zu\n", got);
        }
        for (int unit : {32, 64, 128}) {
            float r[6] = {run<0>(buf, bytes, unit, sink), run<1>(buf, bytes, unit, sink), run<2>(buf, bytes, unit, sink),
                          run<3>(buf, bytes, unit, sink), run<4>(buf, bytes, unit, sink), run<5>(buf, bytes, unit, sink)};
            for (int m = 0; m < 6; ++m) printf("  unit %3d B  %-30s %8.1f GB/s useful\n", unit, names[m], r[m]);
        }
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}

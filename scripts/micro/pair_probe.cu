// CTA-pair (cta_group::2) tcgen05.mma probe: semantics check + issue rate, before building a kernel on it.
// Cluster of 2 CTAs.  CTA c holds rows 128c .. 128c+127 of A [256 x 64] and rows (N/2)c .. of B [N x 64] in its own
// shared memory (K-major, 128-byte swizzle, same offsets in both CTAs); the LEADER (rank 0) issues
// tcgen05.mma.cta_group::2 (M = 256, N, K = 16) x 4 and a multicast commit; each CTA then reads ITS 128 accumulator rows
// from its own TMEM.  Part 1 checks D = A B^T against the host; part 2 times back-to-back MMAs.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/pair_probe scripts/micro/pair_probe.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_sw128_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
__device__ __forceinline__ void umma2_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n}\n"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}
__device__ __forceinline__ void commit2(uint64_t* bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile("{\n.reg .pred p;\nWL:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra WD;\nbra WL;\nWD:\n}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n.reg .pred P;\nelect.sync _|P, 0xFFFFFFFF;\nselp.b32 %0, 1, 0, P;\n}\n" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__host__ __device__ inline float a_val(int r, int k) { return (float)((r * 3 + k * 5) % 7 - 3); }
__host__ __device__ inline float b_val(int n, int k) { return (float)((n * 5 + k * 3) % 5 - 2); }

// element (row, k) of a [rows x 64] fp16 block, K-major, 128-byte swizzle
__device__ __forceinline__ uint32_t sw128_off(int row, int k) { return (uint32_t)row * 128u + ((((uint32_t)k >> 3) ^ ((uint32_t)row & 7u)) << 4) + ((uint32_t)k & 7u) * 2u; }

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(192, 1) probe(int N, int iters, float* D, long long* cycles) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t done_bar;
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t rank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    const int half_n = N / 2;
    uint8_t* sA = smem;                        // [128 x 64]
    uint8_t* sB = smem + 128 * 128;            // [N/2 x 64]
    for (int i = threadIdx.x; i < 128 * 64; i += blockDim.x) {
        const int r = i >> 6, k = i & 63;
        *reinterpret_cast<__half*>(sA + sw128_off(r, k)) = __float2half(a_val(128 * (int)rank + r, k));
    }
    for (int i = threadIdx.x; i < half_n * 64; i += blockDim.x) {
        const int n = i >> 6, k = i & 63;
        *reinterpret_cast<__half*>(sB + sw128_off(n, k)) = __float2half(b_val(half_n * (int)rank + n, k));
    }
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&done_bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync_all();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base_s;
    long long t0 = 0, t1 = 0;
    if (warp == 1) {
        if (rank == 0) {
            // instruction descriptor: D fp32, A / B fp16 K-major, N, M = 256
            const uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
            const uint64_t ad = make_sw128_desc(smem_u32(sA)), bd = make_sw128_desc(smem_u32(sB));
            t0 = clock64();
            for (int it = 0; it < iters; ++it) {
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma2_f16(tmem, ad + k * 2, bd + k * 2, idesc, (it | k) ? 1u : 0u);
                }
                __syncwarp();
            }
            if (elect_one()) commit2(&done_bar, 3);
            __syncwarp();
        }
        mbar_wait(&done_bar, 0);           // both CTAs: the leader's multicast commit arrives here
        t1 = clock64();
        if (rank == 0 && lane == 0) cycles[blockIdx.x / 2] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (warp >= 2 && D != nullptr && blockIdx.x < 2) {      // 4 warps: warp w reads TMEM lanes 32 (w % 4) ..
        const int q = warp & 3;
        const int row = 128 * (int)rank + q * 32 + lane;
        for (int c = 0; c < N; c += 16) {
            uint32_t r[16];
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                           "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                         : "r"(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)c));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            for (int i = 0; i < 16; ++i) D[(size_t)row * N + c + i] = __uint_as_float(r[i]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync_all();
    if (warp == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
    }
}

int main() {
    float* D;
    long long* cyc;
    cudaMalloc(&D, 256 * 256 * sizeof(float));
    cudaMalloc(&cyc, 1024 * 8);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    float* h = (float*)malloc(256 * 256 * sizeof(float));
    for (int N : {256, 128, 192}) {
        const size_t smem = 1024 + 128 * 128 + (N / 2) * 128;
        cudaMemset(D, 0xff, 256 * 256 * sizeof(float));
        probe<<<2, 192, smem>>>(N, 1, D, cyc);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("N=%d: %s\n", N, cudaGetErrorString(e)); return 1; }
        cudaMemcpy(h, D, 256 * N * sizeof(float), cudaMemcpyDeviceToHost);
        int bad = 0, first = -1;
        for (int r = 0; r < 256; ++r)
            for (int n = 0; n < N; ++n) {
                float want = 0;
                for (int k = 0; k < 64; ++k) want += a_val(r, k) * b_val(n, k);
                if (h[r * N + n] != want) { if (first < 0) first = r * N + n; ++bad; }
            }
        printf("check M=256 N=%d K=64: %d wrong of %d", N, bad, 256 * N);
        if (bad) printf("  (first at row %d col %d: got %g)", first / N, first % N, h[first]);
        printf("\n");
    }
    for (int N : {64, 128, 192, 256}) {
        const size_t smem = 1024 + 128 * 128 + (N / 2) * 128;
        const int iters = 4000;
        probe<<<148, 192, smem>>>(N, 100, nullptr, cyc);
        probe<<<148, 192, smem>>>(N, iters, nullptr, cyc);
        cudaError_t e = cudaDeviceSynchronize();
        long long hc[74];
        cudaMemcpy(hc, cyc, sizeof(hc), cudaMemcpyDeviceToHost);
        long long mx = 0;
        for (int i = 0; i < 74; ++i) mx = hc[i] > mx ? hc[i] : mx;
        const double per = (double)mx / (iters * 4.0);
        printf("pair MMA 256 x %3d x 16: %6.1f cycles per MMA (ideal %5.1f = N/2, %3.0f%% of peak)  %s\n", N, per, N / 2.0, 100.0 * (N / 2.0) / per,
               e == cudaSuccess ? "" : cudaGetErrorString(e));
    }
    return 0;
}

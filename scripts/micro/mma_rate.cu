// tcgen05.mma issue-rate probe: cycles per 128 x N x 16 fp16 MMA (operands in shared memory, SS mode, cta_group::1) with
// the operands static in smem (no loads), optionally with a second warp hammering shared memory (st.shared) to see how
// much operand bandwidth the tensor pipe needs.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/mma_rate scripts/micro/mma_rate.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

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
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}

__global__ void __launch_bounds__(160, 1) probe(int N, int iters, int stages, int hammer, long long* out) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar, bar2;
    __shared__ uint32_t tmem_base_s;
    __shared__ int stop;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t a_bytes = 128 * 64 * 2, b_bytes = (uint32_t)N * 64 * 2;
    for (uint32_t i = threadIdx.x; i < (uint32_t)stages * (a_bytes + b_bytes) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar2)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        stop = 0;
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base_s;
    if (warp == 0 && lane == 0) {
        const uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            const int s = it % stages;
            const uint64_t ad = make_sw128_desc(smem_u32(smem + (size_t)s * a_bytes));
            const uint64_t bd = make_sw128_desc(smem_u32(smem + (size_t)stages * a_bytes + (size_t)s * b_bytes));
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_f16(tmem + (it & 1) * 256u, ad + k * 2, bd + k * 2, idesc, 1u);
            if (hammer == 3)     // the real kernel frees the operand stage with one commit per k-block
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar2)) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(smem_u32(&bar)) : "memory");
        const long long t1 = clock64();
        if (blockIdx.x == 0) out[0] = t1 - t0;
        *reinterpret_cast<volatile int*>(&stop) = 1;
    } else if (warp >= 1 && hammer == 2) {
        // TMEM readers (the epilogue's tcgen05.ld) on the accumulator buffer the MMAs are not writing... here: both buffers
        // are written alternately, the readers walk all 512 columns of their lane quarter
        const uint32_t trow = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        long long n = 0;
        uint32_t acc = 0;
        while (*reinterpret_cast<volatile int*>(&stop) == 0) {
            for (uint32_t c = 0; c < 512; c += 16) {
                uint32_t r[16];
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                               "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) : "r"(trow + c));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                acc ^= r[0] ^ r[15];
            }
            n += 32;
        }
        if (acc == 0x12345678u) out[1] = 0;
        if (blockIdx.x == 0 && lane == 0 && warp == 1) out[1] = n * 4 * 32 * 16 * 4;     // bytes read by the 4 reader warps
    } else if (warp >= 2 && hammer == 1) {
        // shared-memory traffic from the other warps: 16-byte stores into a scratch area behind the operands
        uint32_t dst = smem_u32(smem + (size_t)stages * (a_bytes + b_bytes)) + threadIdx.x * 16;
        long long n = 0;
        while (*reinterpret_cast<volatile int*>(&stop) == 0) {
#pragma unroll
            for (int u = 0; u < 8; ++u)
                asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(dst + (u & 3) * 2048), "r"((uint32_t)n) : "memory");
            n += 8;
        }
        if (blockIdx.x == 0 && threadIdx.x == 64) out[1] = n * 64 * 16;     // bytes stored by the 2 hammer warps
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
    }
}

int main() {
    long long* d;
    cudaMalloc(&d, 16);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    const int iters = 2000;
    printf("128 x N x 64 k-block (4 MMAs of K=16), %d k-blocks back to back, grid 148; ideal = N/2 cycles per MMA\n", iters);
    for (int hammer = 0; hammer < 4; ++hammer)
        for (int N : {64, 96, 128, 160, 176, 192, 256}) {
            const int stages = 3;
            const size_t smem = 1024 + stages * (128 * 128 + N * 128) + 16 * 1024;
            cudaMemset(d, 0, 16);
            probe<<<148, 160, smem>>>(N, iters, stages, hammer, d);
            cudaError_t e = cudaDeviceSynchronize();
            long long h[2];
            cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
            const double per = (double)h[0] / (iters * 4);
            printf("N=%3d hammer=%d: %7.1f cycles per MMA (ideal %5.1f, %4.0f%% of peak), operand bytes/clk %.0f, %s bytes/clk %.0f  %s\n", N, hammer, per,
                   N / 2.0, 100.0 * (N / 2.0) / per, (128 + N) * 32.0 / per, hammer == 3 ? "commit-per-k-block" : hammer == 2 ? "tmem-read" : "st.shared", hammer ? (double)h[1] / h[0] : 0.0, e == cudaSuccess ? "" : cudaGetErrorString(e));
        }
    return 0;
}

// tcgen05.mma issue-loop probe: why does the MMA warp of gemm_tc_kernel need ~2x the N/2 cycles per 128 x N x 16 MMA that
// scripts/micro/mma_rate.cu measures?  Same loop shape as the kernel (full / empty mbarrier ring between a producer thread and the
// MMA thread, one commit per k-block), no TMA, with each ingredient switchable:
//   mode bit 0: tcgen05.commit per k-block onto empty[s]
//   mode bit 1: producer thread (waits empty[s], arrives full[s]) + MMA thread waits full[s]     (needs bit 0)
//   mode bit 2: tcgen05.fence::after_thread_sync after the wait
//   mode bit 4: ring / tile counters kept incrementally instead of it % stages, it / stages, it % kb_per_tile (runtime divisors)
//   mode bit 3: 8 more warps polling an mbarrier that completes at the very end (the idle epilogue warps)
//   fill 0: operands all zero, 1: pseudo-random fp16 in [-2, 2) (tensor-pipe power)
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/mma_loop scripts/micro/mma_loop.cu
#include <cuda_runtime.h>
#include <cuda_fp16.h>
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
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile("{\n.reg .pred p;\nWL:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra WD;\nbra WL;\nWD:\n}\n"
                 ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n.reg .pred P;\nelect.sync _|P, 0xFFFFFFFF;\nselp.b32 %0, 1, 0, P;\n}\n" : "=r"(pred));
    return pred != 0;
}

constexpr int MAX_STAGES = 8;

__global__ void __launch_bounds__(320, 1) probe(int N, int iters, int stages, int mode, int fill, int kb_per_tile, long long* out) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t full_bar[MAX_STAGES], empty_bar[MAX_STAGES], done_bar;
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t a_bytes = 128 * 64 * 2, b_bytes = (uint32_t)N * 64 * 2;
    const uint32_t words = (uint32_t)stages * (a_bytes + b_bytes) / 4;
    for (uint32_t i = threadIdx.x; i < words; i += blockDim.x) {
        uint32_t v = 0;
        if (fill) {
            uint32_t h = (i + 1u) * 2654435761u + blockIdx.x * 40503u;
            h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
            const float f0 = ((h & 0xFFFF) / 65536.0f - 0.5f) * 4.0f, f1 = ((h >> 16) / 65536.0f - 0.5f) * 4.0f;
            const __half2 hh = __floats2half2_rn(f0, f1);
            v = *reinterpret_cast<const uint32_t*>(&hh);
        }
        reinterpret_cast<uint32_t*>(smem)[i] = v;
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(&done_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base_s;
    if (warp == 0) {
        if (lane == 0 && (mode & 2)) {
            for (int it = 0; it < iters; ++it) {
                const int s = it % stages;
                const uint32_t ph = (it / stages) & 1u;
                mbar_wait(&empty_bar[s], ph ^ 1u);
                mbar_arrive(&full_bar[s]);
            }
        }
    } else if (warp == 1 && (mode & 32)) {
        // the kernels' style: the whole warp runs the loop (addresses and counters stay warp-uniform), only the issue is elected;
        // mode bit 6: the K=16 steps of TWO k-blocks issued back to back before their two commits
        const uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        const long long t0 = clock64();
        int s = 0, kb = 0;
        uint32_t ph = 0, buf = 0;
        const uint64_t ad0 = make_sw128_desc(smem_u32(smem)), bd0 = make_sw128_desc(smem_u32(smem + (size_t)stages * a_bytes));
        const uint64_t a_step = a_bytes >> 4, b_step = b_bytes >> 4;
        uint64_t ad = ad0, bd = bd0;
        for (int it = 0; it < iters; ++it) {
            if (mode & 2) mbar_wait(&full_bar[s], ph);
            if (mode & 4) asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_f16(tmem + buf * 256u, ad + k * 2, bd + k * 2, idesc, (kb | k) ? 1u : 0u);
                if (mode & 1) umma_commit(&empty_bar[s]);
            }
            __syncwarp();
            ad += a_step; bd += b_step;
            if (++s == stages) { s = 0; ph ^= 1u; ad = ad0; bd = bd0; }
            if (++kb == kb_per_tile) { kb = 0; buf ^= 1u; }
        }
        if (elect_one()) umma_commit(&done_bar);
        __syncwarp();
        mbar_wait(&done_bar, 0);
        const long long t1 = clock64();
        if (blockIdx.x == 0 && lane == 0) out[0] = t1 - t0;
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
            const long long t0 = clock64();
            int s_i = 0, kb_i = 0;
            uint32_t ph_i = 0, buf_i = 0;
            for (int it = 0; it < iters; ++it) {
                int s, kb;
                uint32_t ph, buf;
                if (mode & 16) {      // strength-reduced ring / tile counters: no integer division on the issue path
                    s = s_i; ph = ph_i; kb = kb_i; buf = buf_i;
                    if (++s_i == stages) { s_i = 0; ph_i ^= 1u; }
                    if (++kb_i == kb_per_tile) { kb_i = 0; buf_i ^= 1u; }
                } else {
                    s = it % stages; ph = (it / stages) & 1u; kb = it % kb_per_tile; buf = (it / kb_per_tile) & 1u;
                }
                if (mode & 2) mbar_wait(&full_bar[s], ph);
                if (mode & 4) asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint64_t ad = make_sw128_desc(smem_u32(smem + (size_t)s * a_bytes));
                const uint64_t bd = make_sw128_desc(smem_u32(smem + (size_t)stages * a_bytes + (size_t)s * b_bytes));
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_f16(tmem + buf * 256u, ad + k * 2, bd + k * 2, idesc, (kb | k) ? 1u : 0u);
                if (mode & 1) umma_commit(&empty_bar[s]);
            }
            umma_commit(&done_bar);
            mbar_wait(&done_bar, 0);
            const long long t1 = clock64();
            if (blockIdx.x == 0) out[0] = t1 - t0;
        }
    } else if (mode & 8) {
        mbar_wait(&done_bar, 0);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
    }
}

int main() {
    long long* d;
    cudaMalloc(&d, 16);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    const int iters = 4000;
    printf("128 x N x 64 k-blocks (4 MMAs of K=16 each), %d k-blocks, 6 per tile, grid 148; ideal = N/2 cycles per MMA\n", iters);
    for (int fill = 1; fill < 2; ++fill)
        for (int mode : {16, 23, 32, 33, 39})
            for (int N : {64, 96, 128, 160, 176, 256}) {
                const int stages = 4;
                const size_t smem = 1024 + stages * (128 * 128 + N * 128);
                float ms = 0;
                cudaEvent_t e0, e1;
                cudaEventCreate(&e0); cudaEventCreate(&e1);
                probe<<<148, 320, smem>>>(N, 200, stages, mode, fill, 6, d);     // warm-up
                cudaEventRecord(e0);
                probe<<<148, 320, smem>>>(N, iters, stages, mode, fill, 6, d);
                cudaEventRecord(e1);
                cudaError_t e = cudaDeviceSynchronize();
                cudaEventElapsedTime(&ms, e0, e1);
                long long h[2];
                cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
                const double per = (double)h[0] / (iters * 4.0);
                printf("fill=%d mode=%2d N=%3d: %7.1f cycles per MMA (ideal %5.1f, %4.0f%% of peak)  %.0f us -> %.2f GHz, %.0f TFLOP/s  %s\n", fill, mode, N, per,
                       N / 2.0, 100.0 * (N / 2.0) / per, ms * 1e3, h[0] / (ms * 1e6), 148.0 * iters * 4 * 2.0 * 128 * N * 16 / (ms * 1e-3) * 1e-12,
                       e == cudaSuccess ? "" : cudaGetErrorString(e));
            }
    return 0;
}

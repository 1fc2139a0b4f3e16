// fp16 tensor-core tier of the pointwise-convolution / matmul op: TMA -> shared (128B swizzle) ->
// tcgen05.mma (fp32 accumulators in TMEM) -> tcgen05.ld -> fused epilogue -> global.
//
// D[m, n] = epi(alpha * sum_k A[m, k] * B[n, k]); A = activations in rows layout (K contiguous),
// B = conv weight [Cout, Cin].  Same contract as the FFMA tier (gemm_ffma.cu); the reference runs
// these layers under fp16 autocast on CUDA (MOFNetStack/network.py:149-192).
//
// One CTA = one 128 x BN output tile (BN runtime, multiple of 16, <= 256).  Warp 0 lane 0 issues
// TMA loads of 128x64 (A) and BNx64 (B) fp16 boxes into a ring of stages; warp 1 lane 0 issues
// tcgen05.mma (M=128, N=BN, K=16, 4 per stage) and commits stage release / accumulator-ready
// mbarriers; warps 2-5 read the accumulator (each warp its own 32-lane TMEM quarter), apply
// bias / residual / GELU / ReLU / the k=1 depthwise stage and store.  Out-of-range rows, columns
// and the K tail are zero-filled by TMA, so no shape needs padding.  Shared memory is sized so
// that two CTAs are resident per SM: one CTA's epilogue overlaps the other's MMA main loop.
#include <cuda.h>

#include "common.cuh"

namespace vfml {

constexpr int TC_BM = 128, TC_BK = 64;
constexpr int TC_THREADS = 192;
constexpr int TC_MAX_STAGES = 6;

struct TcParams {
    void* D;
    const float* bias;
    const __half* res;
    const float* dw_w;
    const float* dw_b;
    int64_t ldd, ldr, strideD, strideR;
    int M, N, K;
    int BN, stages;
    float alpha;
    uint32_t epi;
    int out_f32;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_3d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// shared-memory matrix descriptor, K-major, 128-byte swizzle: 8-row groups of 128 B rows, 1024 B apart
__device__ __forceinline__ uint64_t make_sw128_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);          // start address
    d |= (uint64_t)1 << 16;                          // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;                // stride byte offset between 8-row groups
    d |= (uint64_t)1 << 46;                          // descriptor version (Blackwell)
    d |= (uint64_t)2 << 61;                          // SWIZZLE_128B
    return d;
}

__global__ void __launch_bounds__(TC_THREADS)
gemm_tc_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, const TcParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int BN = p.BN, stages = p.stages;
    const uint32_t a_bytes = TC_BM * TC_BK * 2, b_bytes = (uint32_t)BN * TC_BK * 2;
    const uint32_t b_stride = (b_bytes + 1023) & ~1023u;
    uint8_t* sA = smem;
    uint8_t* sB = smem + (size_t)stages * a_bytes;
    __shared__ uint64_t full_bar[TC_MAX_STAGES], empty_bar[TC_MAX_STAGES], acc_bar;
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.y * TC_BM, n0 = blockIdx.x * BN, bz = blockIdx.z;
    const int nkb = (p.K + TC_BK - 1) / TC_BK;

    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&acc_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&mapA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&mapB) : "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(256u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;

    if (warp == 0) {
        if (lane == 0) {
            for (int kb = 0; kb < nkb; ++kb) {
                const int s = kb % stages;
                const uint32_t ph = (uint32_t)(kb / stages) & 1u;
                mbar_wait(&empty_bar[s], ph ^ 1u);
                mbar_expect_tx(&full_bar[s], a_bytes + b_bytes);
                tma_load_3d(&mapA, &full_bar[s], sA + (size_t)s * a_bytes, kb * TC_BK, m0, bz);
                tma_load_3d(&mapB, &full_bar[s], sB + (size_t)s * b_stride, kb * TC_BK, n0, bz);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // instruction descriptor: D fp32, A/B fp16, both K-major, N = BN, M = 128
            const uint32_t idesc = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
            for (int kb = 0; kb < nkb; ++kb) {
                const int s = kb % stages;
                const uint32_t ph = (uint32_t)(kb / stages) & 1u;
                mbar_wait(&full_bar[s], ph);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint64_t adesc = make_sw128_desc(smem_u32(sA + (size_t)s * a_bytes));
                const uint64_t bdesc = make_sw128_desc(smem_u32(sB + (size_t)s * b_stride));
#pragma unroll
                for (int k = 0; k < TC_BK / 16; ++k) {
                    // +32 bytes per K=16 step inside the 128-byte swizzle atom
                    umma_f16(tmem_base, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) ? 1u : 0u);
                }
                umma_commit(&empty_bar[s]);
            }
            umma_commit(&acc_bar);
        }
    } else {
        const int q = warp & 3;                   // TMEM lane quarter this warp may access
        const int m = m0 + q * 32 + lane;
        mbar_wait(&acc_bar, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const bool row_ok = m < p.M;
        const __half* rrow = p.res ? p.res + bz * p.strideR + (int64_t)m * p.ldr : nullptr;
        __half* d16 = reinterpret_cast<__half*>(p.D) + bz * p.strideD + (int64_t)m * p.ldd;
        float* d32 = reinterpret_cast<float*>(p.D) + bz * p.strideD + (int64_t)m * p.ldd;
        for (int c = 0; c < BN; c += 16) {
            const int n = n0 + c;
            if (n >= p.N) break;                  // warp-uniform
            float v[16];
            tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c, v);
            if (!row_ok) continue;
            const bool full = n + 16 <= p.N;
            float r[16];
            if (p.epi & VFML_EPI_RES) {
                if (full && ((reinterpret_cast<uintptr_t>(rrow + n) & 15) == 0)) {
                    const uint4 t0 = *reinterpret_cast<const uint4*>(rrow + n);
                    const uint4 t1 = *reinterpret_cast<const uint4*>(rrow + n + 8);
                    const __half2* h0 = reinterpret_cast<const __half2*>(&t0);
                    const __half2* h1 = reinterpret_cast<const __half2*>(&t1);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float2 f0 = __half22float2(h0[i]), f1 = __half22float2(h1[i]);
                        r[2 * i] = f0.x; r[2 * i + 1] = f0.y; r[8 + 2 * i] = f1.x; r[8 + 2 * i + 1] = f1.y;
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 16; ++i) r[i] = (n + i < p.N) ? __half2float(rrow[n + i]) : 0.0f;
                }
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const int nn = min(n + i, p.N - 1);
                float x = v[i] * p.alpha;
                if (p.epi & VFML_EPI_BIAS) x += __ldg(p.bias + nn);
                if (p.epi & VFML_EPI_RES) x += r[i];
                if (p.epi & VFML_EPI_GELU) x = gelu_erf(x);
                if (p.epi & VFML_EPI_RELU) x = fmaxf(x, 0.0f);
                if (p.epi & VFML_EPI_DW1) x = gelu_erf(x + (__ldg(p.dw_w + nn) * x + __ldg(p.dw_b + nn)));
                v[i] = x;
            }
            if (p.out_f32) {
                if (full && ((reinterpret_cast<uintptr_t>(d32 + n) & 15) == 0)) {
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        *reinterpret_cast<float4*>(d32 + n + 4 * i) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
                } else {
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        if (n + i < p.N) d32[n + i] = v[i];
                }
            } else {
                if (full && ((reinterpret_cast<uintptr_t>(d16 + n) & 15) == 0)) {
                    uint4 o0, o1;
                    __half2* h0 = reinterpret_cast<__half2*>(&o0);
                    __half2* h1 = reinterpret_cast<__half2*>(&o1);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        h0[i] = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
                        h1[i] = __floats2half2_rn(v[8 + 2 * i], v[8 + 2 * i + 1]);
                    }
                    *reinterpret_cast<uint4*>(d16 + n) = o0;
                    *reinterpret_cast<uint4*>(d16 + n + 8) = o1;
                } else {
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        if (n + i < p.N) d16[n + i] = __float2half_rn(v[i]);
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(256u) : "memory");
    }
}

// ------------------------------------------------------------------------------------------ host
static int make_map(CUtensorMap* map, const void* base, int64_t inner, int64_t rows, int64_t batch, int64_t ld,
                    int64_t stride, int box_rows) {
    cuuint64_t dims[3] = {(cuuint64_t)inner, (cuuint64_t)rows, (cuuint64_t)batch};
    cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)(batch > 1 ? stride : ld * rows) * 2};
    cuuint32_t box[3] = {(cuuint32_t)TC_BK, (cuuint32_t)box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    // libcuda is resolved at run time through the runtime (no link-time dependency: the library must
    // load on a CPU-only build box for the symbol check)
    typedef CUresult (*encode_fn_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static encode_fn_t encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
            set_error("gemm_tc: cuTensorMapEncodeTiled is not available from the driver (%s)", cudaGetErrorString(e));
            return VFML_ERR_CUDA;
        }
        encode = reinterpret_cast<encode_fn_t>(fn);
    }
    CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(base), dims, strides, box,
                                        estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("gemm_tc: cuTensorMapEncodeTiled failed: CUresult %d (inner=%lld rows=%lld batch=%lld ld=%lld)", (int)r,
                  (long long)inner, (long long)rows, (long long)batch, (long long)ld);
        return VFML_ERR_CUDA;
    }
    return VFML_OK;
}

// can this descriptor go through the TMA path? (16-byte aligned bases and pitches, fp16 operands)
bool gemm_tc_supported(const vfml_gemm_desc& d) {
    if (d.in_dtype != VFML_F16) return false;
    if (d.epi & VFML_EPI_ADDTO) return false;
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    if (!al16(d.A) || !al16(d.B)) return false;
    if ((d.lda % 8) || (d.ldb % 8)) return false;
    if (d.batch > 1 && ((d.strideA % 8) || (d.strideB % 8))) return false;
    return d.K >= 1;
}

static int pick_bn(int N) {
    if (N <= 256) return (N + 15) / 16 * 16;
    const int tiles = (N + 255) / 256;
    return ((N + tiles - 1) / tiles + 15) / 16 * 16;
}

int gemm_tc(const vfml_gemm_desc& d, cudaStream_t st) {
    const int BN = pick_bn(d.N);
    CUtensorMap mapA, mapB;
    int rc = make_map(&mapA, d.A, d.K, d.M, d.batch, d.lda, d.strideA, TC_BM);
    if (rc) return rc;
    rc = make_map(&mapB, d.B, d.K, d.N, d.batch, d.ldb, d.strideB, BN);
    if (rc) return rc;
    const int nkb = (d.K + TC_BK - 1) / TC_BK;
    const size_t a_bytes = TC_BM * TC_BK * 2, b_stride = ((size_t)BN * TC_BK * 2 + 1023) & ~(size_t)1023;
    int stages = (int)((110 * 1024 - 1024) / (a_bytes + b_stride));     // two CTAs per SM
    stages = std::max(2, std::min(std::min(stages, TC_MAX_STAGES), std::max(nkb, 2)));
    const size_t smem = 1024 + stages * (a_bytes + b_stride);
    static bool attr_done = false;
    if (!attr_done) {
        VFML_CUDA(cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_done = true;
    }
    TcParams p{};
    p.D = d.D; p.bias = d.bias; p.res = reinterpret_cast<const __half*>(d.res); p.dw_w = d.dw_w; p.dw_b = d.dw_b;
    p.ldd = d.ldd; p.ldr = d.ldr; p.strideD = d.strideD; p.strideR = d.strideR;
    p.M = d.M; p.N = d.N; p.K = d.K; p.BN = BN; p.stages = stages;
    p.alpha = d.alpha; p.epi = d.epi; p.out_f32 = d.out_dtype == VFML_F32;
    dim3 grid(cdiv(d.N, BN), cdiv(d.M, TC_BM), d.batch);
    gemm_tc_kernel<<<grid, TC_THREADS, smem, st>>>(mapA, mapB, p);
    VFML_LAUNCH_CHECK("gemm_tc");
    return VFML_OK;
}

}  // namespace vfml

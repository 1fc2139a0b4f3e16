// fp32-exact tier of the pointwise-convolution / matmul op: register-tiled FFMA GEMM with the
// PCBlock epilogues fused (bias, residual, erf-GELU, ReLU, the k=1 depthwise stage).
//
// D[m, n] = epi(alpha * sum_k A[m, k] * B[n, k]);  A = activations in rows layout, B = conv weight
// [Cout, Cin] (MOFNetStack/stack.py:8-32 ffn1/pw/ffn2, gma.py to_qk / to_v, q.k^T, attn.v).
// This tier reproduces the reference's true-fp32 arithmetic (needed for the 1e-2 px EPE
// bound after 32 recurrent iterations); the fp16 tensor-core tier lives in gemm_tc.cu.
#include "common.cuh"

namespace vfml {

constexpr int G_BM = 128, G_BK = 16;

template <typename TI, typename TO, int BN>
__global__ void __launch_bounds__(256)
gemm_ffma_kernel(const vfml_gemm_desc d) {
    constexpr int TN = BN / 16;
    constexpr int APITCH = G_BM + 4, BPITCH = BN + 4;
    __shared__ __align__(16) float As[G_BK][APITCH];
    __shared__ __align__(16) float Bs[G_BK][BPITCH];
    const int bz = blockIdx.z;
    const TI* A = reinterpret_cast<const TI*>(d.A) + bz * d.strideA;
    const TI* B = reinterpret_cast<const TI*>(d.B) + bz * d.strideB;
    TO* D = reinterpret_cast<TO*>(d.D) + bz * d.strideD;
    const TI* R = d.res ? reinterpret_cast<const TI*>(d.res) + bz * d.strideR : nullptr;
    const int m0 = blockIdx.y * G_BM, n0 = blockIdx.x * BN;
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;

    float acc[8][TN];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.0f;

    // loader mapping: 4 consecutive k per thread, rows tid/4 (+64)
    const int lrow = tid >> 2, lk = (tid & 3) * 4;
    for (int k0 = 0; k0 < d.K; k0 += G_BK) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const int r = lrow + half * 64;
            const int m = m0 + r;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int k = k0 + lk + i;
                As[lk + i][r] = (m < d.M && k < d.K) ? to_f32<TI>(A[(int64_t)m * d.lda + k]) : 0.0f;
            }
        }
#pragma unroll
        for (int half = 0; half < (BN + 63) / 64; ++half) {
            const int r = lrow + half * 64;
            if (r < BN) {
                const int n = n0 + r;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int k = k0 + lk + i;
                    Bs[lk + i][r] = (n < d.N && k < d.K) ? to_f32<TI>(B[(int64_t)n * d.ldb + k]) : 0.0f;
                }
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < G_BK; ++k) {
            float a[8], b[TN];
            *reinterpret_cast<float4*>(&a[0]) = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            *reinterpret_cast<float4*>(&a[4]) = *reinterpret_cast<const float4*>(&As[k][64 + ty * 4]);
            if constexpr (TN % 4 == 0) {
#pragma unroll
                for (int q = 0; q < TN / 4; ++q)
                    *reinterpret_cast<float4*>(&b[4 * q]) = *reinterpret_cast<const float4*>(&Bs[k][tx * TN + 4 * q]);
            } else {
#pragma unroll
                for (int j = 0; j < TN; ++j) b[j] = Bs[k][tx * TN + j];
            }
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }

#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (m >= d.M) continue;
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            const int n = n0 + tx * TN + j;
            if (n >= d.N) continue;
            float v = acc[i][j] * d.alpha;
            if (d.epi & VFML_EPI_BIAS) v += d.bias[n];
            if (d.epi & VFML_EPI_RES) v += to_f32<TI>(R[(int64_t)m * d.ldr + n]);
            if (d.epi & VFML_EPI_GELU) v = gelu_erf(v);
            if (d.epi & VFML_EPI_RELU) v = fmaxf(v, 0.0f);
            if (d.epi & VFML_EPI_DW1) v = gelu_erf(v + (d.dw_w[n] * v + d.dw_b[n]));
            TO* dst = D + (int64_t)m * d.ldd + n;
            if (d.epi & VFML_EPI_ADDTO) v += to_f32<TO>(*dst);
            *dst = from_f32<TO>(v);
        }
    }
}

template <typename TI, typename TO>
static int launch_gemm_ffma(const vfml_gemm_desc& d, cudaStream_t st) {
    dim3 grid(1, cdiv(d.M, G_BM), d.batch);
    if (d.N > 64) {
        grid.x = cdiv(d.N, 128);
        gemm_ffma_kernel<TI, TO, 128><<<grid, 256, 0, st>>>(d);
    } else if (d.N > 16) {
        grid.x = cdiv(d.N, 64);
        gemm_ffma_kernel<TI, TO, 64><<<grid, 256, 0, st>>>(d);
    } else {
        grid.x = 1;
        gemm_ffma_kernel<TI, TO, 16><<<grid, 256, 0, st>>>(d);
    }
    VFML_LAUNCH_CHECK("gemm_ffma");
    return VFML_OK;
}

int gemm_ffma(const vfml_gemm_desc& d, cudaStream_t st) {
    if (d.in_dtype == VFML_F32 && d.out_dtype == VFML_F32) return launch_gemm_ffma<float, float>(d, st);
    if (d.in_dtype == VFML_F32 && d.out_dtype == VFML_F16) return launch_gemm_ffma<float, __half>(d, st);
    if (d.in_dtype == VFML_F16 && d.out_dtype == VFML_F16) return launch_gemm_ffma<__half, __half>(d, st);
    return launch_gemm_ffma<__half, float>(d, st);
}

int validate_gemm(const vfml_gemm_desc* d) {
    VFML_REQUIRE(d, "gemm: null descriptor");
    VFML_REQUIRE(d->A && d->B && d->D, "gemm: null operand");
    VFML_REQUIRE(d->M >= 0 && d->N >= 0 && d->K >= 0 && d->batch >= 1, "gemm: bad shape M=%d N=%d K=%d batch=%d",
                 d->M, d->N, d->K, d->batch);
    VFML_REQUIRE(d->lda >= d->K && d->ldb >= d->K && d->ldd >= d->N, "gemm: leading dimension too small");
    VFML_REQUIRE(valid_dtype(d->in_dtype) && valid_dtype(d->out_dtype), "gemm: bad dtype");
    if (d->epi & VFML_EPI_BIAS) VFML_REQUIRE(d->bias, "gemm: EPI_BIAS without bias");
    if (d->epi & VFML_EPI_RES) VFML_REQUIRE(d->res && d->ldr >= d->N, "gemm: EPI_RES without residual");
    if (d->epi & VFML_EPI_DW1) VFML_REQUIRE(d->dw_w && d->dw_b, "gemm: EPI_DW1 without dw weights");
    if (d->epi & VFML_EPI_ADDTO) VFML_REQUIRE(d->out_dtype == VFML_F32, "gemm: EPI_ADDTO needs fp32 output");
    return VFML_OK;
}

}  // namespace vfml

// All-pairs correlation volume and its average-pool pyramid (fp32 tier: exact fp32 FFMA).
//
// Replaces CorrBlock.corr / CorrBlock.__init__ (MOFNetStack/corr.py:60-74, 99-107):
//   vol[b, q, t] = sum_c fmap1[b, c, q] * fmap2[b, c, t] / sqrt(C),  then 3x avg_pool2d(2, 2) over t.
// Both operands are stored [C][P] (pixel contiguous), which is exactly the k-outer layout a
// register-tiled SGEMM wants: 128x128 CTA tile, 8x8 per thread, BK = 8, double-buffered smem.
#include "common.cuh"

namespace vfml {

constexpr int CV_BM = 128, CV_BN = 128, CV_BK = 8;

template <typename TV>
__global__ void __launch_bounds__(256)
corr_volume_kernel(const float* __restrict__ f1, const float* __restrict__ f2, TV* __restrict__ vol,
                   int C, int P, float scale) {
    __shared__ __align__(16) float As[2][CV_BK][CV_BM];
    __shared__ __align__(16) float Bs[2][CV_BK][CV_BN];
    const int b = blockIdx.z;
    const float* A = f1 + (int64_t)b * C * P;
    const float* B = f2 + (int64_t)b * C * P;
    TV* D = vol + (int64_t)b * P * P;
    const int m0 = blockIdx.y * CV_BM, n0 = blockIdx.x * CV_BN;
    const int tid = threadIdx.x;
    const int lk = tid >> 5;          // 0..7: k row loaded by this thread
    const int lc = (tid & 31) * 4;    // 0..124: column group
    const bool vec = (P % 4) == 0;

    auto load_tile = [&](const float* src, int c0, int k0, float (&dst)[4]) {
        const int k = k0 + lk;
        const int col = c0 + lc;
        if (k < C && vec && col + 3 < P) {
            const float4 t = *reinterpret_cast<const float4*>(src + (int64_t)k * P + col);
            dst[0] = t.x; dst[1] = t.y; dst[2] = t.z; dst[3] = t.w;
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) dst[i] = (k < C && col + i < P) ? src[(int64_t)k * P + col + i] : 0.0f;
        }
    };

    const int tx = tid & 15, ty = tid >> 4;   // 16 x 16 threads, each 8 x 8 outputs
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.0f;

    float ra[4], rb[4];
    load_tile(A, m0, 0, ra);
    load_tile(B, n0, 0, rb);
    *reinterpret_cast<float4*>(&As[0][lk][lc]) = make_float4(ra[0], ra[1], ra[2], ra[3]);
    *reinterpret_cast<float4*>(&Bs[0][lk][lc]) = make_float4(rb[0], rb[1], rb[2], rb[3]);
    __syncthreads();
    const int nk = (C + CV_BK - 1) / CV_BK;
    for (int kt = 0; kt < nk; ++kt) {
        const int cur = kt & 1;
        if (kt + 1 < nk) {
            load_tile(A, m0, (kt + 1) * CV_BK, ra);
            load_tile(B, n0, (kt + 1) * CV_BK, rb);
        }
#pragma unroll
        for (int k = 0; k < CV_BK; ++k) {
            float a[8], bb[8];
            *reinterpret_cast<float4*>(&a[0]) = *reinterpret_cast<const float4*>(&As[cur][k][ty * 4]);
            *reinterpret_cast<float4*>(&a[4]) = *reinterpret_cast<const float4*>(&As[cur][k][64 + ty * 4]);
            *reinterpret_cast<float4*>(&bb[0]) = *reinterpret_cast<const float4*>(&Bs[cur][k][tx * 4]);
            *reinterpret_cast<float4*>(&bb[4]) = *reinterpret_cast<const float4*>(&Bs[cur][k][64 + tx * 4]);
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
        }
        if (kt + 1 < nk) {
            *reinterpret_cast<float4*>(&As[cur ^ 1][lk][lc]) = make_float4(ra[0], ra[1], ra[2], ra[3]);
            *reinterpret_cast<float4*>(&Bs[cur ^ 1][lk][lc]) = make_float4(rb[0], rb[1], rb[2], rb[3]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (m >= P) continue;
#pragma unroll
        for (int jh = 0; jh < 2; ++jh) {
            const int n = n0 + jh * 64 + tx * 4;
            TV* dst = D + (int64_t)m * P + n;
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (n + j < P) dst[j] = from_f32<TV>(acc[i][jh * 4 + j] * scale);
        }
    }
}

// one avg_pool2d(2, stride 2) level: in [rows, hi, wi] -> out [rows, hi/2, wi/2]
template <typename TV>
__global__ void corr_pool_kernel(const TV* __restrict__ in, TV* __restrict__ out, int64_t rows, int hi, int wi) {
    const int ho = hi / 2, wo = wi / 2;
    const int64_t total = rows * ho * wo;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int x = (int)(t % wo);
        const int y = (int)((t / wo) % ho);
        const int64_t r = t / ((int64_t)wo * ho);
        const TV* src = in + (r * hi + 2 * y) * (int64_t)wi + 2 * x;
        const float s = ((to_f32<TV>(src[0]) + to_f32<TV>(src[1])) + to_f32<TV>(src[wi])) + to_f32<TV>(src[wi + 1]);
        out[t] = from_f32<TV>(s * 0.25f);
    }
}

template <typename TV>
static int run_volume(const float* f1, const float* f2, int n_pairs, int C, int h, int w, void* const* levels,
                      int num_levels, cudaStream_t st) {
    const int P = h * w;
    dim3 grid(cdiv(P, CV_BN), cdiv(P, CV_BM), n_pairs);
    corr_volume_kernel<TV><<<grid, 256, 0, st>>>(f1, f2, reinterpret_cast<TV*>(levels[0]), C, P, 1.0f / sqrtf((float)C));
    VFML_LAUNCH_CHECK("corr_volume");
    int hi = h, wi = w;
    for (int l = 1; l < num_levels; ++l) {
        const int64_t rows = (int64_t)n_pairs * P;
        const int64_t total = rows * (hi / 2) * (wi / 2);
        if (total > 0) {
            const int blocks = (int)std::min<int64_t>((total + 255) / 256, device_sms() * 32);
            corr_pool_kernel<TV><<<blocks, 256, 0, st>>>(reinterpret_cast<const TV*>(levels[l - 1]),
                                                        reinterpret_cast<TV*>(levels[l]), rows, hi, wi);
            VFML_LAUNCH_CHECK("corr_pool");
        }
        hi /= 2;
        wi /= 2;
    }
    return VFML_OK;
}

}  // namespace vfml

using namespace vfml;

// fp16 tier: the volume is a dense contraction -> tensor cores.  Both feature maps are transposed to
// pixel-major fp16 rows ([P, C], K contiguous) in the caller's workspace and multiplied by the tcgen05
// GEMM (batch = pairs, alpha = 1/sqrt(C), fp16 output straight into level 0); the pyramid follows.
static bool volume_uses_tc(int C, int vol_dtype) { return vol_dtype == VFML_F16 && (C % 8) == 0; }

extern "C" size_t vfml_corr_volume_workspace(int n_pairs, int C, int h, int w, int vol_dtype) {
    if (!volume_uses_tc(C, vol_dtype)) return 0;
    return (size_t)2 * n_pairs * h * w * C * sizeof(__half) + 256;
}

static int run_volume_tc(const float* f1, const float* f2, int n_pairs, int C, int h, int w, void* const* levels,
                         int num_levels, void* workspace, size_t workspace_bytes, vfml_stream_t s) {
    const int P = h * w;
    const size_t need = vfml_corr_volume_workspace(n_pairs, C, h, w, VFML_F16);
    if (!workspace || workspace_bytes < need) {
        set_error("corr_volume_pyramid: workspace of %zu bytes needed, %zu given", need, workspace_bytes);
        return VFML_ERR_WORKSPACE;
    }
    __half* r1 = reinterpret_cast<__half*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
    __half* r2 = r1 + (size_t)n_pairs * P * C;
    int rc = vfml_nchw_to_rows(f1, r1, n_pairs, C, P, C, 0, VFML_F16, s);
    if (rc) return rc;
    rc = vfml_nchw_to_rows(f2, r2, n_pairs, C, P, C, 0, VFML_F16, s);
    if (rc) return rc;
    vfml_gemm_desc d{};
    d.A = r1; d.B = r2; d.D = levels[0];
    d.lda = C; d.ldb = C; d.ldd = P;
    d.M = P; d.N = P; d.K = C; d.batch = n_pairs;
    d.strideA = (int64_t)P * C; d.strideB = (int64_t)P * C; d.strideD = (int64_t)P * P;
    d.alpha = 1.0f / sqrtf((float)C);
    d.in_dtype = VFML_F16; d.out_dtype = VFML_F16;
    rc = vfml_gemm(&d, s);
    if (rc) return rc;
    cudaStream_t st = to_stream(s);
    int hi = h, wi = w;
    for (int l = 1; l < num_levels; ++l) {
        const int64_t rows = (int64_t)n_pairs * P;
        const int64_t total = rows * (hi / 2) * (wi / 2);
        if (total > 0) {
            const int blocks = (int)std::min<int64_t>((total + 255) / 256, device_sms() * 32);
            corr_pool_kernel<__half><<<blocks, 256, 0, st>>>(reinterpret_cast<const __half*>(levels[l - 1]),
                                                            reinterpret_cast<__half*>(levels[l]), rows, hi, wi);
            VFML_LAUNCH_CHECK("corr_pool");
        }
        hi /= 2;
        wi /= 2;
    }
    return VFML_OK;
}

extern "C" int vfml_corr_volume_pyramid(const float* fmap1, const float* fmap2, int n_pairs, int C, int h, int w,
                                        void* const* levels, int num_levels, int vol_dtype, void* workspace,
                                        size_t workspace_bytes, vfml_stream_t s) {
    (void)workspace; (void)workspace_bytes;
    VFML_REQUIRE(n_pairs >= 0 && C > 0 && h > 0 && w > 0, "corr_volume_pyramid: bad shape");
    if (n_pairs == 0) return VFML_OK;
    VFML_REQUIRE(fmap1 && fmap2 && levels, "corr_volume_pyramid: null pointer");
    VFML_REQUIRE(num_levels >= 1 && num_levels <= 4, "corr_volume_pyramid: num_levels %d not in 1..4", num_levels);
    VFML_REQUIRE(valid_dtype(vol_dtype), "corr_volume_pyramid: bad dtype");
    for (int l = 0; l < num_levels; ++l) VFML_REQUIRE(levels[l], "corr_volume_pyramid: level %d is null", l);
    if (volume_uses_tc(C, vol_dtype))
        return run_volume_tc(fmap1, fmap2, n_pairs, C, h, w, levels, num_levels, workspace, workspace_bytes, s);
    if (vol_dtype == VFML_F32) return run_volume<float>(fmap1, fmap2, n_pairs, C, h, w, levels, num_levels, to_stream(s));
    return run_volume<__half>(fmap1, fmap2, n_pairs, C, h, w, levels, num_levels, to_stream(s));
}

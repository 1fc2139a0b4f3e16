// Volume-free correlation lookup: the arithmetic of alt_cuda_corr.forward
// (VideoFlow/alt_cuda_corr/correlation_kernel.cu:18-119, bound at correlation.cpp:51-54) as called by
// AlternateCorrBlock.__call__ (MOFNetStack/corr.py:160-177).
//
//   s[iy][ix]  = <fmap1[b, h1, w1, :], fmap2[b, floor(y)-r+iy, floor(x)-r+ix, :]>   (0 outside fmap2), iy, ix in 0..2r+1
//   corr[b, n, ix*(2r+1) + iy, h1, w1] = (1-dy)(1-dx) s[iy][ix] + (1-dy) dx s[iy][ix+1] + dy (1-dx) s[iy+1][ix] + dy dx s[iy+1][ix+1]
//
// with (x, y) = coords[b, n, h1, w1, :], dx = x - floor(x), dy = y - floor(y): the bilinear sample of the
// correlation surface at (x - r + ix, y - r + iy) with zeros padding, no 1/sqrt(C) factor (the caller applies it).
//
// Shape of the work: (2r+2)^2 = 100 dot products of C = 256 per query pixel against texels of an fmap2 level
// that fits in L2 (3.6 MB fp16 at 55 x 128), i.e. L2->SM gather traffic, not HBM.  A CTA owns 32 consecutive query
// pixels of one (b, n) so that the [81][32] result block is written as 128-byte rows; a warp owns a pixel at a
// time: its 32 lanes split the channels (one 16-byte load per lane and texel = one coalesced 512-byte row),
// 32 texel partials are kept in registers and reduced across lanes by a 31-shuffle transposing butterfly
// (instead of 32 x 5), after which lane k holds s of texel k.
#include "common.cuh"

namespace vfml {

constexpr int AC_PIX = 32;       // query pixels per CTA
constexpr int AC_WARPS = 8;
constexpr int AC_MAXR = 4;
constexpr int AC_MAXT = (2 * AC_MAXR + 2) * (2 * AC_MAXR + 2);   // 100 texels
constexpr int AC_MAXO = (2 * AC_MAXR + 1) * (2 * AC_MAXR + 1);   // 81 outputs

struct AltParams {
    const void* fmap1;
    const void* fmap2;
    const float* coords;
    float* corr;
    int B, N, H1, W1, H2, W2, C, r;
};

// v[k] of every lane -> lane k's return value = sum over lanes of v[k]
__device__ __forceinline__ float transpose_reduce32(float (&v)[32], int lane) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        const bool upper = (lane & off) != 0;
#pragma unroll
        for (int k = 0; k < off; ++k) {
            const float send = upper ? v[k] : v[k + off];
            const float keep = upper ? v[k + off] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
    }
    return v[0];
}

template <typename T, int VN> struct AltVec;
template <> struct AltVec<float, 4> {
    float f[4];
    __device__ __forceinline__ void load(const float* p) {
        const float4 q = __ldg(reinterpret_cast<const float4*>(p));
        f[0] = q.x; f[1] = q.y; f[2] = q.z; f[3] = q.w;
    }
};
template <> struct AltVec<__half, 8> {
    float f[8];
    __device__ __forceinline__ void load(const __half* p) {
        const uint4 q = __ldg(reinterpret_cast<const uint4*>(p));
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float2 t = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
            f[2 * i] = t.x; f[2 * i + 1] = t.y;
        }
    }
};
template <> struct AltVec<__half, 4> {
    float f[4];
    __device__ __forceinline__ void load(const __half* p) {
        const uint2 q = __ldg(reinterpret_cast<const uint2*>(p));
        const uint32_t w[2] = {q.x, q.y};
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const float2 t = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
            f[2 * i] = t.x; f[2 * i + 1] = t.y;
        }
    }
};

// VN = channels per lane and load, NCH = number of per-lane channel chunks: C = NCH * 32 * VN
template <typename T, int VN, int NCH>
__global__ void __launch_bounds__(AC_WARPS * 32)
alt_corr_forward_kernel(const AltParams p) {
    __shared__ float s_dot[AC_WARPS][AC_MAXT + 28];     // per warp: s of the current pixel (padded to 4 x 32)
    __shared__ float s_out[AC_MAXO][AC_PIX + 1];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int HW = p.H1 * p.W1;
    const int bn = blockIdx.y, b = bn / p.N;
    const int pix0 = blockIdx.x * AC_PIX;
    const int r = p.r, rd = 2 * r + 1, side = rd + 1, n_tex = side * side, n_out = rd * rd;
    const T* f1 = static_cast<const T*>(p.fmap1) + (size_t)b * HW * p.C;
    const T* f2 = static_cast<const T*>(p.fmap2) + (size_t)b * p.H2 * p.W2 * p.C;

    for (int q = warp; q < AC_PIX; q += AC_WARPS) {
        const int pix = pix0 + q;
        if (pix >= HW) break;                                        // warp-uniform
        const float x = __ldg(p.coords + ((size_t)bn * HW + pix) * 2);
        const float y = __ldg(p.coords + ((size_t)bn * HW + pix) * 2 + 1);
        const bool finite = fabsf(x) < 1e6f && fabsf(y) < 1e6f;     // NaN / inf / absurd -> every texel out of range
        const float flx = finite ? floorf(x) : -1e6f, fly = finite ? floorf(y) : -1e6f;
        const int x0 = (int)flx - r, y0 = (int)fly - r;
        const float dx = finite ? x - flx : 0.f, dy = finite ? y - fly : 0.f;

        AltVec<T, VN> a[NCH];
#pragma unroll
        for (int c = 0; c < NCH; ++c) a[c].load(f1 + (size_t)pix * p.C + (c * 32 + lane) * VN);

        for (int g = 0; g < n_tex; g += 32) {
            float v[32];
#pragma unroll
            for (int k = 0; k < 32; ++k) {
                const int t = g + k;
                const int iy = t / side, ix = t - iy * side;
                const int h2 = y0 + iy, w2 = x0 + ix;
                float acc = 0.f;
                if (t < n_tex && h2 >= 0 && h2 < p.H2 && w2 >= 0 && w2 < p.W2) {      // warp-uniform
                    const T* row = f2 + ((size_t)h2 * p.W2 + w2) * p.C;
#pragma unroll
                    for (int c = 0; c < NCH; ++c) {
                        AltVec<T, VN> bv;
                        bv.load(row + (c * 32 + lane) * VN);
#pragma unroll
                        for (int e = 0; e < VN; ++e) acc = fmaf(a[c].f[e], bv.f[e], acc);
                    }
                }
                v[k] = acc;
            }
            s_dot[warp][g + lane] = transpose_reduce32(v, lane);
        }
        __syncwarp();
        const float w00 = (1.f - dy) * (1.f - dx), w01 = (1.f - dy) * dx, w10 = dy * (1.f - dx), w11 = dy * dx;
        for (int o = lane; o < n_out; o += 32) {
            const int ix = o / rd, iy = o - ix * rd;
            const float* sd = &s_dot[warp][iy * side + ix];
            s_out[o][q] = w00 * sd[0] + w01 * sd[1] + w10 * sd[side] + w11 * sd[side + 1];
        }
        __syncwarp();
    }
    __syncthreads();
    // corr[b, n, o, pix]: 32 consecutive pixels per output channel
    float* out = p.corr + (size_t)bn * n_out * HW;
    for (int o = warp; o < n_out; o += AC_WARPS)
        if (pix0 + lane < HW) out[(size_t)o * HW + pix0 + lane] = s_out[o][lane];
}

template <typename T>
static int alt_launch(const AltParams& p, cudaStream_t st) {
    constexpr int VMAX = 16 / (int)sizeof(T);            // 16-byte loads
    const dim3 grid(cdiv(p.H1 * p.W1, AC_PIX), p.B * p.N);
    const int per_lane = p.C % 32 == 0 ? p.C / 32 : 0;   // channels per lane
    if (per_lane == VMAX) alt_corr_forward_kernel<T, VMAX, 1><<<grid, AC_WARPS * 32, 0, st>>>(p);
    else if (per_lane == 2 * VMAX) alt_corr_forward_kernel<T, VMAX, 2><<<grid, AC_WARPS * 32, 0, st>>>(p);
    else if (per_lane == 4 * VMAX) alt_corr_forward_kernel<T, VMAX, 4><<<grid, AC_WARPS * 32, 0, st>>>(p);
    else if (per_lane == 4) alt_corr_forward_kernel<T, 4, 1><<<grid, AC_WARPS * 32, 0, st>>>(p);
    else {
        set_error("vfml_alt_corr_forward: C=%d unsupported (C must be 128, 256, 512%s)", p.C, sizeof(T) == 2 ? " or 1024" : "");
        return VFML_ERR_ARG;
    }
    VFML_LAUNCH_CHECK("alt_corr_forward_kernel");
    return VFML_OK;
}

}  // namespace vfml

using namespace vfml;

extern "C" int vfml_alt_corr_forward(const void* fmap1, const void* fmap2, const float* coords, float* corr, int B,
                                     int N, int H1, int W1, int H2, int W2, int C, int radius, int fmap_dtype,
                                     vfml_stream_t s) {
    VFML_REQUIRE(fmap1 && fmap2 && coords && corr, "vfml_alt_corr_forward: null pointer");
    VFML_REQUIRE(valid_dtype(fmap_dtype), "vfml_alt_corr_forward: bad dtype %d", fmap_dtype);
    VFML_REQUIRE(B > 0 && N > 0 && H1 > 0 && W1 > 0 && H2 > 0 && W2 > 0 && C > 0, "vfml_alt_corr_forward: bad shape");
    VFML_REQUIRE(radius >= 1 && radius <= AC_MAXR, "vfml_alt_corr_forward: radius %d outside 1..%d", radius, AC_MAXR);
    VFML_REQUIRE((int64_t)B * N <= 65535, "vfml_alt_corr_forward: B*N=%lld exceeds 65535", (long long)B * N);
    AltParams p{fmap1, fmap2, coords, corr, B, N, H1, W1, H2, W2, C, radius};
    return fmap_dtype == VFML_F16 ? alt_launch<__half>(p, to_stream(s)) : alt_launch<float>(p, to_stream(s));
}

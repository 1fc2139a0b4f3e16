// Non-GEMM pieces of the MOF update step in rows (NHWC) layout: layout conversion, depthwise
// convolution with fused residual + GELU, row softmax, the motion-encoder glue (flow
// differences, temporal hidden-state warp, convf1_), coordinate update, im2col for the mask head.
//
// MOFNetStack/stack.py:8-32 (PCBlock depthwise stage), :69-103 (motion encoder),
// gma.py:74 (softmax), network.py:183-197 (flow = coords1 - coords0, coords += delta).
#include "common.cuh"

namespace vfml {

// ---------------------------------------------------------------- layout conversion
template <typename T>
__global__ void nchw_to_rows_kernel(const float* __restrict__ src, T* __restrict__ dst, int c, int hw,
                                    int64_t pitch, int col) {
    __shared__ float tile[32][33];
    const int n = blockIdx.z;
    const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;   // 32 x 8
    for (int i = ty; i < 32; i += 8) {
        const int ch = c0 + i, p = p0 + tx;
        tile[i][tx] = (ch < c && p < hw) ? src[((int64_t)n * c + ch) * hw + p] : 0.0f;
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
        const int p = p0 + i, ch = c0 + tx;
        if (p < hw && ch < c) dst[((int64_t)n * hw + p) * pitch + col + ch] = from_f32<T>(tile[tx][i]);
    }
}

template <typename T>
__global__ void rows_to_nchw_kernel(const T* __restrict__ src, float* __restrict__ dst, int c, int hw,
                                    int64_t pitch, int col) {
    __shared__ float tile[32][33];
    const int n = blockIdx.z;
    const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;
    for (int i = ty; i < 32; i += 8) {
        const int p = p0 + i, ch = c0 + tx;
        tile[i][tx] = (p < hw && ch < c) ? to_f32<T>(src[((int64_t)n * hw + p) * pitch + col + ch]) : 0.0f;
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
        const int ch = c0 + i, p = p0 + tx;
        if (ch < c && p < hw) dst[((int64_t)n * c + ch) * hw + p] = tile[tx][i];
    }
}

// ---------------------------------------------------------------- depthwise conv + residual + GELU
// thread per (pixel, channel), channel fastest; weights tap-major [k*k][C] so both loads coalesce.
template <typename T>
__global__ void __launch_bounds__(256)
dwconv_res_gelu_kernel(const T* __restrict__ x, T* __restrict__ y, int n_img, int h, int w, int C,
                       int64_t px, int64_t py, const float* __restrict__ wgt, const float* __restrict__ bias, int k) {
    const int64_t total = (int64_t)n_img * h * w * C;
    const int r = k / 2;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(t % C);
        const int64_t row = t / C;
        const int xx = (int)(row % w);
        const int yy = (int)((row / w) % h);
        const int64_t img_row0 = row - ((int64_t)yy * w + xx);
        float acc = 0.0f;
        for (int ky = 0; ky < k; ++ky) {
            const int sy = yy + ky - r;
            if (sy < 0 || sy >= h) continue;
            for (int kx = 0; kx < k; ++kx) {
                const int sx = xx + kx - r;
                if (sx < 0 || sx >= w) continue;
                acc = fmaf(to_f32<T>(x[(img_row0 + (int64_t)sy * w + sx) * px + c]), __ldg(wgt + (ky * k + kx) * C + c), acc);
            }
        }
        const float v = to_f32<T>(x[row * px + c]) + (acc + __ldg(bias + c));
        y[row * py + c] = from_f32<T>(gelu_erf(v));
    }
}

// ---------------------------------------------------------------- row softmax
template <typename TO>
__global__ void __launch_bounds__(256)
softmax_rows_kernel(const float* __restrict__ in, TO* __restrict__ out, int n, int64_t ld_in, int64_t ld_out) {
    __shared__ float red[8];
    __shared__ float bcast;
    const float* src = in + blockIdx.x * ld_in;
    TO* dst = out + blockIdx.x * ld_out;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float mx = -INFINITY;
    for (int i = threadIdx.x; i < n; i += 256) mx = fmaxf(mx, src[i]);
    for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) red[warp] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        float m = red[0];
        for (int i = 1; i < 8; ++i) m = fmaxf(m, red[i]);
        bcast = m;
    }
    __syncthreads();
    mx = bcast;
    float sum = 0.0f;
    for (int i = threadIdx.x; i < n; i += 256) sum += expf(src[i] - mx);
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    __syncthreads();
    if (lane == 0) red[warp] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        float s = 0.0f;
        for (int i = 0; i < 8; ++i) s += red[i];
        bcast = s;
    }
    __syncthreads();
    const float inv = 1.0f / bcast;
    for (int i = threadIdx.x; i < n; i += 256) dst[i] = from_f32<TO>(expf(src[i] - mx) * inv);
}

// ---------------------------------------------------------------- motion-encoder glue
struct GlueParams {
    const float* fc; const float* bc;
    void* x400; int64_t p400;
    void* f128; int64_t p128;
    void* mf; int64_t pmf; int mf_col;
    const float* w1; const float* b1;   // convf1_: [128, 4], [128]
    int F, h, w;
    int rot;          // slot f holds the window's frame (f - rot) mod F (ring order, see vfml.h)
};

template <typename T>
__device__ __forceinline__ void warp_state(const T* x400, int64_t p400, int frame, int F, int h, int w, float sx, float sy,
                                           T* dst, int lane) {
    // bilinear_sampler, zeros padding, align_corners=True (core/utils/utils.py:62-76) over the 48-channel
    // hidden state of `frame` (zero state outside the window, stack.py:75-80)
    float acc0 = 0.0f, acc1 = 0.0f;   // channels lane, lane + 32 (< 48)
    if (frame >= 0 && frame < F && sx == sx && sy == sy) {
        sx = fminf(fmaxf(sx, -1.0e6f), 1.0e6f);
        sy = fminf(fmaxf(sy, -1.0e6f), 1.0e6f);
        const float flx = floorf(sx), fly = floorf(sy);
        const float wx1 = sx - flx, wy1 = sy - fly;
        const int x0 = (int)flx, y0 = (int)fly;
        const T* base = x400 + (int64_t)frame * h * w * p400 + 352;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int gx = x0 + (t & 1), gy = y0 + (t >> 1);
            if (gx < 0 || gx >= w || gy < 0 || gy >= h) continue;
            const float wgt = ((t & 1) ? wx1 : 1.0f - wx1) * ((t >> 1) ? wy1 : 1.0f - wy1);
            const T* src = base + ((int64_t)gy * w + gx) * p400;
            acc0 += wgt * to_f32<T>(src[lane]);
            if (lane < 16) acc1 += wgt * to_f32<T>(src[lane + 32]);
        }
    }
    dst[lane] = from_f32<T>(acc0);
    if (lane < 16) dst[lane + 32] = from_f32<T>(acc1);
}

template <typename T>
__global__ void __launch_bounds__(256) motion_glue_kernel(const GlueParams p) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // lets a PDL-launched successor start its prologue
    const int lane = threadIdx.x & 31;
    const int P = p.h * p.w;
    const int64_t total = (int64_t)p.F * P;
    const int64_t warp0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    T* x400 = reinterpret_cast<T*>(p.x400);
    for (int64_t row = warp0; row < total; row += nwarps) {
        const int f = (int)(row / P);
        const int pix = (int)(row - (int64_t)f * P);
        const int y = pix / p.w, x = pix - y * p.w;
        const float fcx = __ldg(p.fc + ((int64_t)f * 2) * P + pix), fcy = __ldg(p.fc + ((int64_t)f * 2 + 1) * P + pix);
        const float bcx = __ldg(p.bc + ((int64_t)f * 2) * P + pix), bcy = __ldg(p.bc + ((int64_t)f * 2 + 1) * P + pix);
        const float fl[4] = {fcx - (float)x, fcy - (float)y, bcx - (float)x, bcy - (float)y};
        T* xrow = x400 + row * p.p400;
        // temporal neighbours in window order; the window's first / last frame has none on that side (zero state)
        const int j = f >= p.rot ? f - p.rot : f - p.rot + p.F;
        const int f_next = j + 1 < p.F ? (f + 1 == p.F ? 0 : f + 1) : -1;
        const int f_prev = j > 0 ? (f == 0 ? p.F - 1 : f - 1) : -1;
        warp_state<T>(x400, p.p400, f_next, p.F, p.h, p.w, fcx, fcy, xrow + 256, lane);
        warp_state<T>(x400, p.p400, f_prev, p.F, p.h, p.w, bcx, bcy, xrow + 304, lane);
        T* frow = reinterpret_cast<T*>(p.f128) + row * p.p128;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int o = lane + 32 * q;
            const float4 wv = __ldg(reinterpret_cast<const float4*>(p.w1) + o);
            float v = __ldg(p.b1 + o);
            v = fmaf(wv.x, fl[0], v);
            v = fmaf(wv.y, fl[1], v);
            v = fmaf(wv.z, fl[2], v);
            v = fmaf(wv.w, fl[3], v);
            frow[o] = from_f32<T>(v);
        }
        if (lane < 4) reinterpret_cast<T*>(p.mf)[row * p.pmf + p.mf_col + 124 + lane] = from_f32<T>(fl[lane]);
    }
}

// out[r, c] = res[r, c] + sum_s partial[s, r, c]: combines the split-K partial products of the aggregate GEMM
// (gma.py:113-115 / memory read-out) with the residual.  One thread per 4 columns.
template <typename T>
__global__ void splitk_reduce_kernel(const float* __restrict__ partial, int S, int64_t M, int N, int64_t slab, const T* __restrict__ res,
                                     int64_t ldr, T* __restrict__ out, int64_t ldo) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // lets a PDL-launched successor start its prologue
    const int n4 = N >> 2;
    const int64_t total = M * n4;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = t / n4;
        const int c = (int)(t - r * n4) << 2;
        const float* pp = partial + r * N + c;
        float4 acc = *reinterpret_cast<const float4*>(pp);
        for (int s = 1; s < S; ++s) {
            const float4 v = *reinterpret_cast<const float4*>(pp + s * slab);
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        if (res) {
            const T* rr = res + r * ldr + c;
            acc.x += to_f32<T>(rr[0]); acc.y += to_f32<T>(rr[1]); acc.z += to_f32<T>(rr[2]); acc.w += to_f32<T>(rr[3]);
        }
        T* o = out + r * ldo + c;
        o[0] = from_f32<T>(acc.x); o[1] = from_f32<T>(acc.y); o[2] = from_f32<T>(acc.z); o[3] = from_f32<T>(acc.w);
    }
}

// LayerNorm over the last dimension of a rows buffer (Twins blocks: twins_ft.py norm1 / norm2, C = 128 or 256): one warp
// per row, the row lives in registers (VPL 8-element vectors per lane), two-pass mean / variance in fp32 like torch.
template <typename T, int VPL>
__global__ void __launch_bounds__(256) layer_norm_rows_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t rows, int C, int64_t ldx,
                                                              int64_t ldy, const T* __restrict__ gamma, const T* __restrict__ beta, float eps) {
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5, nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    constexpr int E = 16 / (int)sizeof(T);        // elements per 16-byte vector
    for (int64_t r = warp0; r < rows; r += nwarps) {
        float v[VPL * E];
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int c = (i * 32 + lane) * E;
            if (c < C) {
                const uint4 q = *reinterpret_cast<const uint4*>(x + r * ldx + c);
                const T* e = reinterpret_cast<const T*>(&q);
#pragma unroll
                for (int j = 0; j < E; ++j) v[i * E + j] = to_f32<T>(e[j]);
            } else {
#pragma unroll
                for (int j = 0; j < E; ++j) v[i * E + j] = 0.f;
            }
        }
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < VPL * E; ++i) s += v[i];
#pragma unroll
        for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        const float mean = s / (float)C;
        float q2 = 0.f;
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int c = (i * 32 + lane) * E;
            if (c < C) {
#pragma unroll
                for (int j = 0; j < E; ++j) { const float d = v[i * E + j] - mean; q2 += d * d; }
            }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) q2 += __shfl_xor_sync(0xffffffffu, q2, o);
        const float rstd = rsqrtf(q2 / (float)C + eps);
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int c = (i * 32 + lane) * E;
            if (c < C) {
                const uint4 g4 = *reinterpret_cast<const uint4*>(gamma + c), b4 = *reinterpret_cast<const uint4*>(beta + c);
                const T* g = reinterpret_cast<const T*>(&g4);
                const T* b = reinterpret_cast<const T*>(&b4);
                uint4 o4;
                T* o = reinterpret_cast<T*>(&o4);
#pragma unroll
                for (int j = 0; j < E; ++j) o[j] = from_f32<T>((v[i * E + j] - mean) * rstd * to_f32<T>(g[j]) + to_f32<T>(b[j]));
                *reinterpret_cast<uint4*>(y + r * ldy + c) = o4;
            }
        }
    }
}

// Positional-encoding generator of the Twins encoder (PosConv, twins_ft.py:341-355; encoders.py:27-29,35-37):
// y = x + dwconv3x3(x) + b on token-major rows [n * h * w, C], zero padding.  One thread = one pixel x one 16-byte channel
// vector, fp32 accumulation; weights tap-major [9, C] in the activations' dtype.  (Also replaces a cuDNN depthwise kernel
// that returned wrong values for batch 1 / 20x20 / 256 channels in fp16 on this stack.)
template <typename T>
__global__ void __launch_bounds__(256) peg3x3_kernel(const T* __restrict__ x, T* __restrict__ y, int n_img, int h, int w, int C,
                                                     int64_t ldx, int64_t ldy, const T* __restrict__ wgt, const T* __restrict__ bias) {
    constexpr int E = 16 / (int)sizeof(T);
    const int cv = C / E;
    const int64_t total = (int64_t)n_img * h * w * cv;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % cv) * E;
        const int64_t pix = i / cv;
        const int px = (int)(pix % w), py = (int)((pix / w) % h);
        float acc[E];
        {
            const uint4 b4 = *reinterpret_cast<const uint4*>(bias + c);
            const T* b = reinterpret_cast<const T*>(&b4);
#pragma unroll
            for (int j = 0; j < E; ++j) acc[j] = to_f32<T>(b[j]);
        }
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
            const int yy = py + ky - 1;
            if (yy < 0 || yy >= h) continue;
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
                const int xx = px + kx - 1;
                if (xx < 0 || xx >= w) continue;
                const uint4 v4 = *reinterpret_cast<const uint4*>(x + (pix + (int64_t)(ky - 1) * w + (kx - 1)) * ldx + c);
                const uint4 w4 = *reinterpret_cast<const uint4*>(wgt + (ky * 3 + kx) * C + c);
                const T* v = reinterpret_cast<const T*>(&v4);
                const T* wv = reinterpret_cast<const T*>(&w4);
#pragma unroll
                for (int j = 0; j < E; ++j) acc[j] = fmaf(to_f32<T>(v[j]), to_f32<T>(wv[j]), acc[j]);
            }
        }
        const uint4 c4 = *reinterpret_cast<const uint4*>(x + pix * ldx + c);
        const T* cv_ = reinterpret_cast<const T*>(&c4);
        uint4 o4;
        T* o = reinterpret_cast<T*>(&o4);
#pragma unroll
        for (int j = 0; j < E; ++j) o[j] = from_f32<T>(to_f32<T>(cv_[j]) + acc[j]);
        *reinterpret_cast<uint4*>(y + pix * ldy + c) = o4;
    }
}

__global__ void coords_add_kernel(float* __restrict__ fc, float* __restrict__ bc, const float* __restrict__ delta,
                                  int64_t ld, int F, int hw) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // lets a PDL-launched successor start its prologue
    const int64_t total = (int64_t)F * hw;
    for (int64_t row = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; row < total; row += (int64_t)gridDim.x * blockDim.x) {
        const int f = (int)(row / hw);
        const int pix = (int)(row - (int64_t)f * hw);
        const float4 d = *reinterpret_cast<const float4*>(delta + row * ld);
        fc[((int64_t)f * 2) * hw + pix] += d.x;
        fc[((int64_t)f * 2 + 1) * hw + pix] += d.y;
        bc[((int64_t)f * 2) * hw + pix] += d.z;
        bc[((int64_t)f * 2 + 1) * hw + pix] += d.w;
    }
}

__global__ void coords_grid_kernel(float* __restrict__ c, int n, int h, int w) {
    const int64_t total = (int64_t)n * 2 * h * w;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int x = (int)(t % w);
        const int y = (int)((t / w) % h);
        const int ch = (int)((t / ((int64_t)w * h)) % 2);
        c[t] = ch == 0 ? (float)x : (float)y;
    }
}

template <typename T>
__global__ void rows_fill_kernel(T* __restrict__ rows, int64_t pitch, int col, int C, int64_t n_rows, const float* __restrict__ v) {
    const int64_t total = n_rows * C;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(t % C);
        rows[(t / C) * pitch + col + c] = from_f32<T>(v[c]);
    }
}

template <typename T>
__global__ void im2col3x3_kernel(const T* __restrict__ x, T* __restrict__ out, int n_img, int h, int w, int C,
                                 int64_t px, int64_t po) {
    const int64_t total = (int64_t)n_img * h * w * 9 * C;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(t % C);
        const int tap = (int)((t / C) % 9);
        const int64_t row = t / ((int64_t)9 * C);
        const int xx = (int)(row % w), yy = (int)((row / w) % h);
        const int sy = yy + tap / 3 - 1, sx = xx + tap % 3 - 1;
        T v = from_f32<T>(0.0f);
        if (sy >= 0 && sy < h && sx >= 0 && sx < w) v = x[(row + (int64_t)(sy - yy) * w + (sx - xx)) * px + c];
        out[row * po + tap * C + c] = v;
    }
}

static inline int ew_blocks(int64_t n) { return (int)std::min<int64_t>((n + 255) / 256, device_sms() * 32); }

int dwconv_tiled(const void* x, void* y, int n_img, int h, int w, int C, int64_t px, int64_t py, const float* wgt,
                 const float* bias, int k, int dtype, cudaStream_t st);
int gemm_ffma(const vfml_gemm_desc& d, cudaStream_t st);
int gemm_tc(const vfml_gemm_desc& d, cudaStream_t st);
bool gemm_tc_supported(const vfml_gemm_desc& d);
int gemm_ws_run_choice(const vfml_gemm_desc& d, cudaStream_t st, int choice);
bool gemm_ws_supported(const vfml_gemm_desc& d);
int gemm_ws_choice(const vfml_gemm_desc& d);
int gemm_autotune(const vfml_gemm_desc& d, cudaStream_t st);
int validate_gemm(const vfml_gemm_desc* d);

}  // namespace vfml

using namespace vfml;

#define DISPATCH_T(dtype, CALL_F32, CALL_F16) \
    do {                                      \
        if ((dtype) == VFML_F32) { CALL_F32; } \
        else { CALL_F16; }                    \
    } while (0)

extern "C" int vfml_nchw_to_rows(const float* src, void* dst, int n, int c, int hw, int dst_pitch, int dst_col,
                                 int dtype, vfml_stream_t s) {
    VFML_REQUIRE(n >= 0 && c >= 0 && hw >= 0 && dst_col >= 0 && dst_pitch >= dst_col + c, "nchw_to_rows: bad shape");
    VFML_REQUIRE(valid_dtype(dtype), "nchw_to_rows: bad dtype");
    if ((int64_t)n * c * hw == 0) return VFML_OK;
    VFML_REQUIRE(src && dst, "nchw_to_rows: null pointer");
    dim3 grid(cdiv(hw, 32), cdiv(c, 32), n), block(32, 8);
    DISPATCH_T(dtype, (nchw_to_rows_kernel<float><<<grid, block, 0, to_stream(s)>>>(src, (float*)dst, c, hw, dst_pitch, dst_col)),
               (nchw_to_rows_kernel<__half><<<grid, block, 0, to_stream(s)>>>(src, (__half*)dst, c, hw, dst_pitch, dst_col)));
    VFML_LAUNCH_CHECK("nchw_to_rows");
    return VFML_OK;
}

extern "C" int vfml_rows_to_nchw(const void* src, float* dst, int n, int c, int hw, int src_pitch, int src_col,
                                 int dtype, vfml_stream_t s) {
    VFML_REQUIRE(n >= 0 && c >= 0 && hw >= 0 && src_col >= 0 && src_pitch >= src_col + c, "rows_to_nchw: bad shape");
    VFML_REQUIRE(valid_dtype(dtype), "rows_to_nchw: bad dtype");
    if ((int64_t)n * c * hw == 0) return VFML_OK;
    VFML_REQUIRE(src && dst, "rows_to_nchw: null pointer");
    dim3 grid(cdiv(hw, 32), cdiv(c, 32), n), block(32, 8);
    DISPATCH_T(dtype, (rows_to_nchw_kernel<float><<<grid, block, 0, to_stream(s)>>>((const float*)src, dst, c, hw, src_pitch, src_col)),
               (rows_to_nchw_kernel<__half><<<grid, block, 0, to_stream(s)>>>((const __half*)src, dst, c, hw, src_pitch, src_col)));
    VFML_LAUNCH_CHECK("rows_to_nchw");
    return VFML_OK;
}

extern "C" int vfml_gemm(const vfml_gemm_desc* d, vfml_stream_t s) {
    int rc = validate_gemm(d);
    if (rc) return rc;
    if (d->M == 0 || d->N == 0) return VFML_OK;
    // fp16 operands with TMA-compatible alignment -> tcgen05 tensor-core kernel; fp32 (exact tier) and
    // oddly aligned fp16 views -> FFMA kernel.  Both are libvfml kernels with identical semantics.
    if (gemm_ws_supported(*d)) {           // the update block's big pointwise layers: the cluster kernel where it measured faster
        const int choice = gemm_ws_choice(*d);          // 0: gemm_tc (also: never tuned), > 0: a gemm_ws mode + cluster size, -1: pinned by the environment
        if (choice != 0) {
            rc = gemm_ws_run_choice(*d, to_stream(s), choice);
            if (rc <= 0) return rc;        // > 0: declined (weight slice does not fit), fall through
        }
    }
    if (gemm_tc_supported(*d)) return gemm_tc(*d, to_stream(s));
    return gemm_ffma(*d, to_stream(s));
}

extern "C" int vfml_gemm_autotune(const vfml_gemm_desc* d, vfml_stream_t s) {
    int rc = validate_gemm(d);
    if (rc) return rc;
    if (d->M == 0 || d->N == 0) return 0;
    return gemm_autotune(*d, to_stream(s));
}

extern "C" int vfml_dwconv_res_gelu(const void* x, void* y, int n_img, int h, int w, int C, int64_t pitch_x,
                                    int64_t pitch_y, const float* weight, const float* bias, int k, int dtype,
                                    vfml_stream_t s) {
    VFML_REQUIRE(n_img >= 0 && h > 0 && w > 0 && C > 0 && pitch_x >= C && pitch_y >= C, "dwconv: bad shape");
    VFML_REQUIRE(k >= 1 && (k & 1), "dwconv: kernel size %d must be odd", k);
    VFML_REQUIRE(valid_dtype(dtype), "dwconv: bad dtype");
    if (n_img == 0) return VFML_OK;
    VFML_REQUIRE(x && y && weight && bias && x != y, "dwconv: null or aliased pointer");
    {   // k = 15 / 7 with vector-aligned channels: shared-memory tiled kernel (dwconv.cu)
        const int rc = dwconv_tiled(x, y, n_img, h, w, C, pitch_x, pitch_y, weight, bias, k, dtype, to_stream(s));
        if (rc < 0) return rc;
        if (rc == 1) return VFML_OK;
    }
    const int64_t total = (int64_t)n_img * h * w * C;
    DISPATCH_T(dtype, (dwconv_res_gelu_kernel<float><<<ew_blocks(total), 256, 0, to_stream(s)>>>((const float*)x, (float*)y, n_img, h, w, C, pitch_x, pitch_y, weight, bias, k)),
               (dwconv_res_gelu_kernel<__half><<<ew_blocks(total), 256, 0, to_stream(s)>>>((const __half*)x, (__half*)y, n_img, h, w, C, pitch_x, pitch_y, weight, bias, k)));
    VFML_LAUNCH_CHECK("dwconv_res_gelu");
    return VFML_OK;
}

extern "C" int vfml_softmax_rows(const float* in, void* out, int64_t rows, int n, int64_t ld_in, int64_t ld_out,
                                 int out_dtype, vfml_stream_t s) {
    VFML_REQUIRE(rows >= 0 && n > 0 && ld_in >= n && ld_out >= n, "softmax_rows: bad shape");
    VFML_REQUIRE(valid_dtype(out_dtype), "softmax_rows: bad dtype");
    if (rows == 0) return VFML_OK;
    VFML_REQUIRE(in && out, "softmax_rows: null pointer");
    VFML_REQUIRE(rows <= 2147483647LL, "softmax_rows: too many rows");
    DISPATCH_T(out_dtype, (softmax_rows_kernel<float><<<(unsigned)rows, 256, 0, to_stream(s)>>>(in, (float*)out, n, ld_in, ld_out)),
               (softmax_rows_kernel<__half><<<(unsigned)rows, 256, 0, to_stream(s)>>>(in, (__half*)out, n, ld_in, ld_out)));
    VFML_LAUNCH_CHECK("softmax_rows");
    return VFML_OK;
}

extern "C" int vfml_splitk_reduce(const float* partial, int n_splits, int64_t rows, int n, const void* res, int64_t ld_res,
                                  void* out, int64_t ld_out, int dtype, vfml_stream_t s) {
    VFML_REQUIRE(n_splits >= 1 && rows >= 0 && n > 0 && (n % 4) == 0 && ld_out >= n && (!res || ld_res >= n), "splitk_reduce: bad shape (n %% 4 == 0)");
    VFML_REQUIRE(valid_dtype(dtype), "splitk_reduce: bad dtype");
    if (rows == 0) return VFML_OK;
    VFML_REQUIRE(partial && out, "splitk_reduce: null pointer");
    VFML_REQUIRE((reinterpret_cast<uintptr_t>(partial) & 15) == 0, "splitk_reduce: partial must be 16-byte aligned");
    const int64_t total = rows * (n / 4), slab = rows * n;
    DISPATCH_T(dtype, (splitk_reduce_kernel<float><<<ew_blocks(total), 256, 0, to_stream(s)>>>(partial, n_splits, rows, n, slab, (const float*)res, ld_res, (float*)out, ld_out)),
               (splitk_reduce_kernel<__half><<<ew_blocks(total), 256, 0, to_stream(s)>>>(partial, n_splits, rows, n, slab, (const __half*)res, ld_res, (__half*)out, ld_out)));
    VFML_LAUNCH_CHECK("splitk_reduce");
    return VFML_OK;
}

extern "C" int vfml_layer_norm_rows(const void* x, void* y, int64_t rows, int C, int64_t ld_x, int64_t ld_y, const void* gamma,
                                    const void* beta, float eps, int dtype, vfml_stream_t s) {
    VFML_REQUIRE(valid_dtype(dtype), "layer_norm_rows: bad dtype");
    const int E = dtype == VFML_F16 ? 8 : 4;
    VFML_REQUIRE(rows >= 0 && C > 0 && (C % E) == 0 && C <= 32 * 4 * E && ld_x >= C && ld_y >= C && (ld_x % E) == 0 && (ld_y % E) == 0,
                 "layer_norm_rows: C must be a multiple of %d and <= %d, pitches multiples of %d", E, 32 * 4 * E, E);
    if (rows == 0) return VFML_OK;
    VFML_REQUIRE(x && y && gamma && beta, "layer_norm_rows: null pointer");
    auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    VFML_REQUIRE(al(x) && al(y) && al(gamma) && al(beta), "layer_norm_rows: pointers must be 16-byte aligned");
    const int vpl = (C / E + 31) / 32;           // 16-byte vectors per lane
    const int blocks = (int)std::min<int64_t>(cdiv(rows, 8), device_sms() * 16);
#define VFML_LN_CASE(V)                                                                                                             \
    case V:                                                                                                                         \
        DISPATCH_T(dtype, (layer_norm_rows_kernel<float, V><<<blocks, 256, 0, to_stream(s)>>>((const float*)x, (float*)y, rows, C, ld_x, ld_y, (const float*)gamma, (const float*)beta, eps)), \
                   (layer_norm_rows_kernel<__half, V><<<blocks, 256, 0, to_stream(s)>>>((const __half*)x, (__half*)y, rows, C, ld_x, ld_y, (const __half*)gamma, (const __half*)beta, eps))); \
        break;
    switch (vpl) {
        VFML_LN_CASE(1)
        VFML_LN_CASE(2)
        VFML_LN_CASE(3)
        VFML_LN_CASE(4)
    }
#undef VFML_LN_CASE
    VFML_LAUNCH_CHECK("layer_norm_rows");
    return VFML_OK;
}

extern "C" int vfml_peg3x3(const void* x, void* y, int n_img, int h, int w, int C, int64_t ld_x, int64_t ld_y, const void* weight,
                           const void* bias, int dtype, vfml_stream_t s) {
    VFML_REQUIRE(valid_dtype(dtype), "peg3x3: bad dtype");
    const int E = dtype == VFML_F16 ? 8 : 4;
    VFML_REQUIRE(n_img >= 0 && h > 0 && w > 0 && C > 0 && (C % E) == 0 && ld_x >= C && ld_y >= C && (ld_x % E) == 0 && (ld_y % E) == 0,
                 "peg3x3: C and the pitches must be multiples of %d", E);
    if (n_img == 0) return VFML_OK;
    VFML_REQUIRE(x && y && weight && bias && x != y, "peg3x3: null pointer or in-place call");
    auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    VFML_REQUIRE(al(x) && al(y) && al(weight) && al(bias), "peg3x3: pointers must be 16-byte aligned");
    const int64_t total = (int64_t)n_img * h * w * (C / E);
    const int blocks = (int)std::min<int64_t>(cdiv(total, 256), device_sms() * 16);
    DISPATCH_T(dtype, (peg3x3_kernel<float><<<blocks, 256, 0, to_stream(s)>>>((const float*)x, (float*)y, n_img, h, w, C, ld_x, ld_y, (const float*)weight, (const float*)bias)),
               (peg3x3_kernel<__half><<<blocks, 256, 0, to_stream(s)>>>((const __half*)x, (__half*)y, n_img, h, w, C, ld_x, ld_y, (const __half*)weight, (const __half*)bias)));
    VFML_LAUNCH_CHECK("peg3x3");
    return VFML_OK;
}

extern "C" int vfml_motion_glue(const float* fcoords, const float* bcoords, void* x400, int64_t pitch400, void* f128,
                                int64_t pitch128, void* mf, int64_t pitch_mf, int mf_col, const float* convf1_w,
                                const float* convf1_b, int n_frames, int h, int w, int rot, int dtype, vfml_stream_t s) {
    VFML_REQUIRE(n_frames >= 0 && h > 0 && w > 0, "motion_glue: bad shape");
    VFML_REQUIRE(rot >= 0 && (rot < n_frames || rot == 0), "motion_glue: rot=%d not in [0, n_frames)", rot);
    VFML_REQUIRE(pitch400 >= 400 && pitch128 >= 128 && mf_col >= 0 && pitch_mf >= mf_col + 128, "motion_glue: bad pitch");
    VFML_REQUIRE(valid_dtype(dtype), "motion_glue: bad dtype");
    if (n_frames == 0) return VFML_OK;
    VFML_REQUIRE(fcoords && bcoords && x400 && f128 && mf && convf1_w && convf1_b, "motion_glue: null pointer");
    GlueParams p{fcoords, bcoords, x400, pitch400, f128, pitch128, mf, pitch_mf, mf_col, convf1_w, convf1_b, n_frames, h, w, rot};
    const int64_t total = (int64_t)n_frames * h * w;
    const int blocks = (int)std::min<int64_t>(cdiv(total, 8), device_sms() * 16);
    DISPATCH_T(dtype, (motion_glue_kernel<float><<<blocks, 256, 0, to_stream(s)>>>(p)),
               (motion_glue_kernel<__half><<<blocks, 256, 0, to_stream(s)>>>(p)));
    VFML_LAUNCH_CHECK("motion_glue");
    return VFML_OK;
}

extern "C" int vfml_coords_add(float* fcoords, float* bcoords, const float* delta, int64_t ld_delta, int n_frames,
                               int hw, vfml_stream_t s) {
    VFML_REQUIRE(n_frames >= 0 && hw >= 0 && ld_delta >= 4 && (ld_delta % 4) == 0, "coords_add: bad shape (ld %% 4 == 0)");
    if ((int64_t)n_frames * hw == 0) return VFML_OK;
    VFML_REQUIRE(fcoords && bcoords && delta, "coords_add: null pointer");
    coords_add_kernel<<<ew_blocks((int64_t)n_frames * hw), 256, 0, to_stream(s)>>>(fcoords, bcoords, delta, ld_delta, n_frames, hw);
    VFML_LAUNCH_CHECK("coords_add");
    return VFML_OK;
}

extern "C" int vfml_coords_grid(float* coords, int n, int h, int w, vfml_stream_t s) {
    VFML_REQUIRE(n >= 0 && h >= 0 && w >= 0, "coords_grid: bad shape");
    if ((int64_t)n * h * w == 0) return VFML_OK;
    VFML_REQUIRE(coords, "coords_grid: null pointer");
    coords_grid_kernel<<<ew_blocks((int64_t)n * 2 * h * w), 256, 0, to_stream(s)>>>(coords, n, h, w);
    VFML_LAUNCH_CHECK("coords_grid");
    return VFML_OK;
}

extern "C" int vfml_rows_fill(void* rows, int64_t pitch, int col, int C, int64_t n_rows, const float* values, int dtype,
                              vfml_stream_t s) {
    VFML_REQUIRE(n_rows >= 0 && C >= 0 && col >= 0 && pitch >= col + C, "rows_fill: bad shape");
    VFML_REQUIRE(valid_dtype(dtype), "rows_fill: bad dtype");
    if (n_rows * C == 0) return VFML_OK;
    VFML_REQUIRE(rows && values, "rows_fill: null pointer");
    DISPATCH_T(dtype, (rows_fill_kernel<float><<<ew_blocks(n_rows * C), 256, 0, to_stream(s)>>>((float*)rows, pitch, col, C, n_rows, values)),
               (rows_fill_kernel<__half><<<ew_blocks(n_rows * C), 256, 0, to_stream(s)>>>((__half*)rows, pitch, col, C, n_rows, values)));
    VFML_LAUNCH_CHECK("rows_fill");
    return VFML_OK;
}

extern "C" int vfml_im2col3x3(const void* x, void* out, int n_img, int h, int w, int C, int64_t pitch_x, int64_t pitch_out,
                              int dtype, vfml_stream_t s) {
    VFML_REQUIRE(n_img >= 0 && h > 0 && w > 0 && C > 0 && pitch_x >= C && pitch_out >= 9 * C, "im2col3x3: bad shape");
    VFML_REQUIRE(valid_dtype(dtype), "im2col3x3: bad dtype");
    if (n_img == 0) return VFML_OK;
    VFML_REQUIRE(x && out, "im2col3x3: null pointer");
    const int64_t total = (int64_t)n_img * h * w * 9 * C;
    DISPATCH_T(dtype, (im2col3x3_kernel<float><<<ew_blocks(total), 256, 0, to_stream(s)>>>((const float*)x, (float*)out, n_img, h, w, C, pitch_x, pitch_out)),
               (im2col3x3_kernel<__half><<<ew_blocks(total), 256, 0, to_stream(s)>>>((const __half*)x, (__half*)out, n_img, h, w, C, pitch_x, pitch_out)));
    VFML_LAUNCH_CHECK("im2col3x3");
    return VFML_OK;
}

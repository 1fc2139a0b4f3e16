// Flow post-pass: gamedev / motion-vector RG8 encoding (bit-exact with the numpy reference)
// and temporal anti-aliasing.  All arithmetic is spelled with round-to-nearest intrinsics
// (__fmul_rn, __fdiv_rn, __dadd_rn ...) so that nvcc can neither contract to FMA nor
// substitute approximate division: the u8 outputs must equal the reference's.
//
// encoding/flow_encoders.py:90-117 (gamedev), :131-177 (rg8 encode / decode)
// effects/taa_processor.py:41-262
#include "common.cuh"

namespace vfml {

// clip((c + clamp) / (2 clamp), 0, 1) * 255 -> nan_to_num -> truncate     flow_encoders.py:103-117
__device__ __forceinline__ uint8_t encode_channel(float v, float clamp, float two_clamp) {
    // np.clip propagates NaN; nan_to_num(nan=0) then maps it to 0
    if (v != v) return 0;
    v = fminf(fmaxf(v, -clamp), clamp);
    float e = __fdiv_rn(__fadd_rn(v, clamp), two_clamp);
    e = fminf(fmaxf(e, 0.0f), 1.0f);
    return (uint8_t)__fmul_rn(e, 255.0f);
}

__global__ void encode_gamedev_kernel(const float2* __restrict__ flow, uint8_t* __restrict__ rgb, int64_t npix,
                                      float nw, float nh, float scale, float clamp, float two_clamp) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < npix; i += (int64_t)gridDim.x * blockDim.x) {
        const float2 f = __ldg(flow + i);
        const float vx = __fmul_rn(__fdiv_rn(f.x, nw), scale);   // flow /= width; *= scale   (:97-101)
        const float vy = __fmul_rn(__fdiv_rn(f.y, nh), scale);
        uint8_t* o = rgb + 3 * i;
        o[0] = encode_channel(vx, clamp, two_clamp);
        o[1] = encode_channel(vy, clamp, two_clamp);
        o[2] = 0;
    }
}

__global__ void encode_rg8_kernel(const float2* __restrict__ flow, uint8_t* __restrict__ rgb, int64_t npix,
                                  float clamp, float two_clamp) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < npix; i += (int64_t)gridDim.x * blockDim.x) {
        const float2 f = __ldg(flow + i);
        uint8_t* o = rgb + 3 * i;
        o[0] = encode_channel(f.x, clamp, two_clamp);
        o[1] = encode_channel(f.y, clamp, two_clamp);
        o[2] = 0;
    }
}

__global__ void decode_rg8_kernel(const uint8_t* __restrict__ rgb, float2* __restrict__ flow, int64_t npix, float clamp) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < npix; i += (int64_t)gridDim.x * blockDim.x) {
        const uint8_t* s = rgb + 3 * i;
        // (img/255 * 2 * clamp) - clamp          flow_encoders.py:171-175
        const float nx = __fdiv_rn((float)s[0], 255.0f), ny = __fdiv_rn((float)s[1], 255.0f);
        float2 o;
        o.x = __fsub_rn(__fmul_rn(__fmul_rn(nx, 2.0f), clamp), clamp);
        o.y = __fsub_rn(__fmul_rn(__fmul_rn(ny, 2.0f), clamp), clamp);
        flow[i] = o;
    }
}

template <typename CurT>
__global__ void taa_init_kernel(const CurT* __restrict__ cur, double* __restrict__ out, int64_t n) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = (double)(float)cur[i];
}

struct TaaParams {
    const void* cur;        // [h, w, 3] u8, or fp32 (current_frame.astype(np.float32), taa_processor.py:64)
    const float* flow;
    const double* hist;
    double* out;
    int h, w;
    double alpha, one_minus_alpha, denom;   // denom = 2 * (sigma^2 * 0.1) + 1e-6 (taa_processor.py:196-199)
    int mode, hist_f32;
};

// reprojected coordinate: grid + flow, scrub, clamp (taa_processor.py:111-126); int64 grid + fp32 flow = f64
__device__ __forceinline__ double taa_coord(int g, float fl, int dim) {
    double v = __dadd_rn((double)g, (double)fl);
    if (v != v) v = 0.0;
    else if (isinf(v)) v = v > 0 ? (double)(dim - 1) : 0.0;
    return fmin(fmax(v, 0.0), (double)(dim - 1));
}

template <typename CurT>
__global__ void __launch_bounds__(256) taa_kernel(const TaaParams p) {
    const CurT* __restrict__ cur = static_cast<const CurT*>(p.cur);
    const int64_t npix = (int64_t)p.h * p.w;
    const float a32 = (float)p.alpha, b32 = (float)p.one_minus_alpha;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < npix; i += (int64_t)gridDim.x * blockDim.x) {
        const int y = (int)(i / p.w), x = (int)(i - (int64_t)y * p.w);
        float c[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) c[k] = (float)cur[3 * i + k];
        double* o = p.out + 3 * i;
        if (p.mode == 0) {
            // simple EMA (:78-80): alpha * cur + (1 - alpha) * prev in the history's dtype
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const double pv = p.hist[3 * i + k];
                if (p.hist_f32) o[k] = (double)__fadd_rn(__fmul_rn(a32, c[k]), __fmul_rn(b32, (float)pv));
                else o[k] = __dadd_rn((double)__fmul_rn(a32, c[k]), __dmul_rn(p.one_minus_alpha, pv));
            }
            continue;
        }
        const double px = taa_coord(x, p.flow[2 * i], p.w);
        const double py = taa_coord(y, p.flow[2 * i + 1], p.h);
        if (p.mode == 1) {
            // bilinear (:233-262): weights in f64, result buffer fp32, blend fp32
            int x0 = (int)floor(px), y0 = (int)floor(py);
            int x1 = min(x0 + 1, p.w - 1), y1 = min(y0 + 1, p.h - 1);
            x0 = min(max(x0, 0), p.w - 1); x1 = min(max(x1, 0), p.w - 1);
            y0 = min(max(y0, 0), p.h - 1); y1 = min(max(y1, 0), p.h - 1);
            const double wx = __dsub_rn(px, (double)x0), wy = __dsub_rn(py, (double)y0);
            const double ux = __dsub_rn(1.0, wx), uy = __dsub_rn(1.0, wy);
            const double* h00 = p.hist + 3 * ((int64_t)y0 * p.w + x0);
            const double* h01 = p.hist + 3 * ((int64_t)y0 * p.w + x1);
            const double* h10 = p.hist + 3 * ((int64_t)y1 * p.w + x0);
            const double* h11 = p.hist + 3 * ((int64_t)y1 * p.w + x1);
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                double r = __dmul_rn(__dmul_rn(h00[k], ux), uy);
                r = __dadd_rn(r, __dmul_rn(__dmul_rn(h01[k], wx), uy));
                r = __dadd_rn(r, __dmul_rn(__dmul_rn(h10[k], ux), wy));
                r = __dadd_rn(r, __dmul_rn(__dmul_rn(h11[k], wx), wy));
                const float rep = (float)r;
                o[k] = (double)__fadd_rn(__fmul_rn(a32, c[k]), __fmul_rn(b32, rep));
            }
            continue;
        }
        // bilateral reprojection (:167-216)
        const int x0 = min(max((int)floor(px), 0), p.w - 2), y0 = min(max((int)floor(py), 0), p.h - 2);
        const double wx = __dsub_rn(px, (double)x0), wy = __dsub_rn(py, (double)y0);
        const double ux = __dsub_rn(1.0, wx), uy = __dsub_rn(1.0, wy);
        const double* nb[4] = {p.hist + 3 * ((int64_t)y0 * p.w + x0), p.hist + 3 * ((int64_t)y0 * p.w + x0 + 1),
                               p.hist + 3 * ((int64_t)(y0 + 1) * p.w + x0), p.hist + 3 * ((int64_t)(y0 + 1) * p.w + x0 + 1)};
        const double bw[4] = {__dmul_rn(ux, uy), __dmul_rn(wx, uy), __dmul_rn(ux, wy), __dmul_rn(wx, wy)};
        // np.mean(float32 frame, axis=2): fp32 pairwise-free sum of 3, divided by 3
        const float lum = __fdiv_rn(__fadd_rn(__fadd_rn(c[0], c[1]), c[2]), 3.0f);
        double wgt[4], v[4][3];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
#pragma unroll
            for (int k = 0; k < 3; ++k) v[t][k] = nb[t][k];
            double cw;
            if (p.hist_f32) {
                const float ln = __fdiv_rn(__fadd_rn(__fadd_rn((float)v[t][0], (float)v[t][1]), (float)v[t][2]), 3.0f);
                const float d = __fsub_rn(lum, ln);
                cw = (double)expf(__fdiv_rn(-__fmul_rn(d, d), (float)p.denom));
            } else {
                const double ln = __ddiv_rn(__dadd_rn(__dadd_rn(v[t][0], v[t][1]), v[t][2]), 3.0);
                const double d = __dsub_rn((double)lum, ln);
                cw = exp(__ddiv_rn(-__dmul_rn(d, d), p.denom));
            }
            wgt[t] = __dmul_rn(bw[t], cw);
        }
        double tot = __dadd_rn(__dadd_rn(__dadd_rn(wgt[0], wgt[1]), wgt[2]), wgt[3]);
        if (tot == 0.0) tot = 1e-6;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            double r = __dmul_rn(v[0][k], wgt[0]);
            r = __dadd_rn(r, __dmul_rn(v[1][k], wgt[1]));
            r = __dadd_rn(r, __dmul_rn(v[2][k], wgt[2]));
            r = __dadd_rn(r, __dmul_rn(v[3][k], wgt[3]));
            r = __ddiv_rn(r, tot);
            o[k] = __dadd_rn((double)__fmul_rn(a32, c[k]), __dmul_rn(p.one_minus_alpha, r));
        }
    }
}

static inline int ew_blocks(int64_t n) { return (int)std::min<int64_t>((n + 255) / 256, device_sms() * 16); }

}  // namespace vfml

using namespace vfml;

extern "C" int vfml_encode_gamedev(const float* flow, uint8_t* rgb, int height, int width, float norm_width,
                                   float norm_height, float scale, float clamp, vfml_stream_t s) {
    VFML_REQUIRE(height >= 0 && width >= 0, "encode_gamedev: bad shape");
    const int64_t n = (int64_t)height * width;
    if (n == 0) return VFML_OK;
    VFML_REQUIRE(flow && rgb, "encode_gamedev: null pointer");
    encode_gamedev_kernel<<<ew_blocks(n), 256, 0, to_stream(s)>>>(reinterpret_cast<const float2*>(flow), rgb, n,
                                                                   norm_width, norm_height, scale, clamp, 2.0f * clamp);
    VFML_LAUNCH_CHECK("encode_gamedev");
    return VFML_OK;
}

extern "C" int vfml_encode_mv_rg8(const float* flow, uint8_t* rgb, int height, int width, float clamp, vfml_stream_t s) {
    VFML_REQUIRE(height >= 0 && width >= 0, "encode_mv_rg8: bad shape");
    const int64_t n = (int64_t)height * width;
    if (n == 0) return VFML_OK;
    VFML_REQUIRE(flow && rgb, "encode_mv_rg8: null pointer");
    encode_rg8_kernel<<<ew_blocks(n), 256, 0, to_stream(s)>>>(reinterpret_cast<const float2*>(flow), rgb, n, clamp, 2.0f * clamp);
    VFML_LAUNCH_CHECK("encode_mv_rg8");
    return VFML_OK;
}

extern "C" int vfml_decode_mv_rg8(const uint8_t* rgb, float* flow, int height, int width, float clamp, vfml_stream_t s) {
    VFML_REQUIRE(height >= 0 && width >= 0, "decode_mv_rg8: bad shape");
    const int64_t n = (int64_t)height * width;
    if (n == 0) return VFML_OK;
    VFML_REQUIRE(flow && rgb, "decode_mv_rg8: null pointer");
    decode_rg8_kernel<<<ew_blocks(n), 256, 0, to_stream(s)>>>(rgb, reinterpret_cast<float2*>(flow), n, clamp);
    VFML_LAUNCH_CHECK("decode_mv_rg8");
    return VFML_OK;
}

extern "C" int vfml_taa_init(const void* cur, int cur_f32, double* out, int h, int w, vfml_stream_t s) {
    VFML_REQUIRE(h >= 0 && w >= 0, "taa_init: bad shape");
    const int64_t n = (int64_t)h * w * 3;
    if (n == 0) return VFML_OK;
    VFML_REQUIRE(cur && out, "taa_init: null pointer");
    if (cur_f32) taa_init_kernel<float><<<ew_blocks(n), 256, 0, to_stream(s)>>>(static_cast<const float*>(cur), out, n);
    else taa_init_kernel<uint8_t><<<ew_blocks(n), 256, 0, to_stream(s)>>>(static_cast<const uint8_t*>(cur), out, n);
    VFML_LAUNCH_CHECK("taa_init");
    return VFML_OK;
}

extern "C" int vfml_taa(const void* cur, int cur_f32, const float* flow, const double* history, double* out, int h, int w,
                        double alpha, double sigma_color, int mode, int hist_f32, vfml_stream_t s) {
    VFML_REQUIRE(h >= 0 && w >= 0, "taa: bad shape");
    if ((int64_t)h * w == 0) return VFML_OK;
    VFML_REQUIRE(cur && history && out, "taa: null pointer");
    VFML_REQUIRE(mode >= 0 && mode <= 2, "taa: mode %d not in 0..2", mode);
    VFML_REQUIRE(mode == 0 || flow, "taa: flow modes need a flow field");
    VFML_REQUIRE(mode != 2 || (h >= 2 && w >= 2), "taa: bilateral mode needs at least 2x2 pixels");
    VFML_REQUIRE(history != out, "taa: out may not alias history");
    TaaParams p{cur, flow, history, out, h, w, alpha, 1.0 - alpha, 2.0 * (sigma_color * sigma_color * 0.1) + 1e-6, mode, hist_f32};
    if (cur_f32) taa_kernel<float><<<ew_blocks((int64_t)h * w), 256, 0, to_stream(s)>>>(p);
    else taa_kernel<uint8_t><<<ew_blocks((int64_t)h * w), 256, 0, to_stream(s)>>>(p);
    VFML_LAUNCH_CHECK("taa");
    return VFML_OK;
}

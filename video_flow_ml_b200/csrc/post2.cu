// Remaining flow encoders / decoders, the LOD pyramid and the flow-quality map.
// Same rules as post.cu: every float operation that decides an integer output is spelled with a
// round-to-nearest intrinsic (no FMA contraction, IEEE division and square root), because the u8
// results must equal the numpy / torch / OpenCV results of the reference bit for bit.
//
//   encoding/flow_encoders.py:191-295   MotionVectorsRGB8FlowEncoder.encode ('rgb+': module-level `method`, :10)
//   encoding/flow_encoders.py:297-364   MotionVectorsRGB8FlowEncoder.decode
//   encoding/flow_encoders.py:38-67     HSVFlowEncoder.encode (+ cv2.cvtColor HSV2RGB, third party, see oracle/post.py)
//   encoding/flow_encoders.py:399-427   TorchvisionFlowEncoder.encode (torchvision.utils.flow_to_image + the `* 255` wrap)
//   storage/cache_manager.py:77-161     LODGenerator.generate_lods
//   correction_worker.py:175-208        generate_quality_frame_gpu
#include "common.cuh"

namespace vfml {

static inline int ew_blocks2(int64_t n) { return (int)std::min<int64_t>((n + 255) / 256, (int64_t)device_sms() * 16); }

// rgb * 255 -> nan_to_num(nan=0, +inf=255, -inf=0) -> astype(uint8)          flow_encoders.py:292-295
__device__ __forceinline__ uint8_t to_u8_scrub(float v) {
    float x = __fmul_rn(v, 255.0f);
    if (x != x) return 0;
    if (isinf(x)) return x > 0 ? 255 : 0;
    return (uint8_t)(int)x;
}
// np.clip keeps NaN
__device__ __forceinline__ float clip_nan(float v, float lo, float hi) { return v != v ? v : fminf(fmaxf(v, lo), hi); }

__global__ void encode_rgb8_kernel(const float2* __restrict__ flow, uint8_t* __restrict__ rgb, int64_t npix, float clamp) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < npix; i += (int64_t)gridDim.x * blockDim.x) {
        const float2 f = __ldg(flow + i);
        float dx = __fdiv_rn(f.x, clamp), dy = __fdiv_rn(f.y, clamp);
        const float len = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
        if (len > 1.0f) {                      // NaN compares false: left alone, scrubbed below
            dx = __fdiv_rn(dx, len);
            dy = __fdiv_rn(dy, len);
        }
        const float corr = __fsqrt_rn(__fsub_rn(__fsub_rn(1.0f, __fmul_rn(dx, dx)), __fmul_rn(dy, dy)));   // NaN if negative
        uint8_t* o = rgb + 3 * i;
        o[0] = to_u8_scrub(__fdiv_rn(__fadd_rn(clip_nan(dx, -1.0f, 1.0f), 1.0f), 2.0f));
        o[1] = to_u8_scrub(__fdiv_rn(__fadd_rn(clip_nan(dy, -1.0f, 1.0f), 1.0f), 2.0f));
        o[2] = to_u8_scrub(corr);
    }
}

__global__ void decode_rgb8_kernel(const uint8_t* __restrict__ rgb, float2* __restrict__ flow, int64_t npix, float clamp) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < npix; i += (int64_t)gridDim.x * blockDim.x) {
        const uint8_t* s = rgb + 3 * i;
        const float n0 = __fdiv_rn((float)s[0], 255.0f), n1 = __fdiv_rn((float)s[1], 255.0f), n2 = __fdiv_rn((float)s[2], 255.0f);
        const float dx = __fsub_rn(__fmul_rn(n0, 2.0f), 1.0f), dy = __fsub_rn(__fmul_rn(n1, 2.0f), 1.0f);
        const float len = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(n2, n2)));
        const float mag = __fmul_rn(__fdiv_rn(1.0f, len), clamp);
        flow[i] = make_float2(__fmul_rn(dx, mag), __fmul_rn(dy, mag));
    }
}

// np.nan_to_num(flow, nan=0, posinf=1, neginf=-1)                               flow_encoders.py:41
__device__ __forceinline__ float scrub_hsv(float v) {
    if (v != v) return 0.0f;
    if (isinf(v)) return v > 0 ? 1.0f : -1.0f;
    return v;
}
__device__ __forceinline__ float mag_f32(float x, float y) { return __fsqrt_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y))); }

// max over pixels of the fp32 magnitude (HSV: after the scrub).  Non-negative floats order like their bit patterns and a
// canonical NaN is larger than +inf, so torch's NaN-propagating max (torchvision) falls out of an integer atomicMax.
__global__ void max_norm_kernel(const float2* __restrict__ flow, int64_t npix, int scrub, unsigned int* __restrict__ out) {
    unsigned int m = 0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < npix; i += (int64_t)gridDim.x * blockDim.x) {
        float2 f = __ldg(flow + i);
        if (scrub) { f.x = scrub_hsv(f.x); f.y = scrub_hsv(f.y); }
        const float g = mag_f32(f.x, f.y);
        m = max(m, g != g ? 0x7fc00000u : __float_as_uint(g));
    }
    for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}

// correctly rounded fp32 atan2 (through fp64): numpy / Sleef are within 1 ulp of this
__device__ __forceinline__ float atan2_f32(float y, float x) { return (float)atan2((double)y, (double)x); }

// cv2.cvtColor(HSV2RGB) for V = 255 (oracle/post.py hsv2rgb_u8): FMA where the x86 build fuses, truncation inside a row's
// 32-pixel SIMD blocks, round-half-even in the scalar tail of the row
__global__ void encode_hsv_kernel(const float2* __restrict__ flow, uint8_t* __restrict__ rgb, int height, int width,
                                  const unsigned int* __restrict__ max_bits) {
    const int64_t npix = (int64_t)height * width;
    const float mx = __uint_as_float(*max_bits);
    const int simd_cols = width / 32 * 32;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < npix; i += (int64_t)gridDim.x * blockDim.x) {
        float2 f = __ldg(flow + i);
        f.x = scrub_hsv(f.x); f.y = scrub_hsv(f.y);
        const float mag = mag_f32(f.x, f.y);
        const float ang = atan2_f32(f.y, f.x);
        // (angle + pi) / (2 pi) * 180, clip [0, 180], truncate                 :48-49 (python-float scalars stay fp32)
        float hue = __fmul_rn(__fdiv_rn(__fadd_rn(ang, 3.14159274101257324f), 6.28318548202514648f), 180.0f);
        hue = fminf(fmaxf(hue, 0.0f), 180.0f);
        const int H8 = (int)hue;
        const int S8 = mx > 0.0f ? (int)(uint8_t)(int)__fmul_rn(__fdiv_rn(mag, mx), 255.0f) : 0;       // :52-56
        // HSV -> RGB
        const float h = __fmul_rn((float)H8, 6.0f / 180.0f);
        const float s = __fmul_rn((float)S8, 1.0f / 255.0f);
        const float v = __fmul_rn(255.0f, 1.0f / 255.0f);
        int sec = (int)floorf(h);
        const float fr = __fsub_rn(h, (float)sec);
        if (sec >= 6) sec -= 6;
        float tab[4];
        tab[0] = v;
        tab[1] = __fmul_rn(v, __fsub_rn(1.0f, s));
        tab[2] = __fmul_rn(v, __fmaf_rn(-s, fr, 1.0f));
        tab[3] = __fmul_rn(v, __fmaf_rn(-s, __fsub_rn(1.0f, fr), 1.0f));
        // sector -> tab index of (r, g, b)
        const int ridx[6] = {0, 2, 1, 1, 3, 0}, gidx[6] = {3, 0, 0, 2, 1, 1}, bidx[6] = {1, 1, 3, 0, 0, 2};
        float r = tab[ridx[sec]], g = tab[gidx[sec]], b = tab[bidx[sec]];
        if (S8 == 0) r = g = b = v;
        const bool simd = (int)(i % width) < simd_cols;
        uint8_t* o = rgb + 3 * i;
        const float r255 = __fmul_rn(r, 255.0f), g255 = __fmul_rn(g, 255.0f), b255 = __fmul_rn(b, 255.0f);
        o[0] = (uint8_t)min(max(simd ? __float2int_rz(r255) : __float2int_rn(r255), 0), 255);
        o[1] = (uint8_t)min(max(simd ? __float2int_rz(g255) : __float2int_rn(g255), 0), 255);
        o[2] = (uint8_t)min(max(simd ? __float2int_rz(b255) : __float2int_rn(b255), 0), 255);
    }
}

// torchvision.utils._make_colorwheel: 55 RGB entries
__constant__ uint8_t c_wheel[55][3] = {
    {255, 0, 0}, {255, 17, 0}, {255, 34, 0}, {255, 51, 0}, {255, 68, 0}, {255, 85, 0}, {255, 102, 0}, {255, 119, 0},
    {255, 136, 0}, {255, 153, 0}, {255, 170, 0}, {255, 187, 0}, {255, 204, 0}, {255, 221, 0}, {255, 238, 0},
    {255, 255, 0}, {213, 255, 0}, {170, 255, 0}, {128, 255, 0}, {85, 255, 0}, {43, 255, 0},
    {0, 255, 0}, {0, 255, 63}, {0, 255, 127}, {0, 255, 191},
    {0, 255, 255}, {0, 232, 255}, {0, 209, 255}, {0, 186, 255}, {0, 163, 255}, {0, 140, 255}, {0, 116, 255}, {0, 93, 255},
    {0, 70, 255}, {0, 47, 255}, {0, 24, 255},
    {0, 0, 255}, {19, 0, 255}, {39, 0, 255}, {58, 0, 255}, {78, 0, 255}, {98, 0, 255}, {117, 0, 255}, {137, 0, 255},
    {156, 0, 255}, {176, 0, 255}, {196, 0, 255}, {215, 0, 255}, {235, 0, 255},
    {255, 0, 255}, {255, 0, 213}, {255, 0, 170}, {255, 0, 128}, {255, 0, 85}, {255, 0, 43}};

__global__ void encode_torchvision_kernel(const float2* __restrict__ flow, uint8_t* __restrict__ rgb, int64_t npix,
                                          const unsigned int* __restrict__ max_bits, int ref_wrap) {
    const float den = __fadd_rn(__uint_as_float(*max_bits), 1.1920928955078125e-07f);      // max_norm + finfo(float32).eps
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < npix; i += (int64_t)gridDim.x * blockDim.x) {
        const float2 f = __ldg(flow + i);
        const float u = __fdiv_rn(f.x, den), v = __fdiv_rn(f.y, den);
        const float norm = mag_f32(u, v);
        const float a = __fdiv_rn(atan2_f32(-v, -u), 3.14159274101257324f);
        const float fk = __fmul_rn(__fdiv_rn(__fadd_rn(a, 1.0f), 2.0f), 54.0f);
        const float k0f = floorf(fk);
        int k0 = (int)k0f;
        k0 = min(max(k0, 0), 54);
        const int k1 = k0 + 1 == 55 ? 0 : k0 + 1;
        const float fr = __fsub_rn(fk, k0f);
        uint8_t* o = rgb + 3 * i;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float col0 = __fdiv_rn((float)c_wheel[k0][c], 255.0f), col1 = __fdiv_rn((float)c_wheel[k1][c], 255.0f);
            float col = __fadd_rn(__fmul_rn(__fsub_rn(1.0f, fr), col0), __fmul_rn(fr, col1));
            col = __fsub_rn(1.0f, __fmul_rn(norm, __fsub_rn(1.0f, col)));
            const float q = floorf(__fmul_rn(255.0f, col));
            int b = q != q ? 0 : (int)fminf(fmaxf(q, 0.0f), 255.0f);
            if (ref_wrap) b = (b * 255) & 255;          // the reference multiplies the uint8 image by 255 in uint8 (:425)
            o[c] = (uint8_t)b;
        }
    }
}

// ------------------------------------------------------------------------------------------ LOD pyramid
// One CTA reduces a 16x16 block of level 0 to 8x8, 4x4, 2x2 and 1 cell(s) of levels 1..4.  A cell of level l+1 is the
// mean over the cells of its 2x2 block that exist in level l (sizes round up: the reference pads odd sizes bottom / right
// with weight-0 zeros), times 0.5; summed in numpy's order ((a00 + a01) + a10) + a11 so the result is bit-identical.
struct LodParams {
    const float2* src;
    float2* dst[4];
    int h[5], w[5];
    int levels;          // number of reduced levels to produce (1..4)
};

__global__ void __launch_bounds__(256) lod_pyramid_kernel(const LodParams p) {
    __shared__ float2 buf[2][16 * 16];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int gx = blockIdx.x * 16 + tx, gy = blockIdx.y * 16 + ty;
    buf[0][ty * 16 + tx] = (gx < p.w[0] && gy < p.h[0]) ? __ldg(p.src + (int64_t)gy * p.w[0] + gx) : make_float2(0.0f, 0.0f);
    __syncthreads();
    int side = 16;
    for (int l = 0; l < p.levels; ++l) {
        const int half = side >> 1;
        const float2* in = buf[l & 1];
        float2* out = buf[(l & 1) ^ 1];
        if ((int)threadIdx.x < half * half) {
            const int cx = threadIdx.x % half, cy = threadIdx.x / half;
            const int ox = blockIdx.x * half + cx, oy = blockIdx.y * half + cy;          // cell of level l+1
            const int bx = (blockIdx.x * side) + 2 * cx, by = (blockIdx.y * side) + 2 * cy;   // its block in level l
            float sx = 0.0f, sy = 0.0f, tot = 0.0f;
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const float wgt = (bx + i < p.w[l] && by + j < p.h[l]) ? 1.0f : 0.0f;
                    const float2 v = in[(2 * cy + j) * side + 2 * cx + i];
                    if (i == 0 && j == 0) {
                        sx = __fmul_rn(v.x, wgt); sy = __fmul_rn(v.y, wgt); tot = wgt;
                    } else {
                        sx = __fadd_rn(sx, __fmul_rn(v.x, wgt)); sy = __fadd_rn(sy, __fmul_rn(v.y, wgt)); tot = __fadd_rn(tot, wgt);
                    }
                }
            float2 r = make_float2(0.0f, 0.0f);
            if (tot > 0.0f) r = make_float2(__fmul_rn(__fdiv_rn(sx, tot), 0.5f), __fmul_rn(__fdiv_rn(sy, tot), 0.5f));
            out[cy * half + cx] = r;
            if (ox < p.w[l + 1] && oy < p.h[l + 1]) p.dst[l][(int64_t)oy * p.w[l + 1] + ox] = r;
        }
        __syncthreads();
        side = half;
    }
}

// ------------------------------------------------------------------------------------------ quality map
struct QualityParams {
    const uint8_t* f1;
    const uint8_t* f2;
    const float* flow;
    uint8_t* out;
    int h, w, fh, fw;
    float thr, sx, sy;         // sx = fw / w (source scale), also the vector rescale w / fw is applied
    float vx, vy;
};

__device__ __forceinline__ void resize_axis(int d, float scale, int n_in, int& i0, int& i1, float& l0, float& l1) {
    const float src = fmaxf(__fsub_rn(__fmul_rn(__fadd_rn((float)d, 0.5f), scale), 0.5f), 0.0f);
    i0 = min((int)src, n_in - 1);
    i1 = min(i0 + 1, n_in - 1);
    l1 = __fsub_rn(src, (float)i0);
    l0 = __fsub_rn(1.0f, l1);
}

__global__ void quality_kernel(const QualityParams p) {
    const int64_t npix = (int64_t)p.h * p.w;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < npix; i += (int64_t)gridDim.x * blockDim.x) {
        const int y = (int)(i / p.w), x = (int)(i - (int64_t)y * p.w);
        float fx, fy;
        if (p.fh == p.h && p.fw == p.w) {
            fx = p.flow[2 * i]; fy = p.flow[2 * i + 1];
        } else {                                   // F.interpolate(bilinear, align_corners=False) + vector rescale (:182-185)
            int x0, x1, y0, y1;
            float wx0, wx1, wy0, wy1;
            resize_axis(x, p.sx, p.fw, x0, x1, wx0, wx1);
            resize_axis(y, p.sy, p.fh, y0, y1, wy0, wy1);
            const float* r0 = p.flow + 2 * (int64_t)y0 * p.fw;
            const float* r1 = p.flow + 2 * (int64_t)y1 * p.fw;
            float o[2];
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const float top = __fadd_rn(__fmul_rn(r0[2 * x0 + c], wx0), __fmul_rn(r0[2 * x1 + c], wx1));
                const float bot = __fadd_rn(__fmul_rn(r1[2 * x0 + c], wx0), __fmul_rn(r1[2 * x1 + c], wx1));
                o[c] = __fadd_rn(__fmul_rn(top, wy0), __fmul_rn(bot, wy1));
            }
            fx = __fmul_rn(o[0], p.vx); fy = __fmul_rn(o[1], p.vy);
        }
        const float tx = __fsub_rn((float)x, fx), ty = __fsub_rn((float)y, fy);
        uint8_t* o = p.out + 3 * i;
        if (tx < 0.0f || tx >= (float)p.w || ty < 0.0f || ty >= (float)p.h) {        // NaN: not out of bounds, samples texel 0
            o[0] = 255; o[1] = 0; o[2] = 0;
            continue;
        }
        const int xi = tx != tx ? 0 : min(max((int)tx, 0), p.w - 1), yi = ty != ty ? 0 : min(max((int)ty, 0), p.h - 1);
        const uint8_t* a8 = p.f1 + 3 * i;
        const uint8_t* s8 = p.f2 + 3 * ((int64_t)yi * p.w + xi);
        float a[3], s[3], d[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            a[c] = __fdiv_rn((float)a8[c], 255.0f);
            s[c] = __fdiv_rn((float)s8[c], 255.0f);
            d[c] = __fsub_rn(a[c], s[c]);
        }
        auto sum3 = [](float x0, float x1, float x2) { return __fadd_rn(__fadd_rn(x0, x1), x2); };
        const float dist = __fsqrt_rn(sum3(__fmul_rn(d[0], d[0]), __fmul_rn(d[1], d[1]), __fmul_rn(d[2], d[2])));
        const float rgb_sim = __fsub_rn(1.0f, __fdiv_rn(dist, 1.732f));
        const float abs_sim = __fsub_rn(1.0f, __fdiv_rn(sum3(fabsf(d[0]), fabsf(d[1]), fabsf(d[2])), 3.0f));
        const float na = fmaxf(__fsqrt_rn(sum3(__fmul_rn(a[0], a[0]), __fmul_rn(a[1], a[1]), __fmul_rn(a[2], a[2]))), 1e-8f);
        const float ns = fmaxf(__fsqrt_rn(sum3(__fmul_rn(s[0], s[0]), __fmul_rn(s[1], s[1]), __fmul_rn(s[2], s[2]))), 1e-8f);
        const float cs = sum3(__fmul_rn(__fdiv_rn(a[0], na), __fdiv_rn(s[0], ns)), __fmul_rn(__fdiv_rn(a[1], na), __fdiv_rn(s[1], ns)),
                              __fmul_rn(__fdiv_rn(a[2], na), __fdiv_rn(s[2], ns)));
        const float cos_sim = __fdiv_rn(__fadd_rn(cs, 1.0f), 2.0f);
        const float sim = __fdiv_rn(__fadd_rn(__fadd_rn(rgb_sim, abs_sim), cos_sim), 3.0f);
        const float green = fminf(fmaxf(__fmul_rn(__fsub_rn(sim, 0.5f), 2.0f), 0.0f), 1.0f);
        const float red = fminf(fmaxf(__fsub_rn(1.0f, sim), 0.0f), 1.0f);
        const bool good = sim > p.thr;
        o[0] = good ? 0 : (uint8_t)(int)__fmul_rn(red, 255.0f);
        o[1] = good ? (uint8_t)(int)__fmul_rn(green, 255.0f) : 0;
        o[2] = 0;
    }
}

}  // namespace vfml

using namespace vfml;

#define VFML_EW_PROLOGUE(name)                                        \
    VFML_REQUIRE(height >= 0 && width >= 0, name ": bad shape");      \
    const int64_t n = (int64_t)height * width;                        \
    if (n == 0) return VFML_OK;

extern "C" int vfml_encode_mv_rgb8(const float* flow, uint8_t* rgb, int height, int width, float clamp, vfml_stream_t s) {
    VFML_EW_PROLOGUE("encode_mv_rgb8");
    VFML_REQUIRE(flow && rgb, "encode_mv_rgb8: null pointer");
    encode_rgb8_kernel<<<ew_blocks2(n), 256, 0, to_stream(s)>>>(reinterpret_cast<const float2*>(flow), rgb, n, clamp);
    VFML_LAUNCH_CHECK("encode_mv_rgb8");
    return VFML_OK;
}

extern "C" int vfml_decode_mv_rgb8(const uint8_t* rgb, float* flow, int height, int width, float clamp, vfml_stream_t s) {
    VFML_EW_PROLOGUE("decode_mv_rgb8");
    VFML_REQUIRE(flow && rgb, "decode_mv_rgb8: null pointer");
    decode_rgb8_kernel<<<ew_blocks2(n), 256, 0, to_stream(s)>>>(rgb, reinterpret_cast<float2*>(flow), n, clamp);
    VFML_LAUNCH_CHECK("decode_mv_rgb8");
    return VFML_OK;
}

extern "C" int vfml_encode_hsv(const float* flow, uint8_t* rgb, int height, int width, void* scratch4, vfml_stream_t s) {
    VFML_EW_PROLOGUE("encode_hsv");
    VFML_REQUIRE(flow && rgb && scratch4, "encode_hsv: null pointer");
    VFML_CUDA(cudaMemsetAsync(scratch4, 0, 4, to_stream(s)));
    max_norm_kernel<<<ew_blocks2(n), 256, 0, to_stream(s)>>>(reinterpret_cast<const float2*>(flow), n, 1, static_cast<unsigned int*>(scratch4));
    VFML_LAUNCH_CHECK("encode_hsv(max)");
    encode_hsv_kernel<<<ew_blocks2(n), 256, 0, to_stream(s)>>>(reinterpret_cast<const float2*>(flow), rgb, height, width,
                                                               static_cast<const unsigned int*>(scratch4));
    VFML_LAUNCH_CHECK("encode_hsv");
    return VFML_OK;
}

extern "C" int vfml_encode_torchvision(const float* flow, uint8_t* rgb, int height, int width, int ref_wrap, void* scratch4,
                                       vfml_stream_t s) {
    VFML_EW_PROLOGUE("encode_torchvision");
    VFML_REQUIRE(flow && rgb && scratch4, "encode_torchvision: null pointer");
    VFML_CUDA(cudaMemsetAsync(scratch4, 0, 4, to_stream(s)));
    max_norm_kernel<<<ew_blocks2(n), 256, 0, to_stream(s)>>>(reinterpret_cast<const float2*>(flow), n, 0, static_cast<unsigned int*>(scratch4));
    VFML_LAUNCH_CHECK("encode_torchvision(max)");
    encode_torchvision_kernel<<<ew_blocks2(n), 256, 0, to_stream(s)>>>(reinterpret_cast<const float2*>(flow), rgb, n,
                                                                       static_cast<const unsigned int*>(scratch4), ref_wrap);
    VFML_LAUNCH_CHECK("encode_torchvision");
    return VFML_OK;
}

extern "C" int vfml_lod_pyramid(const float* flow, int height, int width, float* const* levels, int num_levels, vfml_stream_t s) {
    VFML_REQUIRE(height >= 1 && width >= 1, "lod_pyramid: bad shape %dx%d", height, width);
    VFML_REQUIRE(num_levels >= 1 && num_levels <= 4, "lod_pyramid: num_levels %d not in 1..4 (chain calls for more)", num_levels);
    VFML_REQUIRE(flow && levels, "lod_pyramid: null pointer");
    LodParams p{};
    p.src = reinterpret_cast<const float2*>(flow);
    p.h[0] = height; p.w[0] = width;
    p.levels = num_levels;
    for (int l = 0; l < num_levels; ++l) {
        VFML_REQUIRE(levels[l], "lod_pyramid: null level pointer");
        p.dst[l] = reinterpret_cast<float2*>(levels[l]);
        p.h[l + 1] = (p.h[l] + 1) / 2; p.w[l + 1] = (p.w[l] + 1) / 2;
    }
    lod_pyramid_kernel<<<dim3(cdiv(width, 16), cdiv(height, 16)), 256, 0, to_stream(s)>>>(p);
    VFML_LAUNCH_CHECK("lod_pyramid");
    return VFML_OK;
}

extern "C" int vfml_quality_frame(const uint8_t* frame1, const uint8_t* frame2, const float* flow, int flow_h, int flow_w,
                                  uint8_t* out, int height, int width, float threshold, vfml_stream_t s) {
    VFML_EW_PROLOGUE("quality_frame");
    VFML_REQUIRE(frame1 && frame2 && flow && out, "quality_frame: null pointer");
    VFML_REQUIRE(flow_h >= 1 && flow_w >= 1, "quality_frame: bad flow shape");
    QualityParams p{frame1, frame2, flow, out, height, width, flow_h, flow_w, threshold,
                    (float)((double)flow_w / width), (float)((double)flow_h / height),
                    (float)((double)width / flow_w), (float)((double)height / flow_h)};
    quality_kernel<<<ew_blocks2(n), 256, 0, to_stream(s)>>>(p);
    VFML_LAUNCH_CHECK("quality_frame");
    return VFML_OK;
}

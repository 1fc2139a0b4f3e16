// Convex 8x flow upsampling with the 9-tap softmax fused, cropped write (unpad + tile paste).
//
// Replaces MOFNet.upsample_flow (MOFNetStack/network.py:97-108):
//   out[c, 8y+a, 8x+b] = sum_k softmax_k(mask[k*64 + a*8 + b, y, x]) * 8 * flow[c, y + k/3 - 1, x + k%3 - 1]
// HBM-bound on the mask read (576 values per low-res pixel).  One warp per low-res pixel:
// each lane owns two of the 64 sub-pixels, so mask reads (rows layout) are 128-byte
// coalesced and every output row segment of 8 floats is written by 8 adjacent lanes.
#include "common.cuh"

namespace vfml {

struct UpParams {
    const float* flow;
    const void* mask;
    int mask_layout;
    int64_t mask_pitch;
    int mask_col;
    int n, h, w;
    float* out;
    int64_t sn, sc, sy;
    int crop_top, crop_left, out_h, out_w;
    int sub_grid;   // flow holds coords1: subtract the pixel grid (flow = coords1 - coords0, network.py:206)
};

template <typename TM>
__global__ void __launch_bounds__(256)
convex_upsample8_kernel(const UpParams p) {
    const int lane = threadIdx.x & 31;
    const int P = p.h * p.w;
    const int64_t total = (int64_t)p.n * P;
    const int64_t warp0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t q = warp0; q < total; q += nwarps) {
        const int i = (int)(q / P);
        const int pix = (int)(q - (int64_t)i * P);
        const int y = pix / p.w, x = pix - y * p.w;
        // 3x3 neighbourhood of 8*flow, zero padded (F.unfold(8*flow, 3, padding=1))
        float nfx[9], nfy[9];
        const float* fxp = p.flow + (int64_t)i * 2 * P;
        const float* fyp = fxp + P;
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            const int yy = y + k / 3 - 1, xx = x + k % 3 - 1;
            const bool in = yy >= 0 && yy < p.h && xx >= 0 && xx < p.w;
            const float gx = p.sub_grid ? (float)xx : 0.0f, gy = p.sub_grid ? (float)yy : 0.0f;
            nfx[k] = in ? 8.0f * (__ldg(fxp + yy * p.w + xx) - gx) : 0.0f;
            nfy[k] = in ? 8.0f * (__ldg(fyp + yy * p.w + xx) - gy) : 0.0f;
        }
        const TM* m = reinterpret_cast<const TM*>(p.mask);
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const int s = lane + 32 * half;   // sub-pixel a*8+b
            float mv[9];
            float mx = -INFINITY;
#pragma unroll
            for (int k = 0; k < 9; ++k) {
                const int ch = k * 64 + s;
                float v;
                if (p.mask_layout == VFML_NCHW) v = to_f32<TM>(m[((int64_t)i * 576 + ch) * P + pix]);
                else v = to_f32<TM>(m[((int64_t)i * P + pix) * p.mask_pitch + p.mask_col + ch]);
                mv[k] = v;
                mx = fmaxf(mx, v);
            }
            float den = 0.0f;
#pragma unroll
            for (int k = 0; k < 9; ++k) {
                mv[k] = expf(mv[k] - mx);
                den += mv[k];
            }
            float ox = 0.0f, oy = 0.0f;
#pragma unroll
            for (int k = 0; k < 9; ++k) {
                const float wgt = mv[k] / den;
                ox += wgt * nfx[k];
                oy += wgt * nfy[k];
            }
            const int Y = 8 * y + (s >> 3) - p.crop_top;
            const int X = 8 * x + (s & 7) - p.crop_left;
            if (Y >= 0 && Y < p.out_h && X >= 0 && X < p.out_w) {
                float* o = p.out + (int64_t)i * p.sn + (int64_t)Y * p.sy + X;
                o[0] = ox;
                o[p.sc] = oy;
            }
        }
    }
}

// NCHW fp32 masks (the drop-in upsample_flow signature): lanes across x so that each of the
// 576 channel planes is read coalesced; each thread produces the full 8x8 block of a pixel.
__global__ void __launch_bounds__(128)
convex_upsample8_nchw_kernel(const UpParams p) {
    const int P = p.h * p.w;
    const int64_t total = (int64_t)p.n * P;
    const int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (q >= total) return;
    const int i = (int)(q / P);
    const int pix = (int)(q - (int64_t)i * P);
    const int y = pix / p.w, x = pix - y * p.w;
    float nfx[9], nfy[9];
    const float* fxp = p.flow + (int64_t)i * 2 * P;
    const float* fyp = fxp + P;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const int yy = y + k / 3 - 1, xx = x + k % 3 - 1;
        const bool in = yy >= 0 && yy < p.h && xx >= 0 && xx < p.w;
        const float gx = p.sub_grid ? (float)xx : 0.0f, gy = p.sub_grid ? (float)yy : 0.0f;
        nfx[k] = in ? 8.0f * (__ldg(fxp + yy * p.w + xx) - gx) : 0.0f;
        nfy[k] = in ? 8.0f * (__ldg(fyp + yy * p.w + xx) - gy) : 0.0f;
    }
    const float* m = reinterpret_cast<const float*>(p.mask) + (int64_t)i * 576 * P + pix;
    for (int a = 0; a < 8; ++a) {
        float rx[8], ry[8];
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            const int s = a * 8 + b;
            float mv[9];
            float mx = -INFINITY;
#pragma unroll
            for (int k = 0; k < 9; ++k) {
                mv[k] = __ldg(m + (int64_t)(k * 64 + s) * P);
                mx = fmaxf(mx, mv[k]);
            }
            float den = 0.0f;
#pragma unroll
            for (int k = 0; k < 9; ++k) {
                mv[k] = expf(mv[k] - mx);
                den += mv[k];
            }
            float ox = 0.0f, oy = 0.0f;
#pragma unroll
            for (int k = 0; k < 9; ++k) {
                const float wgt = mv[k] / den;
                ox += wgt * nfx[k];
                oy += wgt * nfy[k];
            }
            rx[b] = ox;
            ry[b] = oy;
        }
        const int Y = 8 * y + a - p.crop_top;
        if (Y < 0 || Y >= p.out_h) continue;
        float* o = p.out + (int64_t)i * p.sn + (int64_t)Y * p.sy;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            const int X = 8 * x + b - p.crop_left;
            if (X >= 0 && X < p.out_w) {
                o[X] = rx[b];
                o[p.sc + X] = ry[b];
            }
        }
    }
}

}  // namespace vfml

using namespace vfml;

extern "C" int vfml_convex_upsample8(const float* flow, const void* mask, int mask_layout, int mask_pitch,
                                     int mask_col, int mask_dtype, int n, int h, int w, float* out,
                                     int64_t out_stride_n, int64_t out_stride_c, int64_t out_stride_y,
                                     int crop_top, int crop_left, int out_h, int out_w, int flow_is_coords, vfml_stream_t s) {
    VFML_REQUIRE(n >= 0 && h > 0 && w > 0, "convex_upsample8: bad shape");
    VFML_REQUIRE(mask_layout == VFML_NCHW || mask_layout == VFML_ROWS, "convex_upsample8: bad layout");
    VFML_REQUIRE(valid_dtype(mask_dtype), "convex_upsample8: bad dtype");
    VFML_REQUIRE(crop_top >= 0 && crop_left >= 0 && out_h >= 0 && out_w >= 0 && crop_top + out_h <= 8 * h &&
                     crop_left + out_w <= 8 * w, "convex_upsample8: crop outside the %dx%d result", 8 * h, 8 * w);
    if (mask_layout == VFML_ROWS)
        VFML_REQUIRE(mask_pitch >= mask_col + 576 && mask_col >= 0, "convex_upsample8: mask pitch %d too small", mask_pitch);
    if (mask_layout == VFML_NCHW) VFML_REQUIRE(mask_dtype == VFML_F32, "convex_upsample8: NCHW mask must be fp32");
    if (n == 0 || out_h == 0 || out_w == 0) return VFML_OK;
    VFML_REQUIRE(flow && mask && out, "convex_upsample8: null pointer");
    UpParams p{flow, mask, mask_layout, mask_pitch, mask_col, n, h, w, out, out_stride_n, out_stride_c,
               out_stride_y, crop_top, crop_left, out_h, out_w, flow_is_coords};
    const int64_t total = (int64_t)n * h * w;
    cudaStream_t st = to_stream(s);
    if (mask_layout == VFML_NCHW) {
        convex_upsample8_nchw_kernel<<<cdiv(total, 128), 128, 0, st>>>(p);
    } else {
        const int blocks = (int)std::min<int64_t>(cdiv(total, 8), device_sms() * 16);
        if (mask_dtype == VFML_F32) convex_upsample8_kernel<float><<<blocks, 256, 0, st>>>(p);
        else convex_upsample8_kernel<__half><<<blocks, 256, 0, st>>>(p);
    }
    VFML_LAUNCH_CHECK("convex_upsample8");
    return VFML_OK;
}

// Multi-level (2r+1)^2 bilinear correlation lookup.
//
// Replaces CorrBlock.__call__ (VideoFlow/core/Networks/MOFNetStack/corr.py:76-97) and
// bilinear_sampler (core/utils/utils.py:62-76).  HBM-bound gather: per query pixel and
// level the footprint of the 81 unit-spaced taps is a 10x10 texel window of that pixel's
// own [h_l, w_l] slab, i.e. ten 40-byte row segments.
//
// Mapping: a CTA owns TP consecutive pixels of one frame.  Each warp gathers the four
// 10x10 windows of one pixel with all loads in flight at once (13 independent 4-byte loads
// per lane), stages them in shared memory, interpolates 81 taps per level from shared
// memory and parks the 324 results in a [TP][325] shared tile.  The tile is then written
// with fully coalesced stores in either layout (NCHW: 32 consecutive pixels per channel;
// rows: 324 consecutive channels per pixel).
#include "common.cuh"

namespace vfml {

constexpr int LK_TP = 32;       // pixels per CTA
constexpr int LK_WARPS = 8;
constexpr int LK_MAXL = 4;

struct LookupParams {
    const void* lvl[LK_MAXL];
    int lh[LK_MAXL], lw[LK_MAXL];
    int num_levels;
    int h, w, n_frames;
    const float* coords;
    void* out;
    int layout;
    int64_t out_pitch;
    int out_col;
};

template <typename TV, typename TO, int R>
__global__ void __launch_bounds__(LK_WARPS * 32)
corr_lookup_kernel(const LookupParams p) {
    constexpr int WIN = 2 * R + 1;        // 9
    constexpr int FP = WIN + 1;           // 10 texels per side
    constexpr int NTEX = FP * FP;         // 100
    constexpr int NTEX_PAD = (NTEX + 3) / 4 * 4 + 4;
    constexpr int NTAP = WIN * WIN;       // 81
    constexpr int LOADS = (NTEX + 31) / 32;
    const int CH = p.num_levels * NTAP;
    const int CHP = CH + 1;               // odd pitch: conflict-free transposed reads

    extern __shared__ float smem[];
    float* s_out = smem;                                   // [LK_TP][CHP]
    float* s_tex = smem + LK_TP * (LK_MAXL * NTAP + 1);    // [LK_WARPS][LK_MAXL][NTEX_PAD]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int P = p.h * p.w;
    const int tiles_per_frame = (P + LK_TP - 1) / LK_TP;
    const int f = blockIdx.x / tiles_per_frame;
    const int p0 = (blockIdx.x - f * tiles_per_frame) * LK_TP;
    const float* cx_plane = p.coords + (int64_t)f * 2 * P;
    const float* cy_plane = cx_plane + P;
    float* tex = s_tex + warp * (LK_MAXL * NTEX_PAD);

    for (int pi = warp; pi < LK_TP; pi += LK_WARPS) {
        const int pix = p0 + pi;
        if (pix >= P) break;
        const float cx = __ldg(cx_plane + pix);
        const float cy = __ldg(cy_plane + pix);
        float fx[LK_MAXL], fy[LK_MAXL];
        float v[LK_MAXL][LOADS];
        float scale = 1.0f;
#pragma unroll
        for (int l = 0; l < LK_MAXL; ++l) {
            if (l < p.num_levels) {
                // coords / 2**l (corr.py:88); clamp keeps the int conversion defined
                float sx = fminf(fmaxf(cx * scale, -1.0e6f), 1.0e6f);
                float sy = fminf(fmaxf(cy * scale, -1.0e6f), 1.0e6f);
                if (!(sx == sx)) sx = -1.0e6f;   // NaN coordinate: every tap out of range -> zeros
                if (!(sy == sy)) sy = -1.0e6f;
                const float flx = floorf(sx), fly = floorf(sy);
                fx[l] = sx - flx;
                fy[l] = sy - fly;
                const int x0 = (int)flx - R, y0 = (int)fly - R;
                const int lw = p.lw[l], lh = p.lh[l];
                const TV* slab = reinterpret_cast<const TV*>(p.lvl[l]) +
                                 ((int64_t)f * P + pix) * ((int64_t)lh * lw);
#pragma unroll
                for (int k = 0; k < LOADS; ++k) {
                    const int idx = lane + 32 * k;
                    const int r = idx / FP, c = idx - r * FP;
                    const int gx = x0 + c, gy = y0 + r;
                    float t = 0.0f;
                    if (idx < NTEX && gx >= 0 && gx < lw && gy >= 0 && gy < lh)
                        t = to_f32<TV>(__ldg(slab + (int64_t)gy * lw + gx));
                    v[l][k] = t;
                }
            }
            scale *= 0.5f;
        }
#pragma unroll
        for (int l = 0; l < LK_MAXL; ++l) {
            if (l < p.num_levels) {
#pragma unroll
                for (int k = 0; k < LOADS; ++k) {
                    const int idx = lane + 32 * k;
                    if (idx < NTEX) tex[l * NTEX_PAD + idx] = v[l][k];
                }
            }
        }
        __syncwarp();
        float* orow = s_out + pi * CHP;
#pragma unroll
        for (int l = 0; l < LK_MAXL; ++l) {
            if (l < p.num_levels) {
                const float wx1 = fx[l], wx0 = 1.0f - fx[l];
                const float wy1 = fy[l], wy0 = 1.0f - fy[l];
                const float* T = tex + l * NTEX_PAD;
#pragma unroll
                for (int k = lane; k < NTAP; k += 32) {
                    const int i = k / WIN, j = k - i * WIN;   // i: x offset (slow), j: y offset
                    const float* t0 = T + j * FP + i;
                    const float v00 = t0[0], v01 = t0[1], v10 = t0[FP], v11 = t0[FP + 1];
                    orow[l * NTAP + k] = v00 * (wx0 * wy0) + v01 * (wx1 * wy0) +
                                         v10 * (wx0 * wy1) + v11 * (wx1 * wy1);
                }
            }
        }
        __syncwarp();
    }
    __syncthreads();

    const int npix = min(LK_TP, P - p0);
    TO* out = reinterpret_cast<TO*>(p.out);
    if (p.layout == VFML_NCHW) {
        // out[f][ch][pix]: lanes run over pixels
        TO* base = out + (int64_t)f * CH * P + p0;
        for (int ch = warp; ch < CH; ch += LK_WARPS) {
            if (lane < npix) base[(int64_t)ch * P + lane] = from_f32<TO>(s_out[lane * CHP + ch]);
        }
    } else {
        // rows: out[(f*P + pix) * pitch + col + ch]: lanes run over channels
        for (int pi = warp; pi < npix; pi += LK_WARPS) {
            TO* row = out + ((int64_t)f * P + p0 + pi) * p.out_pitch + p.out_col;
            const float* srow = s_out + pi * CHP;
            for (int ch = lane; ch < CH; ch += 32) row[ch] = from_f32<TO>(srow[ch]);
        }
    }
}

// Any radius / level count: one thread per output element (also the in-library cross-check).
template <typename TV, typename TO>
__global__ void corr_lookup_generic_kernel(const LookupParams p, int radius) {
    const int win = 2 * radius + 1, ntap = win * win;
    const int CH = p.num_levels * ntap;
    const int P = p.h * p.w;
    const int64_t total = (int64_t)p.n_frames * P * CH;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total;
         t += (int64_t)gridDim.x * blockDim.x) {
        int ch, pix, f;
        if (p.layout == VFML_NCHW) {
            pix = (int)(t % P);
            ch = (int)((t / P) % CH);
            f = (int)(t / ((int64_t)P * CH));
        } else {
            ch = (int)(t % CH);
            pix = (int)((t / CH) % P);
            f = (int)(t / ((int64_t)P * CH));
        }
        const int l = ch / ntap, k = ch - l * ntap;
        const int i = k / win, j = k - i * win;
        const float sc = 1.0f / (float)(1 << l);
        float sx = p.coords[((int64_t)f * 2) * P + pix] * sc + (float)(i - radius);
        float sy = p.coords[((int64_t)f * 2 + 1) * P + pix] * sc + (float)(j - radius);
        sx = fminf(fmaxf(sx, -1.0e6f), 1.0e6f);
        sy = fminf(fmaxf(sy, -1.0e6f), 1.0e6f);
        if (!(sx == sx)) sx = -1.0e6f;
        if (!(sy == sy)) sy = -1.0e6f;
        const float flx = floorf(sx), fly = floorf(sy);
        const float wx1 = sx - flx, wy1 = sy - fly;
        const int x0 = (int)flx, y0 = (int)fly;
        const int lw = p.lw[l], lh = p.lh[l];
        const TV* slab = reinterpret_cast<const TV*>(p.lvl[l]) + ((int64_t)f * P + pix) * ((int64_t)lh * lw);
        auto tx = [&](int gx, int gy) -> float {
            return (gx >= 0 && gx < lw && gy >= 0 && gy < lh) ? to_f32<TV>(slab[(int64_t)gy * lw + gx]) : 0.0f;
        };
        const float r = tx(x0, y0) * ((1.0f - wx1) * (1.0f - wy1)) + tx(x0 + 1, y0) * (wx1 * (1.0f - wy1)) +
                        tx(x0, y0 + 1) * ((1.0f - wx1) * wy1) + tx(x0 + 1, y0 + 1) * (wx1 * wy1);
        TO* out = reinterpret_cast<TO*>(p.out);
        if (p.layout == VFML_NCHW) out[t] = from_f32<TO>(r);
        else out[((int64_t)f * P + pix) * p.out_pitch + p.out_col + ch] = from_f32<TO>(r);
    }
}

template <typename TV, typename TO>
static int launch_lookup(const LookupParams& p, int radius, cudaStream_t st) {
    const int P = p.h * p.w;
    if (radius == 4) {
        auto kern = corr_lookup_kernel<TV, TO, 4>;
        const size_t smem = (LK_TP * (LK_MAXL * 81 + 1) + LK_WARPS * LK_MAXL * 108) * sizeof(float);
        static bool attr_done = false;
        if (!attr_done) {
            VFML_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr_done = true;
        }
        const int tiles = (P + LK_TP - 1) / LK_TP;
        kern<<<p.n_frames * tiles, LK_WARPS * 32, smem, st>>>(p);
        VFML_LAUNCH_CHECK("corr_lookup");
    } else {
        const int win = 2 * radius + 1;
        const int64_t total = (int64_t)p.n_frames * P * p.num_levels * win * win;
        const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 16);
        corr_lookup_generic_kernel<TV, TO><<<blocks, 256, 0, st>>>(p, radius);
        VFML_LAUNCH_CHECK("corr_lookup_generic");
    }
    return VFML_OK;
}

}  // namespace vfml

using namespace vfml;

extern "C" int vfml_corr_lookup(const void* const* levels, int num_levels, int vol_dtype, int n_frames,
                                int h, int w, int radius, const float* coords, void* out, int out_layout,
                                int out_pitch, int out_col, int out_dtype, vfml_stream_t s) {
    VFML_REQUIRE(num_levels >= 1 && num_levels <= LK_MAXL, "corr_lookup: num_levels %d not in 1..4", num_levels);
    VFML_REQUIRE(radius >= 0 && radius <= 8, "corr_lookup: radius %d not in 0..8", radius);
    VFML_REQUIRE(n_frames >= 0 && h > 0 && w > 0, "corr_lookup: bad shape");
    VFML_REQUIRE(valid_dtype(vol_dtype) && valid_dtype(out_dtype), "corr_lookup: bad dtype");
    VFML_REQUIRE(out_layout == VFML_NCHW || out_layout == VFML_ROWS, "corr_lookup: bad layout");
    const int CH = num_levels * (2 * radius + 1) * (2 * radius + 1);
    if (out_layout == VFML_ROWS)
        VFML_REQUIRE(out_pitch >= out_col + CH && out_col >= 0, "corr_lookup: pitch %d < col %d + %d", out_pitch, out_col, CH);
    if (n_frames == 0) return VFML_OK;
    VFML_REQUIRE(levels && coords && out, "corr_lookup: null pointer");
    LookupParams p{};
    int lh = h, lw = w;
    for (int l = 0; l < num_levels; ++l) {
        VFML_REQUIRE(levels[l] != nullptr, "corr_lookup: level %d is null", l);
        VFML_REQUIRE(lh > 0 && lw > 0, "corr_lookup: level %d is empty (h=%d, w=%d)", l, h, w);
        p.lvl[l] = levels[l];
        p.lh[l] = lh;
        p.lw[l] = lw;
        lh /= 2;
        lw /= 2;
    }
    p.num_levels = num_levels;
    p.h = h; p.w = w; p.n_frames = n_frames;
    p.coords = coords;
    p.out = out;
    p.layout = out_layout;
    p.out_pitch = out_pitch;
    p.out_col = out_col;
    cudaStream_t st = to_stream(s);
    if (vol_dtype == VFML_F32 && out_dtype == VFML_F32) return launch_lookup<float, float>(p, radius, st);
    if (vol_dtype == VFML_F32 && out_dtype == VFML_F16) return launch_lookup<float, __half>(p, radius, st);
    if (vol_dtype == VFML_F16 && out_dtype == VFML_F16) return launch_lookup<__half, __half>(p, radius, st);
    return launch_lookup<__half, float>(p, radius, st);
}

// Multi-level (2r+1)^2 bilinear correlation lookup.
//
// Replaces CorrBlock.__call__ (VideoFlow/core/Networks/MOFNetStack/corr.py:76-97) and
// bilinear_sampler (core/utils/utils.py:62-76).  HBM-bound gather: per query pixel and level the
// footprint of the 81 unit-spaced taps is a 10x10 texel window of that pixel's own [h_l, w_l]
// slab, i.e. ten 40-byte (fp32) / 20-byte (fp16) row segments.
//
// rows layout (the engine's path): one warp per query pixel, software pipelined.  A warp gathers
// the four 10x10 windows of pixel i+1 into registers (13 independent loads per lane in flight)
// while it interpolates pixel i from a per-warp shared-memory staging buffer and writes the 324
// results with coalesced stores (324 consecutive channels per pixel).  Forward and backward
// lookups of one update iteration go out as a single launch ("sets").
//
// NCHW layout (the drop-in CorrBlock.__call__ signature): a CTA owns 32 consecutive pixels and
// transposes through a [32][325] shared tile so that every channel plane is written 128 bytes at
// a time.
#include "common.cuh"

namespace vfml {

constexpr int LK_TP = 32;       // pixels per CTA (NCHW kernel)
constexpr int LK_WARPS = 8;
constexpr int LK_MAXL = 4;
constexpr int LK_R = 4, LK_WIN = 9, LK_FP = 10, LK_NTEX = 100, LK_NTEX_PAD = 104, LK_NTAP = 81, LK_LOADS = 4;

struct LookupSet {
    const void* lvl[LK_MAXL];
    const float* coords;
    void* out;
};

struct LookupParams {
    LookupSet set[2];
    int n_sets;
    int lh[LK_MAXL], lw[LK_MAXL];
    int num_levels;
    int h, w, n_frames;
    int layout;
    int64_t out_pitch;
    int out_col;
};

// gather of one pixel: 4 levels x 100 texels spread over the warp, plus the fractional offsets
template <typename TV>
struct Gather {
    float v[LK_MAXL][LK_LOADS];
    float fx[LK_MAXL], fy[LK_MAXL];

    __device__ __forceinline__ void load(const LookupParams& p, const LookupSet& s, int64_t frame_pix, int P, int lane) {
        const int64_t f = frame_pix / P;
        const int pix = (int)(frame_pix - f * P);
        const float cx = __ldg(s.coords + f * 2 * P + pix);
        const float cy = __ldg(s.coords + (f * 2 + 1) * P + pix);
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
                const int x0 = (int)flx - LK_R, y0 = (int)fly - LK_R;
                const int lw = p.lw[l], lh = p.lh[l];
                const TV* slab = reinterpret_cast<const TV*>(s.lvl[l]) + frame_pix * ((int64_t)lh * lw);
#pragma unroll
                for (int k = 0; k < LK_LOADS; ++k) {
                    const int idx = lane + 32 * k;
                    const int r = idx / LK_FP, c = idx - r * LK_FP;
                    const int gx = x0 + c, gy = y0 + r;
                    float t = 0.0f;
                    if (idx < LK_NTEX && gx >= 0 && gx < lw && gy >= 0 && gy < lh)
                        t = to_f32<TV>(__ldg(slab + (int64_t)gy * lw + gx));
                    v[l][k] = t;
                }
            }
            scale *= 0.5f;
        }
    }

    __device__ __forceinline__ void stage(float* tex, int num_levels, int lane) const {
#pragma unroll
        for (int l = 0; l < LK_MAXL; ++l) {
            if (l < num_levels) {
#pragma unroll
                for (int k = 0; k < LK_LOADS; ++k) {
                    const int idx = lane + 32 * k;
                    if (idx < LK_NTEX) tex[l * LK_NTEX_PAD + idx] = v[l][k];
                }
            }
        }
    }
};

// tap k = i*9 + j samples (x + i - 4, y + j - 4): i (x offset) is the slow index
__device__ __forceinline__ float interp_tap(const float* T, int k, float wx0, float wx1, float wy0, float wy1) {
    const int i = k / LK_WIN, j = k - i * LK_WIN;
    const float* t0 = T + j * LK_FP + i;
    return t0[0] * (wx0 * wy0) + t0[1] * (wx1 * wy0) + t0[LK_FP] * (wx0 * wy1) + t0[LK_FP + 1] * (wx1 * wy1);
}

template <typename TV, typename TO>
__global__ void __launch_bounds__(128, 8)
corr_lookup_rows_kernel(const LookupParams p) {
    __shared__ float s_tex[4][LK_MAXL * LK_NTEX_PAD];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int P = p.h * p.w;
    const int64_t per_set = (int64_t)p.n_frames * P;
    const int64_t total = per_set * p.n_sets;
    const int64_t stride = (int64_t)gridDim.x * 4;
    float* tex = s_tex[warp];
    int64_t q = (int64_t)blockIdx.x * 4 + warp;
    if (q >= total) return;
    Gather<TV> g;
    {
        const int si = q >= per_set;
        g.load(p, p.set[si], q - si * per_set, P, lane);
    }
    while (q < total) {
        const int si = q >= per_set;
        const int64_t fp = q - si * per_set;
        g.stage(tex, p.num_levels, lane);
        float fx[LK_MAXL], fy[LK_MAXL];
#pragma unroll
        for (int l = 0; l < LK_MAXL; ++l) { fx[l] = g.fx[l]; fy[l] = g.fy[l]; }
        __syncwarp();
        const int64_t qn = q + stride;
        if (qn < total) {                     // next pixel's loads fly while this one is interpolated
            const int sn = qn >= per_set;
            g.load(p, p.set[sn], qn - sn * per_set, P, lane);
        }
        TO* row = reinterpret_cast<TO*>(p.set[si].out) + fp * p.out_pitch + p.out_col;
#pragma unroll
        for (int l = 0; l < LK_MAXL; ++l) {
            if (l < p.num_levels) {
                const float wx1 = fx[l], wx0 = 1.0f - fx[l], wy1 = fy[l], wy0 = 1.0f - fy[l];
                const float* T = tex + l * LK_NTEX_PAD;
#pragma unroll
                for (int k = lane; k < LK_NTAP; k += 32)
                    row[l * LK_NTAP + k] = from_f32<TO>(interp_tap(T, k, wx0, wx1, wy0, wy1));
            }
        }
        __syncwarp();
        q = qn;
    }
}

template <typename TV, typename TO>
__global__ void __launch_bounds__(LK_WARPS * 32)
corr_lookup_nchw_kernel(const LookupParams p) {
    const int CH = p.num_levels * LK_NTAP;
    const int CHP = CH + 1;               // odd pitch: conflict-free transposed reads
    extern __shared__ float smem[];
    float* s_out = smem;                                        // [LK_TP][CHP]
    float* s_tex = smem + LK_TP * (LK_MAXL * LK_NTAP + 1);      // [LK_WARPS][LK_MAXL][NTEX_PAD]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int P = p.h * p.w;
    const int tiles_per_frame = (P + LK_TP - 1) / LK_TP;
    const int f = blockIdx.x / tiles_per_frame;
    const int p0 = (blockIdx.x - f * tiles_per_frame) * LK_TP;
    float* tex = s_tex + warp * (LK_MAXL * LK_NTEX_PAD);
    for (int pi = warp; pi < LK_TP; pi += LK_WARPS) {
        const int pix = p0 + pi;
        if (pix >= P) break;
        Gather<TV> g;
        g.load(p, p.set[0], (int64_t)f * P + pix, P, lane);
        g.stage(tex, p.num_levels, lane);
        __syncwarp();
        float* orow = s_out + pi * CHP;
#pragma unroll
        for (int l = 0; l < LK_MAXL; ++l) {
            if (l < p.num_levels) {
                const float wx1 = g.fx[l], wx0 = 1.0f - g.fx[l], wy1 = g.fy[l], wy0 = 1.0f - g.fy[l];
#pragma unroll
                for (int k = lane; k < LK_NTAP; k += 32)
                    orow[l * LK_NTAP + k] = interp_tap(tex + l * LK_NTEX_PAD, k, wx0, wx1, wy0, wy1);
            }
        }
        __syncwarp();
    }
    __syncthreads();
    const int npix = min(LK_TP, P - p0);
    TO* base = reinterpret_cast<TO*>(p.set[0].out) + (int64_t)f * CH * P + p0;
    for (int ch = warp; ch < CH; ch += LK_WARPS)
        if (lane < npix) base[(int64_t)ch * P + lane] = from_f32<TO>(s_out[lane * CHP + ch]);
}

// Any radius / level count: one thread per output element (also the in-library cross-check).
template <typename TV, typename TO>
__global__ void corr_lookup_generic_kernel(const LookupParams p, int radius) {
    const int win = 2 * radius + 1, ntap = win * win;
    const int CH = p.num_levels * ntap;
    const int P = p.h * p.w;
    const int64_t total = (int64_t)p.n_frames * P * CH;
    const LookupSet& s = p.set[0];
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int ch, pix, f;
        if (p.layout == VFML_NCHW) {
            pix = (int)(t % P);
            ch = (int)((t / P) % CH);
        } else {
            ch = (int)(t % CH);
            pix = (int)((t / CH) % P);
        }
        f = (int)(t / ((int64_t)P * CH));
        const int l = ch / ntap, k = ch - l * ntap;
        const int i = k / win, j = k - i * win;
        const float sc = 1.0f / (float)(1 << l);
        float sx = s.coords[((int64_t)f * 2) * P + pix] * sc + (float)(i - radius);
        float sy = s.coords[((int64_t)f * 2 + 1) * P + pix] * sc + (float)(j - radius);
        sx = fminf(fmaxf(sx, -1.0e6f), 1.0e6f);
        sy = fminf(fmaxf(sy, -1.0e6f), 1.0e6f);
        if (!(sx == sx)) sx = -1.0e6f;
        if (!(sy == sy)) sy = -1.0e6f;
        const float flx = floorf(sx), fly = floorf(sy);
        const float wx1 = sx - flx, wy1 = sy - fly;
        const int x0 = (int)flx, y0 = (int)fly;
        const int lw = p.lw[l], lh = p.lh[l];
        const TV* slab = reinterpret_cast<const TV*>(s.lvl[l]) + ((int64_t)f * P + pix) * ((int64_t)lh * lw);
        auto tx = [&](int gx, int gy) -> float {
            return (gx >= 0 && gx < lw && gy >= 0 && gy < lh) ? to_f32<TV>(slab[(int64_t)gy * lw + gx]) : 0.0f;
        };
        const float r = tx(x0, y0) * ((1.0f - wx1) * (1.0f - wy1)) + tx(x0 + 1, y0) * (wx1 * (1.0f - wy1)) +
                        tx(x0, y0 + 1) * ((1.0f - wx1) * wy1) + tx(x0 + 1, y0 + 1) * (wx1 * wy1);
        TO* out = reinterpret_cast<TO*>(s.out);
        if (p.layout == VFML_NCHW) out[t] = from_f32<TO>(r);
        else out[((int64_t)f * P + pix) * p.out_pitch + p.out_col + ch] = from_f32<TO>(r);
    }
}

template <typename TV, typename TO>
static int launch_lookup(const LookupParams& p, int radius, cudaStream_t st) {
    const int P = p.h * p.w;
    if (radius == LK_R && p.layout == VFML_ROWS) {
        const int64_t total = (int64_t)p.n_frames * P * p.n_sets;
        const int blocks = (int)std::min<int64_t>((total + 3) / 4, device_sms() * 8);
        corr_lookup_rows_kernel<TV, TO><<<blocks, 128, 0, st>>>(p);
        VFML_LAUNCH_CHECK("corr_lookup_rows");
    } else if (radius == LK_R) {
        auto kern = corr_lookup_nchw_kernel<TV, TO>;
        const size_t smem = (LK_TP * (LK_MAXL * LK_NTAP + 1) + LK_WARPS * LK_MAXL * LK_NTEX_PAD) * sizeof(float);
        static PerDevice<bool> attr_done;
        if (!attr_done.get()) {
            VFML_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr_done.get() = true;
        }
        kern<<<p.n_frames * ((P + LK_TP - 1) / LK_TP), LK_WARPS * 32, smem, st>>>(p);
        VFML_LAUNCH_CHECK("corr_lookup_nchw");
    } else {
        const int win = 2 * radius + 1;
        const int64_t total = (int64_t)p.n_frames * P * p.num_levels * win * win;
        const int blocks = (int)std::min<int64_t>((total + 255) / 256, device_sms() * 16);
        corr_lookup_generic_kernel<TV, TO><<<blocks, 256, 0, st>>>(p, radius);
        VFML_LAUNCH_CHECK("corr_lookup_generic");
    }
    return VFML_OK;
}

static int run_lookup(const void* const* lv0, const void* const* lv1, int num_levels, int vol_dtype, int n_frames, int h,
                      int w, int radius, const float* c0, const float* c1, void* o0, void* o1, int out_layout,
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
    VFML_REQUIRE(lv0 && c0 && o0, "corr_lookup: null pointer");
    LookupParams p{};
    p.n_sets = lv1 ? 2 : 1;
    if (lv1) {
        VFML_REQUIRE(c1 && o1, "corr_lookup_pair: null pointer");
        VFML_REQUIRE(out_layout == VFML_ROWS && radius == LK_R, "corr_lookup_pair: rows layout and radius 4 only");
    }
    int lh = h, lw = w;
    for (int l = 0; l < num_levels; ++l) {
        VFML_REQUIRE(lv0[l] != nullptr && (!lv1 || lv1[l] != nullptr), "corr_lookup: level %d is null", l);
        VFML_REQUIRE(lh > 0 && lw > 0, "corr_lookup: level %d is empty (h=%d, w=%d)", l, h, w);
        p.set[0].lvl[l] = lv0[l];
        p.set[1].lvl[l] = lv1 ? lv1[l] : nullptr;
        p.lh[l] = lh;
        p.lw[l] = lw;
        lh /= 2;
        lw /= 2;
    }
    p.set[0].coords = c0; p.set[0].out = o0;
    p.set[1].coords = c1; p.set[1].out = o1;
    p.num_levels = num_levels;
    p.h = h; p.w = w; p.n_frames = n_frames;
    p.layout = out_layout;
    p.out_pitch = out_pitch;
    p.out_col = out_col;
    cudaStream_t st = to_stream(s);
    if (vol_dtype == VFML_F32 && out_dtype == VFML_F32) return launch_lookup<float, float>(p, radius, st);
    if (vol_dtype == VFML_F32 && out_dtype == VFML_F16) return launch_lookup<float, __half>(p, radius, st);
    if (vol_dtype == VFML_F16 && out_dtype == VFML_F16) return launch_lookup<__half, __half>(p, radius, st);
    return launch_lookup<__half, float>(p, radius, st);
}

}  // namespace vfml

using namespace vfml;

extern "C" int vfml_corr_lookup(const void* const* levels, int num_levels, int vol_dtype, int n_frames, int h, int w,
                                int radius, const float* coords, void* out, int out_layout, int out_pitch, int out_col,
                                int out_dtype, vfml_stream_t s) {
    return run_lookup(levels, nullptr, num_levels, vol_dtype, n_frames, h, w, radius, coords, nullptr, out, nullptr,
                      out_layout, out_pitch, out_col, out_dtype, s);
}

extern "C" int vfml_corr_lookup_pair(const void* const* levels_fwd, const void* const* levels_bwd, int num_levels,
                                     int vol_dtype, int n_frames, int h, int w, const float* coords_fwd,
                                     const float* coords_bwd, void* out_fwd, void* out_bwd, int out_pitch, int out_col,
                                     int out_dtype, vfml_stream_t s) {
    VFML_REQUIRE(levels_bwd, "corr_lookup_pair: null pointer");
    return run_lookup(levels_fwd, levels_bwd, num_levels, vol_dtype, n_frames, h, w, LK_R, coords_fwd, coords_bwd, out_fwd,
                      out_bwd, VFML_ROWS, out_pitch, out_col, out_dtype, s);
}

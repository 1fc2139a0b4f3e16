// Depthwise k x k convolution (k = 15 / 7) over rows (NHWC) layout with the PCBlock residual and
// GELU fused:  y = gelu(x + dwconv_k(x) + b)      MOFNetStack/stack.py:25-26
//
// CUDA-core bound (225 MACs per output, no shared operand for tensor cores).  A CTA owns an
// 8 x 16 pixel tile of one image and a group of channels (lane <-> channel, or channel pair for
// fp16 storage, so every global and shared access is a 128-byte row).  The tile plus its halo and
// is staged in shared memory once with cp.async (84 KB for k = 15, so two CTAs share an SM and one
// CTA's staging overlaps the other's math); weights come through L1.  Each warp then produces one
// output row with a register sliding window: 16 outputs x k taps per loaded input value, i.e.
// ~8 FMAs per shared-memory load.
#include "common.cuh"

namespace vfml {

constexpr int DW_TY = 8, DW_TX = 16;

template <typename T> struct DwVec;
template <> struct DwVec<float> {
    static constexpr int VEC = 1;
    using ld_t = float;
    __device__ static void unpack(ld_t v, float (&o)[1]) { o[0] = v; }
    __device__ static ld_t pack(const float (&o)[1]) { return o[0]; }
};
template <> struct DwVec<__half> {
    static constexpr int VEC = 2;
    using ld_t = __half2;
    __device__ static void unpack(ld_t v, float (&o)[2]) { const float2 f = __half22float2(v); o[0] = f.x; o[1] = f.y; }
    __device__ static ld_t pack(const float (&o)[2]) { return __floats2half2_rn(o[0], o[1]); }
};

template <typename T, int K>
__global__ void __launch_bounds__(DW_TY * 32, 2)
dwconv_tile_kernel(const T* __restrict__ x, T* __restrict__ y, int h, int w, int C, int64_t px, int64_t py,
                   const float* __restrict__ wgt, const float* __restrict__ bias, int tiles_x, int tiles_y) {
    using V = DwVec<T>;
    using ld_t = typename V::ld_t;
    constexpr int VEC = V::VEC;
    constexpr int CG = 32 * VEC;                 // channels per CTA
    constexpr int R = K / 2;
    constexpr int TWX = DW_TX + K - 1, TWY = DW_TY + K - 1;

    extern __shared__ __align__(16) uint8_t dw_smem[];
    ld_t* tile = reinterpret_cast<ld_t*>(dw_smem);                                   // [TWY][TWX][32] of ld_t

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tile_id = blockIdx.x;
    const int tx_i = tile_id % tiles_x;
    const int ty_i = (tile_id / tiles_x) % tiles_y;
    const int img = tile_id / (tiles_x * tiles_y);
    const int c0 = blockIdx.y * CG + lane * VEC;
    const int x0 = tx_i * DW_TX, y0 = ty_i * DW_TY;
    const int64_t img_row0 = (int64_t)img * h * w;
    // a lane's channels are valid as a unit when C and the pitch keep the vector aligned (checked by the host)
    const bool ch_ok = c0 < C;

    // ---- stage the input tile with asynchronous 4-byte copies (zero-filled outside the image / channel
    //      range): every copy of the CTA is in flight at once, one wait at the end
    for (int i = warp; i < TWY * TWX; i += DW_TY) {
        const int ty = i / TWX, tx = i - ty * TWX;
        const int gy = y0 + ty - R, gx = x0 + tx - R;
        const bool ok = ch_ok && gy >= 0 && gy < h && gx >= 0 && gx < w;
        const T* src = ok ? x + (img_row0 + (int64_t)gy * w + gx) * px + c0 : x;
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(tile + i * 32 + lane);
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(ok ? 4 : 0) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();

    // ---- one output row per warp, DW_TX outputs x VEC channels per lane
    float acc[DW_TX][VEC];
#pragma unroll
    for (int o = 0; o < DW_TX; ++o)
#pragma unroll
        for (int v = 0; v < VEC; ++v) acc[o][v] = 0.0f;

#pragma unroll 1
    for (int ky = 0; ky < K; ++ky) {
        float wr[K][VEC];
#pragma unroll
        for (int kx = 0; kx < K; ++kx)
#pragma unroll
            for (int v = 0; v < VEC; ++v) wr[kx][v] = ch_ok ? __ldg(wgt + (int64_t)(ky * K + kx) * C + c0 + v) : 0.0f;   // L1-resident
        const ld_t* row = tile + (size_t)(warp + ky) * TWX * 32 + lane;
#pragma unroll
        for (int i = 0; i < TWX; ++i) {
            float xv[VEC];
            V::unpack(row[i * 32], xv);
#pragma unroll
            for (int kx = 0; kx < K; ++kx) {
                const int o = i - kx;
                if (o >= 0 && o < DW_TX) {
                    if constexpr (VEC == 2) {
                        // both channels of the lane in one packed FFMA2 (fma.rn.f32x2): half the issue slots
                        asm("{\n.reg .b64 ra, rb, rc;\nmov.b64 ra, {%2, %3};\nmov.b64 rb, {%4, %5};\nmov.b64 rc, {%0, %1};\n"
                            "fma.rn.f32x2 rc, ra, rb, rc;\nmov.b64 {%0, %1}, rc;\n}"
                            : "+f"(acc[o][0]), "+f"(acc[o][VEC - 1])
                            : "f"(wr[kx][0]), "f"(wr[kx][VEC - 1]), "f"(xv[0]), "f"(xv[VEC - 1]));
                    } else {
#pragma unroll
                        for (int v = 0; v < VEC; ++v) acc[o][v] = fmaf(wr[kx][v], xv[v], acc[o][v]);
                    }
                }
            }
        }
    }

    // ---- residual + bias + GELU, coalesced store
    const int gy = y0 + warp;
    if (gy >= h || !ch_ok) return;
    float b[VEC];
#pragma unroll
    for (int v = 0; v < VEC; ++v) b[v] = (c0 + v < C) ? __ldg(bias + c0 + v) : 0.0f;
    const ld_t* crow = tile + (size_t)(warp + R) * TWX * 32 + lane;
#pragma unroll
    for (int o = 0; o < DW_TX; ++o) {
        const int gx = x0 + o;
        if (gx >= w) break;
        float xc[VEC], out[VEC];
        V::unpack(crow[(o + R) * 32], xc);
#pragma unroll
        for (int v = 0; v < VEC; ++v) out[v] = gelu_erf(xc[v] + (acc[o][v] + b[v]));
        *reinterpret_cast<ld_t*>(y + (img_row0 + (int64_t)gy * w + gx) * py + c0) = V::pack(out);
    }
}

template <typename T, int K>
static int launch_dw_tile(const T* x, T* y, int n_img, int h, int w, int C, int64_t px, int64_t py, const float* wgt,
                          const float* bias, cudaStream_t st) {
    using V = DwVec<T>;
    constexpr int TWX = DW_TX + K - 1, TWY = DW_TY + K - 1;
    const size_t smem = (size_t)TWY * TWX * 32 * sizeof(typename V::ld_t);
    auto kern = dwconv_tile_kernel<T, K>;
    static PerDevice<bool> attr_done;
    if (!attr_done.get()) {
        VFML_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_done.get() = true;
    }
    const int tiles_x = cdiv(w, DW_TX), tiles_y = cdiv(h, DW_TY);
    dim3 grid(tiles_x * tiles_y * n_img, cdiv(C, 32 * V::VEC));
    kern<<<grid, DW_TY * 32, smem, st>>>(x, y, h, w, C, px, py, wgt, bias, tiles_x, tiles_y);
    VFML_LAUNCH_CHECK("dwconv_tile");
    return VFML_OK;
}

// returns 1 if handled by the tiled kernel, 0 if the caller must use the generic kernel, < 0 on error
int dwconv_tiled(const void* x, void* y, int n_img, int h, int w, int C, int64_t px, int64_t py, const float* wgt,
                 const float* bias, int k, int dtype, cudaStream_t st) {
    if (k != 15 && k != 7) return 0;
    if (dtype == VFML_F16) {
        // channel pairs are loaded as half2: needs even C, even pitches and 4-byte aligned bases
        if ((C & 1) || (px & 1) || (py & 1) || (reinterpret_cast<uintptr_t>(x) & 3) || (reinterpret_cast<uintptr_t>(y) & 3)) return 0;
        const __half* xi = reinterpret_cast<const __half*>(x);
        __half* yo = reinterpret_cast<__half*>(y);
        const int rc = k == 15 ? launch_dw_tile<__half, 15>(xi, yo, n_img, h, w, C, px, py, wgt, bias, st)
                               : launch_dw_tile<__half, 7>(xi, yo, n_img, h, w, C, px, py, wgt, bias, st);
        return rc ? rc : 1;
    }
    const float* xi = reinterpret_cast<const float*>(x);
    float* yo = reinterpret_cast<float*>(y);
    const int rc = k == 15 ? launch_dw_tile<float, 15>(xi, yo, n_img, h, w, C, px, py, wgt, bias, st)
                           : launch_dw_tile<float, 7>(xi, yo, n_img, h, w, C, px, py, wgt, bias, st);
    return rc ? rc : 1;
}

}  // namespace vfml

// Depthwise k x k convolution (k = 15 / 7) on the tensor cores, fp16 tier, rows (NHWC) layout, with the
// PCBlock residual and GELU fused:   y = gelu(x + dwconv_k(x) + b)      MOFNetStack/stack.py:25-26
//
// A depthwise convolution has no shared operand across channels, but per channel it is a 2-D
// correlation, and a 1-D correlation along x is a product with a banded Toeplitz matrix:
//     out[y][xo] = sum_dy  sum_xi  in[y + dy - R][xi] * Wband_dy[xi][xo],   Wband_dy[xi][xo] = w[dy][xi - xo + R].
// Per channel and dy this is D[16 rows y][8 xo] += A[16 rows][16 xi] * B[16 xi][8 xo] (mma.m16n8k16, fp16 in,
// fp32 accumulate): A is the channel's input plane read with ldmatrix at a row offset of dy (the shift is
// free), B is the band, whose fragments depend only on (xi block - xo block) -> 4 variants per dy.  An
// 8-wide xo block needs 8 + 2R <= 22 inputs = two 16-wide k blocks, so 32/15 = 2.1x (k = 15) of the MACs are
// padding -- on a pipe that is > 10x faster than the FFMA path the previous kernel used.
//
// CTA = 64 rows x 32 columns x 8 channels of one image (256 threads, warp <-> channel):
//   1. transpose-load: 16-byte channel vectors of the (64+k-1) x 48 input window -> 8 channel planes in smem
//      (row pitch 112 B: conflict-free ldmatrix); the packed band table of the 8 channels (prepacked once by
//      vfml_dwconv_pack_weights, 30 KB for k = 15) is copied next to it;
//   2. k x 32 MMAs per warp, accumulators (4 m-blocks x 4 n-blocks) in registers;
//   3. epilogue in place: centre value + bias + accumulator -> polynomial GELU -> fp16 back into the plane;
//   4. transpose-store: 8 planes -> 16-byte channel vectors, coalesced rows.
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "gelu_poly.cuh"

namespace vfml {

#ifndef VFML_DW_K8
#define VFML_DW_K8 1          // k16 + k8 MMA split (measured -2 %); -DVFML_DW_K8=0 restores two k16 MMAs per block
#endif
constexpr int DT_TW = 32, DT_CG = 8;        // tile height TH (64 or 32 rows) is a template parameter
constexpr int DT_INW = 48;                 // input columns staged: x = n0 - 8 .. n0 + 39
constexpr int DT_PITCH = 112;              // bytes per plane row (48 halfs + 8 pad)
constexpr int DT_WROW = 256;               // bytes per (channel, dy) band row: even copy [64 halfs] + odd copy [64 halfs]
constexpr int DT_WLEFT = 24;               // tap t lives at index t + 24 of the even copy

template <int K, int TH> struct DtCfg {
    static constexpr int R = K / 2;
    static constexpr int INH = TH + K - 1;
    static constexpr int PLANE = INH * DT_PITCH;
    static constexpr int WCH = K * DT_WROW + 16;          // +16 B: channels land in different banks
    static constexpr int SMEM = DT_CG * PLANE + DT_CG * WCH;
};

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void mma1688(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};\n"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a0), "r"(a1), "r"(b0));
}
__device__ __forceinline__ uint32_t dt_lds32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];\n" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ void dt_sts32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;\n" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void dt_sts16(uint32_t a, uint32_t v) {
    asm volatile("st.shared.u16 [%0], %1;\n" ::"r"(a), "h"((unsigned short)v) : "memory");
}
__device__ __forceinline__ uint32_t dt_lds16(uint32_t a) {
    unsigned short v;
    asm volatile("ld.shared.u16 %0, [%1];\n" : "=h"(v) : "r"(a));
    return v;
}

// Packed band table: [C/8 groups][8 channels][K dy][even copy 64 | odd copy 64] halfs, zero outside the K taps.
// even[u] = w[dy][u - 24];  odd[u] = even[u + 1]  (so that a tap pair starting at an odd index is one aligned word).
template <int K>
__global__ void dwconv_pack_kernel(const float* __restrict__ wgt, __half* __restrict__ packed, int C, int groups) {
    const int total = groups * DT_CG * K * 128;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int u = i & 63, copy = (i >> 6) & 1;
        const int row = i >> 7;                  // (group*8 + ch)*K + dy
        const int dy = row % K, c = row / K;
        const int t = u + copy - DT_WLEFT;
        float v = 0.0f;
        if (t >= 0 && t < K && c < C) v = wgt[(int64_t)(dy * K + t) * C + c];
        packed[i] = __float2half_rn(v);
    }
}

// TMA = true: the input window arrives through the tensor memory accelerator instead of per-thread 16-byte loads.  Those
// loads touch one 128-byte line per thread (pixel pitch >= 256 B), i.e. one LSU data-stage wavefront per pixel, and the
// ncu capture (profiles/r1_dwconv_tc_ncu_summary.txt) shows the LSU data pipe as the busiest unit of this kernel (60 % for
// k = 7, with 26 points of it from global accesses).  One 4-D box {8 channels, 48 columns, rows, 1 image} per half window
// lands pixel-major in the (not yet used) plane region, zero-filled outside the image and beyond channel C; the threads
// then pull it into registers with 128-bit shared loads (1/8 wavefront per pixel) and scatter it into the planes as before.
// TMA bit 1: the output tile leaves the same way (planes -> pixel-major chunks staged in the band-table region, which is
// free after the MMA phase -> bulk tensor stores, clipped at the image border and at channel C by the unit).
// TH = 64: 2 CTAs per SM (100 KB, 128 registers).  TH = 32: 3 CTAs per SM (72 KB, 85 registers) and twice the CTAs: the
// grid of the benchmark shapes (600 / 768 / 984 CTAs on 296 slots) loses up to a third of its run time to the last
// partial wave, and a CTA's phases (load, MMA, epilogue, store) only overlap with those of its SM neighbours.
template <int K, int TMA, int TH>
__global__ void __launch_bounds__(256, TH == 64 ? 2 : 3)
dwconv_tc_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapY, const __half* __restrict__ x, __half* __restrict__ y, int h, int w,
                 int C, int64_t px, int64_t py, const __half* __restrict__ packed, const float* __restrict__ bias, int tiles_x,
                 int tiles_y) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // lets a PDL-launched successor start its prologue
    using Cfg = DtCfg<K, TH>;
    constexpr int MB = TH / 16;                       // 16-row MMA blocks per tile
    constexpr int R = Cfg::R, INH = Cfg::INH, PLANE = Cfg::PLANE, WCH = Cfg::WCH;
    extern __shared__ __align__(128) uint8_t dt_smem[];
    const uint32_t s_planes = (uint32_t)__cvta_generic_to_shared(dt_smem);
    const uint32_t s_w = s_planes + DT_CG * PLANE;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int tile_id = blockIdx.x;
    const int tx_i = tile_id % tiles_x;
    const int ty_i = (tile_id / tiles_x) % tiles_y;
    const int img = tile_id / (tiles_x * tiles_y);
    const int grp = blockIdx.y, c0 = grp * DT_CG;
    const int n0 = tx_i * DT_TW, y0 = ty_i * TH;
    const int64_t img_row0 = (int64_t)img * h * w;
    __shared__ uint64_t tma_bar;
    if (TMA & 1) {
        if (tid == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&tma_bar)));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();            // nobody may poll the barrier before it is initialised (the word holds a previous kernel's data)
        if (tid == 0) {
            constexpr int HALF = (INH + 1) / 2;                 // two boxes of HALF rows (the second may overhang the
            const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&tma_bar);      // window: rows the kernel never reads)
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(2u * HALF * DT_INW * 16u) : "memory");
#pragma unroll
            for (int b = 0; b < 2; ++b)
                asm volatile(
                    "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                    ::"r"(s_planes + b * HALF * DT_INW * 16), "l"(&mapX), "r"(bar), "r"(c0), "r"(n0 - 8), "r"(y0 - R + b * HALF), "r"(img)
                    : "memory");
        }
    }

    // ---- 1a. band table of this channel group: K*16 uint4 per channel, asynchronous (lands while 1b runs)
    {
        const uint4* src = reinterpret_cast<const uint4*>(packed + (int64_t)grp * DT_CG * K * 128);
        for (int i = tid; i < DT_CG * K * 16; i += 256) {
            const int ch = i / (K * 16), o = i - ch * (K * 16);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s_w + ch * WCH + o * 16), "l"(src + i) : "memory");
        }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    }
    // ---- 1b. transpose-load the input window: pixel-major 16-byte channel vectors -> 8 channel planes
    {
        constexpr int NPIX = INH * DT_INW;
        constexpr int CH = (NPIX + 255) / 256;             // every load of the thread in flight at once
        if (TMA & 1) {
            const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&tma_bar);
            asm volatile(
                "{\n.reg .pred p;\nDW_WAIT:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra DW_DONE;\nbra DW_WAIT;\nDW_DONE:\n}\n"
                ::"r"(bar) : "memory");
        }
        for (int p0 = 0; p0 < NPIX; p0 += 256 * CH) {
            uint4 v[CH];
#pragma unroll
            for (int i = 0; i < CH; ++i) {
                const int p = p0 + i * 256 + tid;
                v[i] = make_uint4(0u, 0u, 0u, 0u);
                if (TMA & 1) {
                    if (p < NPIX)
                        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];\n"
                                     : "=r"(v[i].x), "=r"(v[i].y), "=r"(v[i].z), "=r"(v[i].w) : "r"(s_planes + p * 16));
                } else {
                    const int row = p / DT_INW, xi = p - row * DT_INW;
                    const int gy = y0 - R + row, gx = n0 - 8 + xi;
                    if (p < NPIX && gy >= 0 && gy < h && gx >= 0 && gx < w)
                        v[i] = __ldg(reinterpret_cast<const uint4*>(x + (img_row0 + (int64_t)gy * w + gx) * px + c0));
                }
            }
            if (TMA & 1) __syncthreads();          // the pixel-major window aliases the planes: everyone has read it
#pragma unroll
            for (int i = 0; i < CH; ++i) {
                const int p = p0 + i * 256 + tid;
                if (p < NPIX) {
                    const int row = p / DT_INW, xi = p - row * DT_INW;
                    const uint32_t a = s_planes + row * DT_PITCH + xi * 2;
                    dt_sts16(a + 0 * PLANE, v[i].x); dt_sts16(a + 1 * PLANE, v[i].x >> 16);
                    dt_sts16(a + 2 * PLANE, v[i].y); dt_sts16(a + 3 * PLANE, v[i].y >> 16);
                    dt_sts16(a + 4 * PLANE, v[i].z); dt_sts16(a + 5 * PLANE, v[i].z >> 16);
                    dt_sts16(a + 6 * PLANE, v[i].w); dt_sts16(a + 7 * PLANE, v[i].w >> 16);
                }
            }
        }
    }
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    __syncthreads();

    // ---- 2. MMAs: warp <-> channel; 4 m-blocks (16 rows) x 4 n-blocks (8 columns)
    const uint32_t plane = s_planes + warp * PLANE;
    // ldmatrix.x4 lane address: matrices (rows 0-7 | 8-15) x (k 0-7 | 8-15)
    const uint32_t a_lane = plane + ((lane & 7) + ((lane >> 3) & 1) * 8) * DT_PITCH + (lane >> 4) * 16;
    // B fragment words of the 4 band variants (delta = 16*kb - 8*s in {0, 16, -8, 8}): b0 = taps of rows
    // k = 2*(lane%4) + {0,1}, column n = lane/4; b1 = rows k + 8.  Tap index t = k - n + delta - 8 + R.
    // All eight (variant, half) addresses differ from w_base by compile-time constants: 8 r + delta is even, so the parity
    // (even / odd copy) is the lane's own and the rest is a byte offset of 2 (8 r + delta).
    const int u_base = 2 * (lane & 3) - (lane >> 2) - 8 + R + DT_WLEFT;
    const uint32_t w_base = s_w + warp * WCH + ((u_base & 1) ? 128 : 0) + (u_base & ~1) * 2;
    constexpr int W_OFF[4][2] = {{0, 16}, {32, 48}, {-16, 0}, {16, 32}};       // 2 * (8 r + {0, 16, -8, 8}[d])
#define w_lane(d, r) (w_base + W_OFF[d][r])
    float acc[MB][4][4];
#pragma unroll
    for (int mb = 0; mb < MB; ++mb)
#pragma unroll
        for (int s = 0; s < 4; ++s)
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[mb][s][i] = 0.0f;

#pragma unroll
    for (int dy = 0; dy < K; ++dy) {
        uint32_t b[4][2];
#pragma unroll
        for (int d = 0; d < 4; ++d) {
            b[d][0] = dt_lds32(w_lane(d, 0) + dy * DT_WROW);
            b[d][1] = dt_lds32(w_lane(d, 1) + dy * DT_WROW);
        }
#pragma unroll
        for (int mb = 0; mb < MB; ++mb) {
            uint32_t a0[4], a1[4], a2[4];
            const uint32_t ar = a_lane + (mb * 16 + dy) * DT_PITCH;
            ldsm_x4(a0, ar);
            ldsm_x4(a1, ar + 32);
            ldsm_x4(a2, ar + 64);
            // n-block s uses k-blocks (s/2, s/2 + 1) with variants (0,1) for even s and (2,3) for odd s
#if VFML_DW_K8
            // An 8-column output block reads 8 + 2R <= 22 inputs, which lie in three 8-column chunks of the plane: one k16
            // MMA on the aligned chunk pair and one k8 MMA on the third chunk (the other half of that k-block only ever
            // met zero band entries): 12 instead of 16 k8-equivalents per row block on the legacy tensor pipe, which the
            // ncu capture puts at ~10 cycles per m16n8k16 and sub-partition, i.e. half of this kernel's run time.
            mma16816(acc[mb][0], a0, b[0][0], b[0][1]); mma1688(acc[mb][0], a1[0], a1[1], b[1][0]);
            mma1688(acc[mb][1], a0[2], a0[3], b[2][1]); mma16816(acc[mb][1], a1, b[3][0], b[3][1]);
            mma16816(acc[mb][2], a1, b[0][0], b[0][1]); mma1688(acc[mb][2], a2[0], a2[1], b[1][0]);
            mma1688(acc[mb][3], a1[2], a1[3], b[2][1]); mma16816(acc[mb][3], a2, b[3][0], b[3][1]);
#else
            mma16816(acc[mb][0], a0, b[0][0], b[0][1]); mma16816(acc[mb][0], a1, b[1][0], b[1][1]);
            mma16816(acc[mb][1], a0, b[2][0], b[2][1]); mma16816(acc[mb][1], a1, b[3][0], b[3][1]);
            mma16816(acc[mb][2], a1, b[0][0], b[0][1]); mma16816(acc[mb][2], a2, b[1][0], b[1][1]);
            mma16816(acc[mb][3], a1, b[2][0], b[2][1]); mma16816(acc[mb][3], a2, b[3][0], b[3][1]);
#endif
        }
    }


    // ---- 3. epilogue in place (this warp is the only user of its plane): gelu(x + conv + bias)
    {
        const int c = c0 + warp;
        const float bv = c < C ? __ldg(bias + c) : 0.0f;
#pragma unroll
        for (int mb = 0; mb < MB; ++mb)
#pragma unroll
            for (int s = 0; s < 4; ++s)
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    const int row = mb * 16 + (lane >> 2) + hh * 8;          // output row within the tile
                    const int col = s * 8 + 2 * (lane & 3);                  // output column pair
                    const uint32_t a = plane + (row + R) * DT_PITCH + (col + 8) * 2;
                    const uint32_t xc = dt_lds32(a);
                    const float2 xf = __half22float2(*reinterpret_cast<const __half2*>(&xc));
                    float2 v = make_float2(xf.x + (acc[mb][s][hh * 2] + bv), xf.y + (acc[mb][s][hh * 2 + 1] + bv));
                    v = gelu_fast2(v);
                    const __half2 o = __floats2half2_rn(v.x, v.y);
                    dt_sts32(a, *reinterpret_cast<const uint32_t*>(&o));
                }
    }
    __syncthreads();

    // ---- 4. transpose-store: 8 planes -> one 16-byte channel vector per pixel
    if (TMA & 2) {
        constexpr int CR = K == 15 ? 16 : 8;                 // rows per chunk: two chunk buffers fit in the band-table region
        constexpr int CHUNK = CR * DT_TW * 16;
        static_assert(2 * CHUNK <= DT_CG * WCH, "store staging does not fit in the band-table region");
#pragma unroll 1
        for (int c = 0; c < TH / CR; ++c) {
            if (y0 + c * CR >= h) break;
            const uint32_t stage = s_w + (c & 1) * CHUNK;
            if (c >= 2) {                                    // the store issued two chunks ago has read this buffer
                if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                __syncthreads();
            }
            for (int p = tid; p < CR * DT_TW; p += 256) {
                const int row = c * CR + (p >> 5), col = p & 31;
                const uint32_t a = s_planes + (row + R) * DT_PITCH + (col + 8) * 2;
                uint4 v;
                v.x = dt_lds16(a + 0 * PLANE) | (dt_lds16(a + 1 * PLANE) << 16);
                v.y = dt_lds16(a + 2 * PLANE) | (dt_lds16(a + 3 * PLANE) << 16);
                v.z = dt_lds16(a + 4 * PLANE) | (dt_lds16(a + 5 * PLANE) << 16);
                v.w = dt_lds16(a + 6 * PLANE) | (dt_lds16(a + 7 * PLANE) << 16);
                asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};\n" ::"r"(stage + p * 16), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            if (tid == 0) {
                asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
                             ::"l"(&mapY), "r"(stage), "r"(c0), "r"(n0), "r"(y0 + c * CR), "r"(img) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        }
        if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        return;
    }
    for (int p = tid; p < TH * DT_TW; p += 256) {
        const int row = p >> 5, col = p & 31;
        const int gy = y0 + row, gx = n0 + col;
        if (gy < h && gx < w) {
            const uint32_t a = s_planes + (row + R) * DT_PITCH + (col + 8) * 2;
            uint4 v;
            v.x = dt_lds16(a + 0 * PLANE) | (dt_lds16(a + 1 * PLANE) << 16);
            v.y = dt_lds16(a + 2 * PLANE) | (dt_lds16(a + 3 * PLANE) << 16);
            v.z = dt_lds16(a + 4 * PLANE) | (dt_lds16(a + 5 * PLANE) << 16);
            v.w = dt_lds16(a + 6 * PLANE) | (dt_lds16(a + 7 * PLANE) << 16);
            *reinterpret_cast<uint4*>(y + (img_row0 + (int64_t)gy * w + gx) * py + c0) = v;
        }
    }
}

// 4-D map of a rows-layout activation [n_img][h][w][C] (pitch px halfs per pixel): box = {8 channels, 48 columns, rows, 1}
static int make_window_map(CUtensorMap* map, const void* base, int n_img, int h, int w, int C, int64_t px, int box_cols,
                           int box_rows) {
    typedef CUresult (*encode_fn_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static encode_fn_t encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
            set_error("dwconv_tc: cuTensorMapEncodeTiled is not available from the driver (%s)", cudaGetErrorString(e));
            return VFML_ERR_CUDA;
        }
        encode = reinterpret_cast<encode_fn_t>(fn);
    }
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n_img};
    cuuint64_t strides[3] = {(cuuint64_t)px * 2, (cuuint64_t)px * 2 * w, (cuuint64_t)px * 2 * w * h};
    cuuint32_t box[4] = {(cuuint32_t)DT_CG, (cuuint32_t)box_cols, (cuuint32_t)box_rows, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("dwconv_tc: cuTensorMapEncodeTiled failed: CUresult %d (n=%d h=%d w=%d C=%d pitch=%lld)", (int)r, n_img, h, w, C,
                  (long long)px);
        return VFML_ERR_CUDA;
    }
    return VFML_OK;
}

template <int K, int TH>
static int launch_dw_tc(const __half* x, __half* y, int n_img, int h, int w, int C, int64_t px, int64_t py,
                        const __half* packed, const float* bias, int mode, cudaStream_t st) {
    typedef void (*kern_t)(const CUtensorMap, const CUtensorMap, const __half*, __half*, int, int, int, int64_t, int64_t, const __half*,
                           const float*, int, int);
    static const kern_t kerns[4] = {dwconv_tc_kernel<K, 0, TH>, dwconv_tc_kernel<K, 1, TH>, dwconv_tc_kernel<K, 2, TH>,
                                    dwconv_tc_kernel<K, 3, TH>};
    using Cfg = DtCfg<K, TH>;
    static PerDevice<bool> attr_done;
    if (!attr_done.get()) {
        for (int i = 0; i < 4; ++i)
            VFML_CUDA(cudaFuncSetAttribute(kerns[i], cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
        attr_done.get() = true;
    }
    const int tiles_x = cdiv(w, DT_TW), tiles_y = cdiv(h, TH);
    dim3 grid(tiles_x * tiles_y * n_img, cdiv(C, DT_CG));
    CUtensorMap mapX{}, mapY{};
    if (mode & 1) {
        const int rc = make_window_map(&mapX, x, n_img, h, w, C, px, DT_INW, (Cfg::INH + 1) / 2);
        if (rc) return rc;
    }
    if (mode & 2) {
        const int rc = make_window_map(&mapY, y, n_img, h, w, C, py, DT_TW, K == 15 ? 16 : 8);
        if (rc) return rc;
    }
    kerns[mode]<<<grid, 256, Cfg::SMEM, st>>>(mapX, mapY, x, y, h, w, C, px, py, packed, bias, tiles_x, tiles_y);
    VFML_LAUNCH_CHECK("dwconv_tc");
    return VFML_OK;
}

// Tile height: VFML_DW_TH=32|64 forces it; otherwise (VFML_DW_TH=0: auto) the one with the cheaper last wave under the model
// "rounds x CTA cost", a 32-row CTA costing 0.56 of a 64-row one (half the MMAs, 46 instead of 78 window rows for k = 15).
template <int K>
static int launch_dw_tc_any(const __half* x, __half* y, int n_img, int h, int w, int C, int64_t px, int64_t py,
                            const __half* packed, const float* bias, cudaStream_t st) {
    // VFML_DW_TMA: bit 0 = window load through TMA, bit 1 = tile store through TMA.  Default 3 and 32-row tiles: measured
    // 181 -> 136 us over the four C1 depthwise shapes (scripts/micro/dw_ab.sh, profiles/r1_s18_dw_ab.txt); 0 / 64 = the
    // previous configuration, kept for A/B.
    static const int mode = getenv("VFML_DW_TMA") ? atoi(getenv("VFML_DW_TMA")) & 3 : 3;
    static const int th_env = getenv("VFML_DW_TH") ? atoi(getenv("VFML_DW_TH")) : 32;
    const int num_sms = device_sms();
    int th = th_env;
    if (th != 32 && th != 64) {
        const int64_t per = (int64_t)cdiv(w, DT_TW) * n_img * cdiv(C, DT_CG);
        const double c64 = (double)cdiv(per * cdiv(h, 64), 2 * num_sms), c32 = 0.56 * cdiv(per * cdiv(h, 32), 3 * num_sms);
        th = c32 < c64 ? 32 : 64;
    }
    return th == 32 ? launch_dw_tc<K, 32>(x, y, n_img, h, w, C, px, py, packed, bias, mode, st)
                    : launch_dw_tc<K, 64>(x, y, n_img, h, w, C, px, py, packed, bias, mode, st);
}

}  // namespace vfml

using namespace vfml;

extern "C" size_t vfml_dwconv_packed_bytes(int C, int k) {
    if (C <= 0 || (k != 7 && k != 15)) return 0;
    return (size_t)cdiv(C, DT_CG) * DT_CG * k * 128 * sizeof(__half);
}

extern "C" int vfml_dwconv_pack_weights(const float* weight, int C, int k, void* packed, vfml_stream_t s) {
    VFML_REQUIRE(weight && packed, "dwconv_pack_weights: null pointer");
    VFML_REQUIRE(C > 0 && (k == 7 || k == 15), "dwconv_pack_weights: C=%d k=%d (k must be 7 or 15)", C, k);
    VFML_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 15) == 0, "dwconv_pack_weights: packed must be 16-byte aligned");
    const int groups = cdiv(C, DT_CG);
    const int total = groups * DT_CG * k * 128;
    const int blocks = std::min(cdiv(total, 256), device_sms() * 8);
    if (k == 15) dwconv_pack_kernel<15><<<blocks, 256, 0, to_stream(s)>>>(weight, reinterpret_cast<__half*>(packed), C, groups);
    else dwconv_pack_kernel<7><<<blocks, 256, 0, to_stream(s)>>>(weight, reinterpret_cast<__half*>(packed), C, groups);
    VFML_LAUNCH_CHECK("dwconv_pack");
    return VFML_OK;
}

extern "C" int vfml_dwconv_tc_res_gelu(const void* x, void* y, int n_img, int h, int w, int C, int64_t pitch_x,
                                       int64_t pitch_y, const void* packed, const float* bias, int k, vfml_stream_t s) {
    VFML_REQUIRE(n_img >= 0 && h > 0 && w > 0 && C > 0, "dwconv_tc: bad shape");
    VFML_REQUIRE(k == 7 || k == 15, "dwconv_tc: k=%d (must be 7 or 15)", k);
    if (n_img == 0) return VFML_OK;
    VFML_REQUIRE(x && y && packed && bias, "dwconv_tc: null pointer");
    VFML_REQUIRE(x != y, "dwconv_tc: in-place operation is not supported (tiles read their neighbours' halo)");
    const int cpad = cdiv(C, DT_CG) * DT_CG;
    VFML_REQUIRE(pitch_x >= cpad && pitch_y >= cpad && (pitch_x % 8) == 0 && (pitch_y % 8) == 0,
                 "dwconv_tc: pitches (%lld, %lld) must be multiples of 8 and >= %d", (long long)pitch_x, (long long)pitch_y, cpad);
    VFML_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(packed)) & 15) == 0,
                 "dwconv_tc: x, y and packed must be 16-byte aligned");
    const __half* xi = reinterpret_cast<const __half*>(x);
    __half* yo = reinterpret_cast<__half*>(y);
    const __half* pk = reinterpret_cast<const __half*>(packed);
    return k == 15 ? launch_dw_tc_any<15>(xi, yo, n_img, h, w, C, pitch_x, pitch_y, pk, bias, to_stream(s))
                   : launch_dw_tc_any<7>(xi, yo, n_img, h, w, C, pitch_x, pitch_y, pk, bias, to_stream(s));
}

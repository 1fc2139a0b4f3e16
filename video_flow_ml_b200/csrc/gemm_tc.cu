// fp16 tensor-core tier of the pointwise-convolution / matmul op: TMA -> shared (128B swizzle) ->
// tcgen05.mma (fp32 accumulators in TMEM) -> tcgen05.ld -> fused epilogue -> global.
//
// D[m, n] = epi(alpha * sum_k A[m, k] * B[n, k]); A = activations in rows layout (K contiguous),
// B = conv weight [Cout, Cin].  Same contract as the FFMA tier (gemm_ffma.cu); the reference runs
// these layers under fp16 autocast on CUDA (MOFNetStack/network.py:149-192).
//
// Persistent kernel, one CTA per SM, static round-robin over 128 x BN output tiles (BN runtime,
// multiple of 16, <= 256).  Warp 0 lane 0 issues TMA loads of 128x64 (A) and BNx64 (B) fp16 boxes
// into a ring of stages that keeps running across tiles; warp 1 lane 0 issues tcgen05.mma (M=128,
// N=BN, K=16, 4 per stage) into one of TWO 256-column TMEM accumulators; warps 2-9 drain the other
// accumulator (each warp its own 32-lane TMEM quarter and half of the columns), apply
// bias / residual / GELU / ReLU / the k=1 depthwise stage and store, so the epilogue of tile i
// overlaps the MMA main loop of tile i+1.  Out-of-range rows, columns and the K tail are
// zero-filled by TMA, so no shape needs padding.
//
// The epilogue is the co-bottleneck of these small-K GEMMs (K = 128..768), so its math is packed:
// fp32 pairs through fma.rn.f32x2, and GELU in its tanh form (gelu_poly.cuh: one MUFU tanh.approx plus six packed
// FMA-pipe operations per pair; the MUFU-free degree-12 polynomial is kept behind -DVFML_GELU_MUFU=0).
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "gelu_poly.cuh"
#include "tcgen05.cuh"

namespace vfml {

constexpr int TC_BM = 128, TC_BK = 64;
#ifndef VFML_TC_EPI_WARPS
#define VFML_TC_EPI_WARPS 12
#endif
#ifndef VFML_TC_EPI_V1
#define VFML_TC_EPI_V1 1          // 1 = lane-per-row epilogue (default); 0 = the row-segment variant (16x256b TMEM loads + quad
                                  // transposes, 8 rows x 128 B per access): measured equal within noise (profiles/r2_gemm_epilogue_ab.txt),
                                  // because the epilogue is bound by its math issue slots and by ~19 B/clk/SM of store bandwidth, not by
                                  // the number of L2 requests
#endif
constexpr int TC_EPI_WARPS = VFML_TC_EPI_WARPS;      // multiple of 4: warp w may only read TMEM lanes 32 (w % 4) ..
constexpr int TC_THREADS = 64 + TC_EPI_WARPS * 32;
constexpr int TC_MAX_STAGES = 8;
constexpr int TC_VEC_MAX = 2048;      // columns of bias / dw vectors staged in shared memory

struct TcParams {
    void* D;
    const float* bias;
    const __half* res;
    const float* dw_w;
    const float* dw_b;
    int64_t ldd, ldr, strideD, strideR;
    int M, N, K, batch;
    int BN, stages, m_tiles, n_tiles;
    float alpha;
    uint32_t epi;
    int out_f32;
    int vec_in_smem;
    int b_res;           // weight-stationary mode: this CTA's B tile (all K blocks) stays in shared memory
    int b_is_const;      // B (and bias / dw vectors) are weights: may be fetched before the preceding kernel has finished
    long long* trace;    // debug (VFML_GEMM_TRACE=1): per-CTA cycle counters, 16 per CTA
};

// SPEC >= 0: the epilogue flag set (a subset of BIAS | RES | GELU | DW1) is a compile-time constant, the output is fp16,
// alpha == 1 and the per-column vectors are in shared memory: the per-chunk epilogue becomes one basic block (no uniform
// branches or constant-bank reloads between its stages), which the 3 epilogue warps per scheduler need to hide their
// fixed-latency dependency stalls.  SPEC < 0: everything decided at run time.
template <int SPEC>
__global__ void __launch_bounds__(TC_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, const TcParams p) {
    // programmatic dependent launch: the next kernel of the stream may start its prologue as soon as every CTA of this
    // grid has got here (this grid is persistent, <= 1 CTA per SM: nothing of it is left to schedule)
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int BN = p.BN, stages = p.stages;
    const uint32_t a_bytes = TC_BM * TC_BK * 2, b_bytes = (uint32_t)BN * TC_BK * 2;
    const uint32_t b_stride = (b_bytes + 1023) & ~1023u;
    const int nkb = (p.K + TC_BK - 1) / TC_BK;
    // b_res: [nkb resident B blocks][A ring]; else [A ring][B ring]
    const int b_slots = p.b_res ? nkb : stages;
    uint8_t* sB = p.b_res ? smem : smem + (size_t)stages * a_bytes;
    uint8_t* sA = p.b_res ? smem + (size_t)b_slots * b_stride : smem;
    float* s_vec = reinterpret_cast<float*>(smem + (size_t)stages * a_bytes + (size_t)b_slots * b_stride);   // [3][vec_cols]: bias, dw_w, dw_b
    __shared__ uint64_t full_bar[TC_MAX_STAGES], empty_bar[TC_MAX_STAGES], acc_full[2], acc_empty[2], b_full;
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_per_batch = p.m_tiles * p.n_tiles;
    const int num_tiles = tiles_per_batch * p.batch;
    // tile walk.  Default: round-robin over all (batch, m, n) tiles.  b_res (batch == 1): the CTA keeps ONE n-tile
    // (its weights never leave shared memory) and strides over the m-tiles; gridDim.x is a multiple of n_tiles.
    const int t_first = p.b_res ? (int)blockIdx.x / p.n_tiles : (int)blockIdx.x;
    const int t_step = p.b_res ? (int)gridDim.x / p.n_tiles : (int)gridDim.x;
    const int t_end = p.b_res ? p.m_tiles : num_tiles;
    const int n_fixed = p.b_res ? ((int)blockIdx.x % p.n_tiles) * BN : 0;
    auto tile_coords = [&](int tile, int& bz, int& m0, int& n0) {
        if (p.b_res) {
            bz = 0; m0 = tile * TC_BM; n0 = n_fixed;
        } else {
            bz = tile / tiles_per_batch;
            const int t = tile - bz * tiles_per_batch;
            m0 = (t / p.n_tiles) * TC_BM; n0 = (t % p.n_tiles) * BN;
        }
    };
    const int vec_cols = p.n_tiles * BN;

    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], TC_EPI_WARPS);
        }
        mbar_init(&b_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&mapA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&mapB) : "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (!p.b_is_const) asm volatile("griddepcontrol.wait;" ::: "memory");     // bias / dw vectors may be activations
    if (p.vec_in_smem && warp >= 2) {
        for (int i = threadIdx.x - 64; i < vec_cols; i += TC_EPI_WARPS * 32) {
            const bool ok = i < p.N;
            s_vec[i] = (ok && (p.epi & VFML_EPI_BIAS)) ? p.bias[i] : 0.0f;
            s_vec[vec_cols + i] = (ok && (p.epi & VFML_EPI_DW1)) ? p.dw_w[i] : 0.0f;
            s_vec[2 * vec_cols + i] = (ok && (p.epi & VFML_EPI_DW1)) ? p.dw_b[i] : 0.0f;
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;

    // Everything above (barriers, TMEM, bias / dw vectors) and the resident weight tile below depend on constants only;
    // activations written by the preceding kernels are touched after griddepcontrol.wait.
    if (warp == 0 && p.b_res && p.b_is_const && t_first < t_end && elect_one()) {
        mbar_expect_tx(&b_full, (uint32_t)nkb * b_bytes);
        for (int kb = 0; kb < nkb; ++kb) tma_load_3d(&mapB, &b_full, sB + (size_t)kb * b_stride, kb * TC_BK, n_fixed, 0);
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");

    if (warp == 0) {
        // ring position kept incrementally: an integer division by the runtime stage count on this issue path costs
        // more than the k-block it feeds (scripts/micro/mma_loop.cu)
        int s = 0;
        uint32_t ph = 0;
        const bool tr = p.trace != nullptr;
        long long t_wait = 0, t0 = tr ? clock64() : 0;
        if (p.b_res && !p.b_is_const && t_first < t_end && elect_one()) {
            mbar_expect_tx(&b_full, (uint32_t)nkb * b_bytes);
            for (int kb = 0; kb < nkb; ++kb) tma_load_3d(&mapB, &b_full, sB + (size_t)kb * b_stride, kb * TC_BK, n_fixed, 0);
        }
        __syncwarp();
        const uint32_t tx_bytes = p.b_res ? a_bytes : a_bytes + b_bytes;
        for (int tile = t_first; tile < t_end; tile += t_step) {
            int bz, m0, n0;
            tile_coords(tile, bz, m0, n0);
            for (int kb = 0; kb < nkb; ++kb) {
                const long long w0 = tr ? clock64() : 0;
                mbar_wait(&empty_bar[s], ph ^ 1u);
                if (tr) t_wait += clock64() - w0;
                if (elect_one()) {
#ifdef VFML_EXP_NOTMA
                    mbar_arrive(&full_bar[s]);
#else
                    mbar_expect_tx(&full_bar[s], tx_bytes);
                    tma_load_3d(&mapA, &full_bar[s], sA + (size_t)s * a_bytes, kb * TC_BK, m0, bz);
                    if (!p.b_res) tma_load_3d(&mapB, &full_bar[s], sB + (size_t)s * b_stride, kb * TC_BK, n0, bz);
#endif
                }
                __syncwarp();
                if (++s == stages) { s = 0; ph ^= 1u; }
            }
        }
        if (tr && lane == 0) { p.trace[blockIdx.x * 16 + 0] = t_wait; p.trace[blockIdx.x * 16 + 1] = clock64() - t0; }
    } else if (warp == 1) {
        // instruction descriptor: D fp32, A/B fp16, both K-major, N = BN, M = 128
        const uint32_t idesc = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
        const bool tr = p.trace != nullptr;
        long long t_we = 0, t_wf = 0, t0 = tr ? clock64() : 0;
        if (p.b_res && t_first < t_end) mbar_wait(&b_full, 0);
        // ring position, accumulator buffer and the operand descriptors are kept incrementally (no division, no
        // descriptor rebuild per k-block): the descriptor's low 14 bits are the smem address >> 4, so stepping a
        // stage is one 64-bit add
        int s = 0;
        uint32_t ph = 0, buf = 0, aph = 0;
        const uint64_t adesc0 = make_sw128_desc(smem_u32(sA)), bdesc0 = make_sw128_desc(smem_u32(sB));
        const uint64_t a_step = (uint64_t)(a_bytes >> 4), b_step = (uint64_t)(b_stride >> 4);
        uint64_t adesc = adesc0, bdesc_s = bdesc0;
        for (int tile = t_first; tile < t_end; tile += t_step) {
            const long long w0 = tr ? clock64() : 0;
            mbar_wait(&acc_empty[buf], aph ^ 1u);          // epilogue has drained this accumulator
            if (tr) t_we += clock64() - w0;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t tmem_d = tmem_base + buf * 256u;
            uint64_t bdesc_k = bdesc0;
            for (int kb = 0; kb < nkb; ++kb) {
                const long long w1 = tr ? clock64() : 0;
                mbar_wait(&full_bar[s], ph);
                if (tr) t_wf += clock64() - w1;
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint64_t bdesc = p.b_res ? bdesc_k : bdesc_s;
                if (elect_one()) {
#ifdef VFML_EXP_NOMMA
                    if (p.alpha == 12345.0f) umma_f16(tmem_d, adesc, bdesc, idesc, 0u);
                    mbar_arrive(&empty_bar[s]);
#else
                    umma_f16(tmem_d, adesc, bdesc, idesc, kb ? 1u : 0u);
#pragma unroll
                    for (int k = 1; k < TC_BK / 16; ++k)   // +32 bytes per K=16 step inside the 128-byte swizzle atom
                        umma_f16(tmem_d, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, 1u);
                    umma_commit(&empty_bar[s]);
#endif
                }
                __syncwarp();
                bdesc_k += b_step;
                adesc += a_step; bdesc_s += b_step;
                if (++s == stages) { s = 0; ph ^= 1u; adesc = adesc0; bdesc_s = bdesc0; }
            }
            if (elect_one()) {
#ifdef VFML_EXP_NOMMA
                mbar_arrive(&acc_full[buf]);
#else
                umma_commit(&acc_full[buf]);
#endif
            }
            __syncwarp();
            buf ^= 1u;
            aph ^= buf ^ 1u;       // flips each time buf wraps back to 0
        }
        if (tr && lane == 0) { p.trace[blockIdx.x * 16 + 2] = t_we; p.trace[blockIdx.x * 16 + 3] = t_wf; p.trace[blockIdx.x * 16 + 4] = clock64() - t0; }
    } else {
#if VFML_TC_EPI_V1
        long long e_wait = 0, e_t0 = clock64();
        const int q = warp & 3;                   // TMEM lane quarter this warp may access
        const int hsel = (warp - 2) >> 2;         // which CONTIGUOUS share of the tile's columns (contiguous: a thread's
                                                  // consecutive 32-byte stores complete whole 128-byte lines of its row)
        constexpr bool SP = SPEC >= 0;
        const uint32_t epi = SP ? (uint32_t)SPEC : p.epi;
        const bool has_bias = epi & VFML_EPI_BIAS, has_res = epi & VFML_EPI_RES, has_dw = epi & VFML_EPI_DW1;
        const bool do_gelu = epi & VFML_EPI_GELU, do_relu = epi & VFML_EPI_RELU;
        const bool out_f32 = SP ? false : (bool)p.out_f32, vec_smem = SP ? true : (bool)p.vec_in_smem;
        const float2 alpha2 = splat(p.alpha);
        uint32_t tl = 0;
        for (int tile = t_first; tile < t_end; tile += t_step, ++tl) {
            int bz, m0, n0;
            tile_coords(tile, bz, m0, n0);
            const uint32_t buf = tl & 1u, aph = (tl >> 1) & 1u;
            const int m = m0 + q * 32 + lane;
            const bool row_ok = m < p.M;
            const __half* rrow = p.res ? p.res + bz * p.strideR + (int64_t)m * p.ldr : nullptr;
            __half* d16 = reinterpret_cast<__half*>(p.D) + bz * p.strideD + (int64_t)m * p.ldd;
            float* d32 = reinterpret_cast<float*>(p.D) + bz * p.strideD + (int64_t)m * p.ldd;
            // software pipeline over this warp's 16-column chunks: the TMEM load and the residual load of
            // chunk i+1 are in flight while chunk i goes through the epilogue math and is stored
            const bool res_vec = has_res && row_ok && ((reinterpret_cast<uintptr_t>(rrow) & 15) == 0) && ((n0 & 7) == 0);
            const bool res_256 = res_vec && ((reinterpret_cast<uintptr_t>(rrow + n0) & 31) == 0);   // chunks start at multiples of 16 columns
            const uint32_t trow = tmem_base + buf * 256u + ((uint32_t)(q * 32) << 16);
            uint32_t tc[16], tn[16];
            uint4 rc0 = make_uint4(0, 0, 0, 0), rc1 = rc0, rn0 = rc0, rn1 = rc0;
            constexpr int NCS = TC_EPI_WARPS / 4;
            const int cols_here = min(BN, p.N - n0);                       // valid columns of this tile
            const int share = ((cols_here + 15) / 16 + NCS - 1) / NCS * 16;   // columns per warp share, multiple of 16
            int c = hsel * share;
#ifdef VFML_EXP_NOEPI
            const int c_end = p.alpha == 12345.0f ? min(c + share, BN) : 0;
#else
            const int c_end = min(c + share, BN);
#endif
            // the first chunk's residual does not depend on the MMA: it is fetched while this warp waits for the accumulator
            if (c < c_end && n0 + c < p.N && res_vec && n0 + c + 16 <= p.N) {
                if (res_256) ld_global_256(rrow + n0 + c, rc0, rc1);
                else {
                    rc0 = *reinterpret_cast<const uint4*>(rrow + n0 + c);
                    rc1 = *reinterpret_cast<const uint4*>(rrow + n0 + c + 8);
                }
            }
            const long long ew0 = clock64();
            mbar_wait(&acc_full[buf], aph);
            e_wait += clock64() - ew0;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (c < c_end && n0 + c < p.N) {
                tmem_ld16_issue(trow + (uint32_t)c, tc);
                tmem_ld_wait(tc);
            }
            for (; c < c_end; c += 16) {
                const int n = n0 + c;
                if (n >= p.N) break;                  // warp-uniform
                const int cn = c + 16;
                const bool has_next = cn < c_end && n0 + cn < p.N;
                if (has_next) {
                    tmem_ld16_issue(trow + (uint32_t)cn, tn);
                    if (res_vec && n0 + cn + 16 <= p.N) {
                        if (res_256) ld_global_256(rrow + n0 + cn, rn0, rn1);
                        else {
                            rn0 = *reinterpret_cast<const uint4*>(rrow + n0 + cn);
                            rn1 = *reinterpret_cast<const uint4*>(rrow + n0 + cn + 8);
                        }
                    }
                }
                if (row_ok) {
                    const bool full = n + 16 <= p.N;
                    float2 x[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        x[i] = make_float2(__uint_as_float(tc[2 * i]), __uint_as_float(tc[2 * i + 1]));
                        if (!SP) x[i] = mul2(x[i], alpha2);
                    }
                    if (has_bias) {
                        if (vec_smem) {
                            const float4* b4 = reinterpret_cast<const float4*>(s_vec + n);
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const float4 b = b4[i];
                                x[2 * i] = add2(x[2 * i], make_float2(b.x, b.y));
                                x[2 * i + 1] = add2(x[2 * i + 1], make_float2(b.z, b.w));
                            }
                        } else {
#pragma unroll
                            for (int i = 0; i < 8; ++i)
                                x[i] = add2(x[i], make_float2(__ldg(p.bias + min(n + 2 * i, p.N - 1)), __ldg(p.bias + min(n + 2 * i + 1, p.N - 1))));
                        }
                    }
                    if (has_res) {
                        if (res_vec && full) {
                            const __half2* h0 = reinterpret_cast<const __half2*>(&rc0);
                            const __half2* h1 = reinterpret_cast<const __half2*>(&rc1);
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                x[i] = add2(x[i], __half22float2(h0[i]));
                                x[4 + i] = add2(x[4 + i], __half22float2(h1[i]));
                            }
                        } else {
#pragma unroll
                            for (int i = 0; i < 8; ++i)
                                x[i] = add2(x[i], make_float2(n + 2 * i < p.N ? __half2float(rrow[n + 2 * i]) : 0.0f,
                                                              n + 2 * i + 1 < p.N ? __half2float(rrow[n + 2 * i + 1]) : 0.0f));
                        }
                    }
                    if (do_gelu) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) x[i] = gelu_fast2(x[i]);
                    }
                    if (do_relu) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) x[i] = make_float2(fmaxf(x[i].x, 0.0f), fmaxf(x[i].y, 0.0f));
                    }
                    if (has_dw) {   // k = 1 depthwise stage: gelu(v + (w * v + b))
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            float2 w2, b2;
                            if (vec_smem) {
                                w2 = *reinterpret_cast<const float2*>(s_vec + vec_cols + n + 2 * i);
                                b2 = *reinterpret_cast<const float2*>(s_vec + 2 * vec_cols + n + 2 * i);
                            } else {
                                const int n1 = min(n + 2 * i, p.N - 1), n2 = min(n + 2 * i + 1, p.N - 1);
                                w2 = make_float2(__ldg(p.dw_w + n1), __ldg(p.dw_w + n2));
                                b2 = make_float2(__ldg(p.dw_b + n1), __ldg(p.dw_b + n2));
                            }
                            x[i] = gelu_fast2(add2(x[i], fma2(w2, x[i], b2)));
                        }
                    }
#ifdef VFML_EXP_NOSTORE
                    if (p.alpha == 12345.0f)
#endif
                    if (out_f32) {
                        if (full && ((reinterpret_cast<uintptr_t>(d32 + n) & 15) == 0)) {
#pragma unroll
                            for (int i = 0; i < 4; ++i)
                                *reinterpret_cast<float4*>(d32 + n + 4 * i) = make_float4(x[2 * i].x, x[2 * i].y, x[2 * i + 1].x, x[2 * i + 1].y);
                        } else {
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                if (n + 2 * i < p.N) d32[n + 2 * i] = x[i].x;
                                if (n + 2 * i + 1 < p.N) d32[n + 2 * i + 1] = x[i].y;
                            }
                        }
                    } else {
                        if (full && ((reinterpret_cast<uintptr_t>(d16 + n) & 15) == 0)) {
                            uint4 o0, o1;
                            __half2* h0 = reinterpret_cast<__half2*>(&o0);
                            __half2* h1 = reinterpret_cast<__half2*>(&o1);
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                h0[i] = __float22half2_rn(x[i]);
                                h1[i] = __float22half2_rn(x[4 + i]);
                            }
                            // a row's 16 columns are one 32-byte sector when aligned: one 256-bit store = one whole-sector
                            // write request to L2 instead of two half-sector ones (the epilogue's limiter: each lane of a
                            // store instruction addresses a different row, so every request is a separate L2 transaction)
                            if ((reinterpret_cast<uintptr_t>(d16 + n) & 31) == 0) st_global_256(d16 + n, o0, o1);
                            else {
                                *reinterpret_cast<uint4*>(d16 + n) = o0;
                                *reinterpret_cast<uint4*>(d16 + n + 8) = o1;
                            }
                        } else {
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                if (n + 2 * i < p.N) d16[n + 2 * i] = __float2half_rn(x[i].x);
                                if (n + 2 * i + 1 < p.N) d16[n + 2 * i + 1] = __float2half_rn(x[i].y);
                            }
                        }
                    }
                }
                if (has_next) {
                    tmem_ld_wait(tn);
#pragma unroll
                    for (int i = 0; i < 16; ++i) tc[i] = tn[i];
                    rc0 = rn0;
                    rc1 = rn1;
                }
            }
            // this warp is done reading the accumulator: hand it back to the MMA warp
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[buf]);
        }
        if (p.trace && warp == 2 && lane == 0) { p.trace[blockIdx.x * 16 + 5] = e_wait; p.trace[blockIdx.x * 16 + 6] = clock64() - e_t0; }
#else
        // ---- epilogue, row-segment version.  Each of this warp's accesses to global memory covers 8 rows x 128 contiguous
        // bytes: the accumulator is read in the 16x256b shape (a quad of lanes holds one row segment), the fused math runs
        // in that fragment layout, and two 4x4 quad transposes per row turn the fp16 result (and the residual on its way in)
        // into 32 contiguous bytes per lane.  The lane-per-row version (VFML_TC_EPI_V1) issued one 32-byte sector per lane
        // and row -- 32 L2 requests per instruction -- and the in-kernel trace showed the MMA warp waiting for accumulators
        // 60 % of the time on the residual layers (profiles/r2_gemm_trace.txt).
        long long e_wait = 0, e_t0 = clock64();
        long long e_ld = 0, e_res = 0, e_math = 0, e_st = 0;       // VFML_GEMM_TRACE: where a unit's time goes (warp 2)
        const bool etr = p.trace != nullptr;
        const int q4 = warp & 3;                  // TMEM lane quarter this warp may access
        const int cg = (warp - 2) >> 2;           // units (64-column chunk, 16-row half) are dealt round-robin over the
        constexpr int NCS = TC_EPI_WARPS / 4;     // NCS warps of a quarter
        constexpr bool SP = SPEC >= 0;
        const uint32_t epi = SP ? (uint32_t)SPEC : p.epi;
        const bool has_bias = epi & VFML_EPI_BIAS, has_res = epi & VFML_EPI_RES, has_dw = epi & VFML_EPI_DW1;
        const bool do_gelu = epi & VFML_EPI_GELU, do_relu = epi & VFML_EPI_RELU;
        const bool out_f32 = SP ? false : (bool)p.out_f32, vec_smem = SP ? true : (bool)p.vec_in_smem;
        const float2 alpha2 = splat(p.alpha);
        const int qd = lane & 3, rl = lane >> 2;
        uint32_t tl = 0;
        for (int tile = t_first; tile < t_end; tile += t_step, ++tl) {
            int bz, m0, n0;
            tile_coords(tile, bz, m0, n0);
            const uint32_t buf = tl & 1u, aph = (tl >> 1) & 1u;
            const int cols_here = min(BN, p.N - n0);
            const int n_units = ((cols_here + 63) >> 6) * 2;
            const int n_lim = n0 + cols_here;          // a 64-column chunk may reach past the tile's last column: never touch those
            const __half* rbase = p.res ? p.res + bz * p.strideR : nullptr;
            __half* d16b = reinterpret_cast<__half*>(p.D) + bz * p.strideD;
            float* d32b = reinterpret_cast<float*>(p.D) + bz * p.strideD;
            // vector path: 16-byte aligned rows (warp-uniform: the quad transposes are collective)
            const bool d_vec = !out_f32 && ((reinterpret_cast<uintptr_t>(d16b) & 15) == 0) && ((p.ldd & 7) == 0) && ((n0 & 7) == 0);
            const bool r_vec = has_res && ((reinterpret_cast<uintptr_t>(rbase) & 15) == 0) && ((p.ldr & 7) == 0) && ((n0 & 7) == 0);
            const long long ew0 = clock64();
            mbar_wait(&acc_full[buf], aph);
            e_wait += clock64() - ew0;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            // residual rows of a unit, row-contiguous words (lane qd: columns c + 16 qd .. + 15 of rows rA, rA + 8)
            auto load_res = [&](int u, uint32_t (&ra)[8], uint32_t (&rb)[8]) {
                const int c = (u >> 1) << 6;
                const int mA = m0 + q4 * 32 + (u & 1) * 16 + rl;
                const int n = n0 + c + 16 * qd;
#pragma unroll
                for (int s2 = 0; s2 < 2; ++s2) {
                    uint32_t (&r)[8] = s2 ? rb : ra;
                    const int m = mA + 8 * s2;
                    const __half* rp = rbase + (int64_t)m * p.ldr + n;
                    if (m < p.M && n + 16 <= n_lim && r_vec) {
                        const uint4 v0 = *reinterpret_cast<const uint4*>(rp), v1 = *reinterpret_cast<const uint4*>(rp + 8);
                        r[0] = v0.x; r[1] = v0.y; r[2] = v0.z; r[3] = v0.w; r[4] = v1.x; r[5] = v1.y; r[6] = v1.z; r[7] = v1.w;
                    } else {
#pragma unroll
                        for (int k = 0; k < 8; ++k) {
                            const __half lo = (m < p.M && n + 2 * k < n_lim) ? rp[2 * k] : __float2half_rn(0.0f);
                            const __half hi = (m < p.M && n + 2 * k + 1 < n_lim) ? rp[2 * k + 1] : __float2half_rn(0.0f);
                            r[k] = (uint32_t)__half_as_ushort(lo) | ((uint32_t)__half_as_ushort(hi) << 16);
                        }
                    }
                }
            };
            uint32_t ra[8], rb[8];
            int u = cg;
            if (has_res && u < n_units) load_res(u, ra, rb);
            for (; u < n_units; u += NCS) {
                const int c = (u >> 1) << 6, hf = u & 1;
                uint32_t acc[32];
                long long tk0 = etr ? clock64() : 0;
                tmem_ld_16x256b_x8_issue(tmem_base + buf * 256u + (uint32_t)c + ((uint32_t)(q4 * 32 + hf * 16) << 16), acc);
                uint32_t fa[8], fb[8];                 // residual in fragment layout
                if (has_res) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) { fa[k] = ra[k]; fb[k] = rb[k]; }
                    quad_rows_to_frag(fa, qd);
                    quad_rows_to_frag(fb, qd);
                    if (u + NCS < n_units) load_res(u + NCS, ra, rb);      // next unit's residual flies during this unit's math
                }
                long long tk1 = etr ? clock64() : 0;
                tmem_ld_wait32(acc);
                long long tk2 = etr ? clock64() : 0;
                e_res += tk1 - tk0; e_ld += tk2 - tk1;
                const int mA = m0 + q4 * 32 + hf * 16 + rl;
                uint32_t oa[8], ob[8];                 // fp16 results, fragment layout
#pragma unroll
                for (int g = 0; g < 8; ++g) {
                    const int n = n0 + c + 8 * g + 2 * qd;           // this lane's two columns of group g
                    const int nv = min(n, vec_cols - 2);             // columns past the tile's last one are computed, never stored
                    float2 xa = make_float2(__uint_as_float(acc[4 * g]), __uint_as_float(acc[4 * g + 1]));
                    float2 xb = make_float2(__uint_as_float(acc[4 * g + 2]), __uint_as_float(acc[4 * g + 3]));
                    if (!SP) { xa = mul2(xa, alpha2); xb = mul2(xb, alpha2); }
                    if (has_bias) {
                        float2 b2;
                        if (vec_smem) b2 = *reinterpret_cast<const float2*>(s_vec + nv);         // indexed by absolute column
                        else b2 = make_float2(__ldg(p.bias + min(n, p.N - 1)), __ldg(p.bias + min(n + 1, p.N - 1)));
                        xa = add2(xa, b2); xb = add2(xb, b2);
                    }
                    if (has_res) {
                        xa = add2(xa, __half22float2(*reinterpret_cast<const __half2*>(&fa[g])));
                        xb = add2(xb, __half22float2(*reinterpret_cast<const __half2*>(&fb[g])));
                    }
                    if (do_gelu) { xa = gelu_fast2(xa); xb = gelu_fast2(xb); }
                    if (do_relu) {
                        xa = make_float2(fmaxf(xa.x, 0.0f), fmaxf(xa.y, 0.0f));
                        xb = make_float2(fmaxf(xb.x, 0.0f), fmaxf(xb.y, 0.0f));
                    }
                    if (has_dw) {   // k = 1 depthwise stage: gelu(v + (w * v + b))
                        float2 w2, b2;
                        if (vec_smem) {
                            w2 = *reinterpret_cast<const float2*>(s_vec + vec_cols + nv);
                            b2 = *reinterpret_cast<const float2*>(s_vec + 2 * vec_cols + nv);
                        } else {
                            const int n1 = min(n, p.N - 1), n2 = min(n + 1, p.N - 1);
                            w2 = make_float2(__ldg(p.dw_w + n1), __ldg(p.dw_w + n2));
                            b2 = make_float2(__ldg(p.dw_b + n1), __ldg(p.dw_b + n2));
                        }
                        xa = gelu_fast2(add2(xa, fma2(w2, xa, b2)));
                        xb = gelu_fast2(add2(xb, fma2(w2, xb, b2)));
                    }
                    if (out_f32) {          // fp32 outputs (logits, partial sums, delta, mask) leave in fragment layout: 8 B per lane
                        float* da = d32b + (int64_t)mA * p.ldd + n;
                        float* db = da + 8 * p.ldd;
                        const bool al = ((reinterpret_cast<uintptr_t>(d32b) & 7) == 0) && ((p.ldd & 1) == 0) && ((n & 1) == 0);
                        if (mA < p.M) {
                            if (n + 1 < n_lim && al) *reinterpret_cast<float2*>(da) = xa;
                            else { if (n < n_lim) da[0] = xa.x; if (n + 1 < n_lim) da[1] = xa.y; }
                        }
                        if (mA + 8 < p.M) {
                            if (n + 1 < n_lim && al) *reinterpret_cast<float2*>(db) = xb;
                            else { if (n < n_lim) db[0] = xb.x; if (n + 1 < n_lim) db[1] = xb.y; }
                        }
                    } else {
                        const __half2 ha = __float22half2_rn(xa), hb = __float22half2_rn(xb);
                        oa[g] = *reinterpret_cast<const uint32_t*>(&ha);
                        ob[g] = *reinterpret_cast<const uint32_t*>(&hb);
                    }
                }
                long long tk3 = etr ? clock64() : 0;
                e_math += tk3 - tk2;
                if (!out_f32) {
                    quad_frag_to_rows(oa, qd);
                    quad_frag_to_rows(ob, qd);
                    const int n = n0 + c + 16 * qd;
#pragma unroll
                    for (int s2 = 0; s2 < 2; ++s2) {
                        const uint32_t (&o)[8] = s2 ? ob : oa;
                        const int m = mA + 8 * s2;
                        if (m >= p.M) continue;
                        __half* dp = d16b + (int64_t)m * p.ldd + n;
                        if (n + 16 <= n_lim && d_vec) {
                            const uint4 v0 = make_uint4(o[0], o[1], o[2], o[3]), v1 = make_uint4(o[4], o[5], o[6], o[7]);
                            if ((reinterpret_cast<uintptr_t>(dp) & 31) == 0) st_global_256(dp, v0, v1);
                            else {
                                *reinterpret_cast<uint4*>(dp) = v0;
                                *reinterpret_cast<uint4*>(dp + 8) = v1;
                            }
                        } else {
#pragma unroll
                            for (int k = 0; k < 8; ++k) {
                                if (n + 2 * k < n_lim) dp[2 * k] = __ushort_as_half((unsigned short)(o[k] & 0xffffu));
                                if (n + 2 * k + 1 < n_lim) dp[2 * k + 1] = __ushort_as_half((unsigned short)(o[k] >> 16));
                            }
                        }
                    }
                }
                if (etr) e_st += clock64() - tk3;
            }
            // this warp is done reading the accumulator: hand it back to the MMA warp
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[buf]);
        }
        if (p.trace && warp == 2 && lane == 0) {
            p.trace[blockIdx.x * 16 + 5] = e_wait; p.trace[blockIdx.x * 16 + 6] = clock64() - e_t0;
            p.trace[blockIdx.x * 16 + 8] = e_ld; p.trace[blockIdx.x * 16 + 9] = e_res; p.trace[blockIdx.x * 16 + 10] = e_math; p.trace[blockIdx.x * 16 + 11] = e_st;
        }
#endif
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// ------------------------------------------------------------------------------------------ host
static int make_map(CUtensorMap* map, const void* base, int64_t inner, int64_t rows, int64_t batch, int64_t ld,
                    int64_t stride, int box_rows) {
    cuuint64_t dims[3] = {(cuuint64_t)inner, (cuuint64_t)rows, (cuuint64_t)batch};
    cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)(batch > 1 ? stride : ld * rows) * 2};
    cuuint32_t box[3] = {(cuuint32_t)TC_BK, (cuuint32_t)box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    // libcuda is resolved at run time through the runtime (no link-time dependency: the library must
    // load on a CPU-only build box for the symbol check)
    typedef CUresult (*encode_fn_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static encode_fn_t encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
            set_error("gemm_tc: cuTensorMapEncodeTiled is not available from the driver (%s)", cudaGetErrorString(e));
            return VFML_ERR_CUDA;
        }
        encode = reinterpret_cast<encode_fn_t>(fn);
    }
    CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("gemm_tc: cuTensorMapEncodeTiled failed: CUresult %d (inner=%lld rows=%lld batch=%lld ld=%lld)", (int)r,
                  (long long)inner, (long long)rows, (long long)batch, (long long)ld);
        return VFML_ERR_CUDA;
    }
    return VFML_OK;
}

// can this descriptor go through the TMA path? (16-byte aligned bases and pitches, fp16 operands)
bool gemm_tc_supported(const vfml_gemm_desc& d) {
    if (d.in_dtype != VFML_F16) return false;
    if (d.epi & VFML_EPI_ADDTO) return false;
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    if (!al16(d.A) || !al16(d.B)) return false;
    if ((d.lda % 8) || (d.ldb % 8)) return false;
    if (d.batch > 1 && ((d.strideA % 8) || (d.strideB % 8))) return false;
    return d.K >= 1;
}

static int pick_bn(int N) {
    if (N <= 256) return (N + 15) / 16 * 16;
    const int tiles = (N + 255) / 256;
    return ((N + tiles - 1) / tiles + 15) / 16 * 16;
}

// kernel instances: [0] generic, the rest specialised on the epilogue flag set
struct TcKernel { int spec; const void* fn; };
#define TC_K(S) {S, reinterpret_cast<const void*>(&gemm_tc_kernel<S>)}
static const TcKernel tc_kernels[] = {
    TC_K(-1), TC_K(0), TC_K(VFML_EPI_BIAS), TC_K(VFML_EPI_BIAS | VFML_EPI_GELU), TC_K(VFML_EPI_BIAS | VFML_EPI_RES),
    TC_K(VFML_EPI_BIAS | VFML_EPI_RES | VFML_EPI_GELU), TC_K(VFML_EPI_BIAS | VFML_EPI_RES | VFML_EPI_GELU | VFML_EPI_DW1)};
#undef TC_K
constexpr int TC_NUM_KERNELS = sizeof(tc_kernels) / sizeof(tc_kernels[0]);

int gemm_tc(const vfml_gemm_desc& d, cudaStream_t st) {
    const int num_sms = device_sms();
    static PerDevice<bool> attr_done;
    if (!attr_done.get()) {
        for (int i = 0; i < TC_NUM_KERNELS; ++i)
            VFML_CUDA(cudaFuncSetAttribute(tc_kernels[i].fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 512));
        attr_done.get() = true;
    }
    const size_t SMEM_MAX = 227 * 1024 - 512;
    const size_t a_bytes = TC_BM * TC_BK * 2;
    const int nkb = cdiv(d.K, TC_BK), m_tiles = cdiv(d.M, TC_BM);
    const bool needs_vec = d.epi & (VFML_EPI_BIAS | VFML_EPI_DW1);
    auto bstride = [](int bn) { return ((size_t)bn * TC_BK * 2 + 1023) & ~(size_t)1023; };
    // (1) streaming mode: B tiles ride the ring with A
    int BN = pick_bn(d.N);
    int n_tiles = cdiv(d.N, BN);
    double best_traffic = (double)n_tiles * d.M * d.K + (double)m_tiles * n_tiles * BN * d.K;      // elements through L2 -> SM
    bool b_res = false;
    // (2) weight-stationary mode (measured: the streaming kernel is bound by L2 -> SM operand traffic, 190 MB for a
    // 42240 x 486 x 324 layer whose operands are 28 MB): each CTA keeps one B tile (all of K) resident and streams
    // only A.  Pick the fewest n-tiles whose B tile fits next to >= 3 A stages, and use it when it moves fewer bytes.
    static const bool no_res = getenv("VFML_GEMM_NO_RESIDENT") != nullptr;
    if (d.batch == 1 && !no_res) {
        for (int nt = 1; nt <= 16; ++nt) {
            const int bn = ((d.N + nt - 1) / nt + 15) / 16 * 16;
            if (bn > 256) continue;
            const size_t vecb = needs_vec && nt * bn <= TC_VEC_MAX ? (size_t)3 * nt * bn * sizeof(float) : 0;
            if (1024 + nkb * bstride(bn) + 3 * a_bytes + vecb > SMEM_MAX) continue;
            const int grid_r = std::max(1, num_sms / nt) * nt;
            const double traffic = (double)nt * d.M * d.K + (double)grid_r * bn * d.K;
            if (m_tiles * nt >= num_sms && traffic < 0.8 * best_traffic) {
                b_res = true; BN = bn; n_tiles = nt; best_traffic = traffic;
            }
            break;
        }
        // (3) wave quantisation (VFML_GEMM_WAVE=<c0>, default 32): each CTA walks ceil(m_tiles / (SMs / nt)) row tiles of its
        // column tile; 165 row tiles on 148 SMs cost two rounds with nt = 1 but three half-width rounds with nt = 2.
        // Cost model: rounds x (c0 + BN) column-equivalents, c0 = fixed cost of a tile (fill, drain, barrier hand-offs).
        static const int wave_c0 = getenv("VFML_GEMM_WAVE") ? atoi(getenv("VFML_GEMM_WAVE")) : 32;   // measured: 29.68 -> 29.36 ms per window; 0 disables
        if (wave_c0 > 0) {
            auto cost = [&](int nt, int bn) {
                const int rounds = b_res ? cdiv(m_tiles, std::max(1, num_sms / nt)) : cdiv((int64_t)m_tiles * nt, num_sms);
                return (double)rounds * (wave_c0 + bn);
            };
            double best = cost(n_tiles, BN);
            const int nt0 = n_tiles;
            for (int nt = nt0 + 1; nt <= nt0 + 3; ++nt) {
                const int bn = ((d.N + nt - 1) / nt + 15) / 16 * 16;
                if (bn < 48 || (nt - 1) * bn >= d.N) continue;
                const double c = cost(nt, bn);
                if (c < 0.95 * best) { best = c; BN = bn; n_tiles = nt; }
            }
        }
    }
    CUtensorMap mapA, mapB;
    int rc = make_map(&mapA, d.A, d.K, d.M, d.batch, d.lda, d.strideA, TC_BM);
    if (rc) return rc;
    rc = make_map(&mapB, d.B, d.K, d.N, d.batch, d.ldb, d.strideB, BN);
    if (rc) return rc;
    TcParams p{};
    p.m_tiles = m_tiles;
    p.n_tiles = n_tiles;
    p.b_res = b_res;
    p.vec_in_smem = needs_vec && p.n_tiles * BN <= TC_VEC_MAX;
    const size_t vec_bytes = p.vec_in_smem ? (size_t)3 * p.n_tiles * BN * sizeof(float) : 0;
    const size_t b_stride = bstride(BN);
    int stages;
    size_t smem;
    if (b_res) {
        stages = (int)((SMEM_MAX - 1024 - vec_bytes - nkb * b_stride) / a_bytes);
        stages = std::max(2, std::min(stages, TC_MAX_STAGES));
        smem = 1024 + nkb * b_stride + stages * a_bytes + vec_bytes;
    } else {
        stages = (int)((200 * 1024 - 1024 - vec_bytes) / (a_bytes + b_stride));
        stages = std::max(2, std::min(stages, TC_MAX_STAGES));
        smem = 1024 + stages * (a_bytes + b_stride) + vec_bytes;
    }
    p.D = d.D; p.bias = d.bias; p.res = reinterpret_cast<const __half*>(d.res); p.dw_w = d.dw_w; p.dw_b = d.dw_b;
    p.ldd = d.ldd; p.ldr = d.ldr; p.strideD = d.strideD; p.strideR = d.strideR;
    p.M = d.M; p.N = d.N; p.K = d.K; p.batch = d.batch; p.BN = BN; p.stages = stages;
    p.alpha = d.alpha; p.epi = d.epi; p.out_f32 = d.out_dtype == VFML_F32;
    const int64_t num_tiles = (int64_t)p.m_tiles * p.n_tiles * d.batch;
    const int grid = b_res ? std::max(1, num_sms / n_tiles) * n_tiles : (int)std::min<int64_t>(num_tiles, num_sms);
    static const bool trace = getenv("VFML_GEMM_TRACE") != nullptr;
    constexpr int TRACE_MAX_CTAS = 1024;      // debug buffer: one 8-counter record per CTA of the persistent grid
    static PerDevice<long long*> trace_bufs;
    long long*& trace_buf = trace_bufs.get();
    if (trace) {
        VFML_REQUIRE(grid <= TRACE_MAX_CTAS, "gemm_tc: VFML_GEMM_TRACE supports at most %d CTAs", TRACE_MAX_CTAS);
        if (!trace_buf) VFML_CUDA(cudaMalloc(&trace_buf, TRACE_MAX_CTAS * 16 * sizeof(long long)));
        VFML_CUDA(cudaMemsetAsync(trace_buf, 0, (size_t)grid * 16 * sizeof(long long), st));
        p.trace = trace_buf;
    }
    // programmatic dependent launch: this kernel's prologue (and its resident weight tile) overlaps the tail of the
    // preceding kernel; it synchronises with griddepcontrol.wait before touching activations
    static const bool no_pdl = getenv("VFML_NO_PDL") != nullptr;
    p.b_is_const = (d.epi & VFML_EPI_WEIGHTS) != 0;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = no_pdl ? 0 : 1;
    const void* fn = tc_kernels[0].fn;
    static const bool no_spec = getenv("VFML_GEMM_NO_SPEC") != nullptr;
    if (!no_spec && !p.out_f32 && d.alpha == 1.0f && (!needs_vec || p.vec_in_smem))
        for (int i = 1; i < TC_NUM_KERNELS; ++i)
            if ((d.epi & ~VFML_EPI_WEIGHTS) == (uint32_t)tc_kernels[i].spec) fn = tc_kernels[i].fn;
    void* args[3] = {&mapA, &mapB, &p};
    cudaError_t le = cudaLaunchKernelExC(&cfg, fn, args);
    if (le != cudaSuccess) {
        set_error("gemm_tc: launch failed: %s", cudaGetErrorString(le));
        cudaGetLastError();
        return VFML_ERR_CUDA;
    }
    VFML_LAUNCH_CHECK("gemm_tc");
    if (trace) {      // debug only: synchronises
        static long long h[TRACE_MAX_CTAS * 16];
        VFML_CUDA(cudaMemcpyAsync(h, trace_buf, (size_t)grid * 16 * sizeof(long long), cudaMemcpyDeviceToHost, st));
        VFML_CUDA(cudaStreamSynchronize(st));
        double a[16] = {0};
        for (int c = 0; c < grid; ++c)
            for (int i = 0; i < 16; ++i) a[i] += (double)h[c * 16 + i] / grid;
        fprintf(stderr, "[gemm_tc trace] M=%d N=%d K=%d BN=%d stages=%d tiles=%lld epi=%u res=%d | producer wait_empty %.0f of %.0f | mma wait_acc_empty %.0f wait_full %.0f of %.0f | epilogue(warp2) wait_acc_full %.0f of %.0f cycles (avg per CTA) | epi phases: tmem_ld %.0f res %.0f math %.0f store %.0f\n",
                d.M, d.N, d.K, BN, stages, (long long)num_tiles, d.epi, (int)b_res, a[0], a[1], a[2], a[3], a[4], a[5], a[6], a[8], a[9], a[10], a[11]);
    }
    return VFML_OK;
}

}  // namespace vfml

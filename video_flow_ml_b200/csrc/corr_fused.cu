// All-pairs correlation volume with the pyramid built in the GEMM's epilogue (fp16 tier).
//
// Reference: CorrBlock.__init__ (VideoFlow/core/Networks/MOFNetStack/corr.py:59-74): corr = fmap1^T fmap2 / sqrt(C), then three
// avg_pool2d(2, 2) over the target image.  The unfused path (corr_tiled.cu) writes a row-major fp16 level-0 volume, reads it back
// and rewrites it as the tiled pyramid.  Here the product never exists in row-major form:
//
//   * tcgen05 GEMM, M = 128 source pixels (TMEM lanes), N = 256 target pixels = FOUR 8 x 8 BLOCKS of the target image (walked
//     down the columns of blocks: the tiles of a tile column are contiguous in the slab, so a lane's stores of one N tile
//     fall into two 256-byte runs of level 0 and one 128-byte run of level 1).  The B
//     operand of a tile is fetched by four TMA boxes {64 channels, 8 x, 8 y} per k-block from a 4-D view [pair][y][x][channel] of
//     fmap2's rows, so accumulator columns 64 g + 8 yy + xx are block g's texel (xx, yy); texels outside the image are zero-filled
//     by TMA, which is exactly the zero padding the tiled layout wants.
//   * An 8 x 8 block holds whole 4 x 4 tiles of level 0 (four of them) AND of level 1 (one, after 2 x 2 pooling), so an epilogue
//     thread (= one source pixel) turns 64 accumulator columns into five finished 32-byte tiles: scale, round to fp16 (the stored
//     level 0), pool the ROUNDED values in fp32 in the order of corr_tile_pyramid_kernel, round again.  Bit-identical to the
//     unfused path (tests/test_gpu_ops.py::test_corr_volume_fused_epilogue_is_bit_identical).
//   * Levels 2 and 3 (6 % of the bytes) need 16 x 16 / 32 x 32 target blocks: corr_tile_pyramid_kernel (corr_tiled.cu, first = 1) makes them
//     from the tiled level 1 in a small second pass.
//   * The A tile (128 source pixels, all of K = C <= 256) stays resident in shared memory; a CTA walks a chunk of the N tiles with
//     a ring of 5-6 k-block slots (32 KB each) for B and two TMEM accumulators (the MMAs of tile i + 1 run under the stores of tile i).
//
// Bound: the epilogue's stores -- 128 lanes x 5 tiles x 32 B per block, one sector per store instruction and lane
// (profiles/r2_store_probe_scaling.txt: 16 B/clk/SM for this pattern), not the tensor pipe (K <= 256).
#include "tcgen05.cuh"

namespace vfml {

constexpr int CF_BM = 128, CF_BN = 256, CF_BK = 64;
constexpr int CF_EPI_WARPS = 8;
constexpr int CF_THREADS = (2 + CF_EPI_WARPS) * 32;
constexpr int CF_A_BYTES = CF_BM * 128;          // one k-block of the resident source-pixel tile
constexpr int CF_B_BYTES = CF_BN * 128;          // one k-block of a target tile
constexpr int CF_BLK_BYTES = 64 * 128;           // one 8 x 8 target block of one k-block
constexpr int CF_MAX_STAGES = 6;          // ring slots of ONE k-block (32 KB) of a target tile

struct CfMaps {
    CUtensorMap A, B;
};

struct CfParams {
    uint8_t* lvl0;
    uint8_t* lvl1;
    int P, nkb, stages;
    int nby, n_blocks, n_tiles, tiles_per_cta, chunks, mt;
    int ntx0, nty0, ntx1, nty1, lw1, lh1;
    uint32_t slab0, slab1;
    float alpha;
};

__device__ __forceinline__ void tma_load_4d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}

// 32 TMEM lanes (thread t = lane t of this warp's quarter) x 32 consecutive fp32 columns
__device__ __forceinline__ void tmem_ld_32x32b_x32_issue(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}

__device__ __forceinline__ uint32_t pack_h2(float lo, float hi) {
    const __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&h);
}

// One 8 x 8 target block of one source pixel: v[half][(yy & 3) * 8 + xx] (half = yy >> 2) holds the accumulators.
// Writes the block's four level-0 tiles and its level-1 tile.
__device__ __forceinline__ void cf_store_block(const uint32_t (&ra)[32], const uint32_t (&rb)[32], const CfParams& p, uint8_t* row0,
                                               uint8_t* row1, int bxg, int by) {
    float v[2][32];
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        v[0][i] = __half2float(__float2half_rn(__uint_as_float(ra[i]) * p.alpha));
        v[1][i] = __half2float(__float2half_rn(__uint_as_float(rb[i]) * p.alpha));
    }
    // level 0: tile (tx, ty), element (xr, y4) at half index xr * 4 + y4; the two tiles of a tile column are adjacent in memory
#pragma unroll
    for (int txl = 0; txl < 2; ++txl) {
        const int tx = 2 * bxg + txl;
        if (tx < p.ntx0) {
#pragma unroll
            for (int tyl = 0; tyl < 2; ++tyl) {
                const int ty = 2 * by + tyl;
                if (ty < p.nty0) {
                    uint32_t w[8];
#pragma unroll
                    for (int wi = 0; wi < 8; ++wi) {
                        const int xr = wi >> 1, y4 = (wi & 1) * 2;
                        w[wi] = pack_h2(v[tyl][y4 * 8 + txl * 4 + xr], v[tyl][(y4 + 1) * 8 + txl * 4 + xr]);
                    }
                    st_global_256(row0 + (size_t)(tx * p.nty0 + ty) * 32, make_uint4(w[0], w[1], w[2], w[3]), make_uint4(w[4], w[5], w[6], w[7]));
                }
            }
        }
    }
    // level 1: avg_pool2d(2, 2) of the rounded level 0, summed as ((a + b) + c) + d like corr_tile_pyramid_kernel
    if (bxg < p.ntx1 && by < p.nty1) {
        uint32_t w[8];
#pragma unroll
        for (int wi = 0; wi < 8; ++wi) {
            const int x1 = wi >> 1;
            float q[2];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int y1 = (wi & 1) * 2 + j;
                const int yy = 2 * y1, hf = yy >> 2, yr = yy & 3;        // rows 2 y1 and 2 y1 + 1 are in the same half
                const float s = ((v[hf][yr * 8 + 2 * x1] + v[hf][yr * 8 + 2 * x1 + 1]) + v[hf][(yr + 1) * 8 + 2 * x1]) + v[hf][(yr + 1) * 8 + 2 * x1 + 1];
                q[j] = (bxg * 4 + x1 < p.lw1 && by * 4 + y1 < p.lh1) ? s * 0.25f : 0.0f;
            }
            w[wi] = pack_h2(q[0], q[1]);
        }
        st_global_256(row1 + (size_t)(bxg * p.nty1 + by) * 32, make_uint4(w[0], w[1], w[2], w[3]), make_uint4(w[4], w[5], w[6], w[7]));
    }
}

__global__ void __launch_bounds__(CF_THREADS, 1)
corr_fused_kernel(const __grid_constant__ CfMaps maps, const CfParams p) {
    extern __shared__ uint8_t cf_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(cf_smem_raw) + 1023) & ~(uintptr_t)1023);
    const int nkb = p.nkb, stages = p.stages;
    uint8_t* sA = smem;                                          // [nkb] resident source-pixel tile
    uint8_t* sB = sA + (size_t)nkb * CF_A_BYTES;                 // [stages] k-blocks of target tiles
    __shared__ uint64_t full_bar[CF_MAX_STAGES], empty_bar[CF_MAX_STAGES], acc_full[2], acc_empty[2], a_full;
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int chunk = (int)blockIdx.x % p.chunks, mrow = (int)blockIdx.x / p.chunks;
    const int pair = mrow / p.mt, m0 = (mrow - pair * p.mt) * CF_BM;
    const int t_begin = chunk * p.tiles_per_cta, t_end = min(t_begin + p.tiles_per_cta, p.n_tiles);

    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], CF_EPI_WARPS);
        }
        mbar_init(&a_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.A) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.B) : "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;

    if (warp == 0) {
        // ---- producer: the resident A tile once, then four 8 x 8 target blocks per k-block and tile
        if (elect_one()) {
            mbar_expect_tx(&a_full, (uint32_t)nkb * CF_A_BYTES);
            for (int kb = 0; kb < nkb; ++kb) tma_load_3d(&maps.A, &a_full, sA + (size_t)kb * CF_A_BYTES, kb * CF_BK, m0, pair);
        }
        __syncwarp();
        int s = 0;
        uint32_t ph = 0;
        for (int t = t_begin; t < t_end; ++t) {
            for (int kb = 0; kb < nkb; ++kb) {
                mbar_spin(&empty_bar[s], ph ^ 1u);
                if (elect_one()) {
                    mbar_expect_tx(&full_bar[s], (uint32_t)CF_B_BYTES);
                    for (int g = 0; g < 4; ++g) {
                        const int b = t * 4 + g, bxg = b / p.nby, by = b - bxg * p.nby;      // blocks past the last one lie right of the image: zeros
                        tma_load_4d(&maps.B, &full_bar[s], sB + (size_t)s * CF_B_BYTES + (size_t)g * CF_BLK_BYTES, kb * CF_BK, bxg * 8, by * 8, pair);
                    }
                }
                __syncwarp();
                if (++s == stages) { s = 0; ph ^= 1u; }
            }
        }
    } else if (warp == 1) {
        // ---- MMA issuer: K / 16 instructions of 128 x 256 x 16 per tile
        const uint32_t idesc = (1u << 4) | ((uint32_t)(CF_BN >> 3) << 17) | ((uint32_t)(CF_BM >> 4) << 24);
        const uint64_t adesc0 = make_sw128_desc(smem_u32(sA)), bdesc0 = make_sw128_desc(smem_u32(sB));
        if (t_begin < t_end) mbar_wait(&a_full, 0);
        int s = 0;
        uint32_t ph = 0, buf = 0, aph = 0;
        for (int t = t_begin; t < t_end; ++t) {
            mbar_spin(&acc_empty[buf], aph ^ 1u);
            const uint32_t tmem_d = tmem_base + buf * (uint32_t)CF_BN;
            for (int kb = 0; kb < nkb; ++kb) {
                mbar_spin(&full_bar[s], ph);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                if (elect_one()) {
                    const uint64_t ad = adesc0 + (uint64_t)((kb * CF_A_BYTES) >> 4), bd = bdesc0 + (uint64_t)((s * CF_B_BYTES) >> 4);
#pragma unroll
                    for (int k = 0; k < CF_BK / 16; ++k) umma_f16(tmem_d, ad + (uint64_t)(k * 2), bd + (uint64_t)(k * 2), idesc, (kb | k) ? 1u : 0u);
                    umma_commit(&empty_bar[s]);                       // the slot is free once these MMAs have read it
                    if (kb == nkb - 1) umma_commit(&acc_full[buf]);
                }
                __syncwarp();
                if (++s == stages) { s = 0; ph ^= 1u; }
            }
            if (++buf == 2u) { buf = 0; aph ^= 1u; }
        }
    } else {
        // ---- epilogue: warp = (TMEM lane quarter it may read, two of the tile's four blocks); lane = source pixel
        const int quarter = warp & 3, hsel = (warp - 2) >> 2;
        const int m = m0 + quarter * 32 + lane;
        const bool m_ok = m < p.P;
        uint8_t* row0 = p.lvl0 + ((size_t)pair * p.P + (m_ok ? m : 0)) * p.slab0;
        uint8_t* row1 = p.lvl1 + ((size_t)pair * p.P + (m_ok ? m : 0)) * p.slab1;
        uint32_t buf = 0, aph = 0;
        for (int t = t_begin; t < t_end; ++t) {
            mbar_wait(&acc_full[buf], aph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll 1
            for (int u = 0; u < 2; ++u) {
                const int g = hsel * 2 + u;
                const uint32_t taddr = tmem_base + buf * (uint32_t)CF_BN + (uint32_t)(g * 64) + ((uint32_t)(quarter * 32) << 16);
                uint32_t ra[32], rb[32];
                tmem_ld_32x32b_x32_issue(taddr, ra);
                tmem_ld_32x32b_x32_issue(taddr + 32, rb);
                tmem_ld_wait32(ra);
                tmem_ld_wait32(rb);
                if (u == 1) {                      // both blocks are in registers: the accumulator can be overwritten
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&acc_empty[buf]);
                }
                const int b = t * 4 + g, bxg = b / p.nby, by = b - bxg * p.nby;
                if (m_ok && b < p.n_blocks) cf_store_block(ra, rb, p, row0, row1, bxg, by);
            }
            if (++buf == 2u) { buf = 0; aph ^= 1u; }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

static int cf_map(CUtensorMap* map, const void* base, int rank, const cuuint64_t* dims, const cuuint64_t* strides, const cuuint32_t* box,
                  const char* what) {
    tensor_map_encode_fn encode = tensor_map_encoder();
    if (!encode) {
        set_error("corr_volume_fused: cuTensorMapEncodeTiled is not available from the driver");
        return VFML_ERR_CUDA;
    }
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, (cuuint32_t)rank, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("corr_volume_fused: cuTensorMapEncodeTiled(%s) failed: CUresult %d", what, (int)r);
        return VFML_ERR_CUDA;
    }
    return VFML_OK;
}

// VFML_CORR_FUSED=0 restores the unfused path (GEMM -> row-major level 0 -> corr_tile_pyramid_kernel); read at every call
bool corr_fused_supported(int C, int h, int w, int num_levels) {
    const char* e = getenv("VFML_CORR_FUSED");
    if (e && e[0] == '0') return false;
    return num_levels == 4 && C >= 64 && C <= 256 && (C % 64) == 0 && h >= 8 && w >= 8;
}

// r1, r2: fp16 rows [n_pairs * P, C] of the two feature maps; writes levels[0] and levels[1] (tiled layout of corr_tiled.cu)
int corr_volume_fused(const __half* r1, const __half* r2, int n_pairs, int C, int h, int w, void* const* levels, cudaStream_t st) {
    const int P = h * w;
    CfMaps maps;
    {
        const cuuint64_t dims[3] = {(cuuint64_t)C, (cuuint64_t)P, (cuuint64_t)n_pairs};
        const cuuint64_t strides[2] = {(cuuint64_t)C * 2, (cuuint64_t)P * C * 2};
        const cuuint32_t box[3] = {CF_BK, CF_BM, 1};
        int rc = cf_map(&maps.A, r1, 3, dims, strides, box, "fmap1 rows");
        if (rc) return rc;
    }
    {
        const cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n_pairs};
        const cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)w * C * 2, (cuuint64_t)P * C * 2};
        const cuuint32_t box[4] = {CF_BK, 8, 8, 1};
        int rc = cf_map(&maps.B, r2, 4, dims, strides, box, "fmap2 image");
        if (rc) return rc;
    }
    CfParams p{};
    p.lvl0 = static_cast<uint8_t*>(levels[0]);
    p.lvl1 = static_cast<uint8_t*>(levels[1]);
    p.P = P;
    p.nkb = C / CF_BK;
    p.stages = (int)std::min<size_t>(CF_MAX_STAGES, (226 * 1024 - (size_t)p.nkb * CF_A_BYTES) / CF_B_BYTES);
    p.nby = (h + 7) / 8;
    p.n_blocks = p.nby * ((w + 7) / 8);
    p.n_tiles = (p.n_blocks + 3) / 4;
    p.mt = (P + CF_BM - 1) / CF_BM;
    p.ntx0 = (w + 3) / 4; p.nty0 = (h + 3) / 4;
    p.lw1 = w / 2; p.lh1 = h / 2;
    p.ntx1 = (p.lw1 + 3) / 4; p.nty1 = (p.lh1 + 3) / 4;
    p.slab0 = (uint32_t)p.ntx0 * p.nty0 * 32;
    p.slab1 = (uint32_t)p.ntx1 * p.nty1 * 32;
    p.alpha = 1.0f / sqrtf((float)C);
    // chunks of the N tiles per (pair, M tile): whole waves of one-CTA-per-SM work, one tile's worth of fixed cost per CTA
    const int64_t units = (int64_t)n_pairs * p.mt;
    const int sms = device_sms();
    double best = 1e30;
    for (int c = 1; c <= std::min(p.n_tiles, 32); ++c) {
        const int per = (p.n_tiles + c - 1) / c, cc = (p.n_tiles + per - 1) / per;
        const double cost = (double)((units * cc + sms - 1) / sms) * (per + 1.0);
        if (cost < best - 1e-9) { best = cost; p.chunks = cc; p.tiles_per_cta = per; }
    }
    const int64_t grid = units * p.chunks;
    VFML_REQUIRE(grid < (1ll << 31), "corr_volume_fused: too many tiles");
    const size_t smem = (size_t)p.nkb * CF_A_BYTES + (size_t)p.stages * CF_B_BYTES + 1024;
    static PerDevice<bool> attr_done;
    if (!attr_done.get()) {
        VFML_CUDA(cudaFuncSetAttribute(corr_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        attr_done.get() = true;
    }
    corr_fused_kernel<<<(unsigned)grid, CF_THREADS, smem, st>>>(maps, p);
    VFML_LAUNCH_CHECK("corr_volume_fused");
    return VFML_OK;
}

}  // namespace vfml

// Gather-friendly ("tiled") correlation pyramid and its 4-level 9x9 bilinear lookup (fp16 tier).
//
// Same arithmetic as CorrBlock.__init__ / CorrBlock.__call__ (MOFNetStack/corr.py:59-97) and
// bilinear_sampler (core/utils/utils.py:62-76), but the pyramid is laid out for the gather:
//
//   level l of source pixel q is a slab of NTX_l x NTY_l tiles (NTX = ceil(w_l/4), NTY = ceil(h_l/4)),
//   tile (tx, ty) at byte ((tx*NTY + ty) * 32); a tile is 4 x 4 fp16 texels = ONE 32-byte DRAM sector,
//   stored x-major: element (xr, yy) = texel (4*tx + xr, 4*ty + yy); texels beyond the level are zero.
//
// The 10x10 footprint of the 81 taps then touches (10+3)/4 = 3.25 tiles per axis = 10.6 sectors
// per level on average, against 15.6 sectors for ten 20-byte row segments of the row-major slab,
// and the 4 tiles of a tile column are one contiguous 128-byte line.  Measured on B200
// (scripts/micro/sector_gather.cu): random gathers are served at ~30 G requests/s whatever the
// request size up to a line, and cp.async.cg / zero-fill lanes cost full requests, so the loads
// are predicated register loads (ld.global.nc.L1::no_allocate.v4, zeros outside the slab: the
// reference's zeros padding) issued two rounds ahead of their use.
//
// Lookup kernel: a warp owns a contiguous range of pixels and walks it in rounds of 3 pixels x 1
// level.  12 header lanes (pixel, level) turn the next batch's coordinates into tile origin,
// sub-tile offset, fractional weights and a 16-bit tile mask in shared memory; every lane then
// loads its half tile of the 4 x 4 tile footprint for each of the 3 pixels into registers, and
// one round later stores them to a ping-pong smem stage.  In the compute phase lane = (pixel v,
// x-column i in 0..9): it reads its x-column (the y run is contiguous in the x-major tile,
// realigned with a funnel shift), interpolates along y, takes column i+1 from the next lane by
// shuffle and interpolates along x.  Its 9 results are taps (i, j=0..8) of the reference's channel
// order l*81 + i*9 + j; they are written as ONE 16-byte store + one 2-byte store using the channel
// order   c' = (l*9 + i)*8 + j  (j < 8),   c' = 288 + l*9 + i  (j = 8),
// a fixed permutation of the 324 correlation channels that the engine folds into the weights of
// the convc1 block (every op of that block is channel-wise or a 1x1 convolution).
#include "common.cuh"

namespace vfml {

bool corr_fused_supported(int C, int h, int w, int num_levels);                 // corr_fused.cu
int corr_volume_fused(const __half* r1, const __half* r2, int n_pairs, int C, int h, int w, void* const* levels, cudaStream_t st);

constexpr int TL_LEVELS = 4;
constexpr int TL_UNITS = 3;            // pixels per round
constexpr int TL_AROW = 160;           // smem bytes per tile row (4 tiles x 32 B + 32 B: bank spread)
constexpr int TL_UNIT_BYTES = 4 * TL_AROW;
constexpr int TL_HDR_BYTES = 32;
constexpr int TL_STAGE_BYTES = TL_UNITS * TL_UNIT_BYTES;          // 1920
constexpr int TL_STAGES = 2;           // smem ping-pong; the loads in flight live in registers
constexpr int TL_BATCH_HDR = TL_UNITS * TL_LEVELS * TL_HDR_BYTES; // 384: headers of one batch (3 pixels x 4 levels)
constexpr int TL_HSLOTS = 2;           // batches whose headers are alive (issue runs < 1 batch ahead)
constexpr int TL_WARP_BYTES = TL_STAGES * TL_STAGE_BYTES + TL_HSLOTS * TL_BATCH_HDR;   // 8448
constexpr int TL_CTAS_PER_SM = 6;

constexpr int TL_WARPS = 4;
constexpr int TL_SMEM = TL_WARPS * TL_WARP_BYTES;

struct TiledSet {
    const uint8_t* lvl[TL_LEVELS];
    const float* coords;
    __half* out;
};

struct TiledParams {
    TiledSet set[2];
    int ntx[TL_LEVELS], nty[TL_LEVELS];
    int h, w, P;
    int per_set;        // n_frames * P
    int total;          // per_set * n_sets
    int n_frames;
    uint32_t slab_bytes[TL_LEVELS];
    int n_batches;      // ceil(total / 3)
    int64_t out_pitch;  // elements
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// streaming 16-byte load: read-only path, no L1 allocation (every texel is used once)
__device__ __forceinline__ uint4 ldg_stream128(const void* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

__device__ __forceinline__ uint32_t lds32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];\n" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ uint4 lds128(uint32_t a) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];\n" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts128(uint32_t a, uint4 v) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};\n" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ __half2 u2h2(uint32_t u) { return *reinterpret_cast<__half2*>(&u); }
__device__ __forceinline__ uint32_t h22u(__half2 h) { return *reinterpret_cast<uint32_t*>(&h); }

template <bool F32MATH>
__global__ void __launch_bounds__(TL_WARPS * 32, TL_CTAS_PER_SM)
corr_lookup_tiled_kernel(const TiledParams p) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // lets a PDL-launched successor start its prologue
    extern __shared__ __align__(128) uint8_t tl_smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t ring = smem_u32(tl_smem) + warp * TL_WARP_BYTES;
    const uint32_t hring = ring + TL_STAGES * TL_STAGE_BYTES;       // TL_HSLOTS batch slots x 12 unit headers
    const int gw = blockIdx.x * TL_WARPS + warp, nw = gridDim.x * TL_WARPS;
    const int b0 = (int)((int64_t)p.n_batches * gw / nw), b1 = (int)((int64_t)p.n_batches * (gw + 1) / nw);
    const int n_rounds = (b1 - b0) * TL_LEVELS;
    if (n_rounds == 0) return;

    // ---- header side: lane u < 12 owns unit (pixel v = u >> 2, level l = u & 3) of a batch.
    // (frame, pixel) advance incrementally by 3 per batch; frames of the second set follow the first's.
    const int hv = (lane >> 2) < TL_UNITS ? (lane >> 2) : 0, hl = lane & 3;
    const bool hlane = lane < TL_UNITS * TL_LEVELS;
    const int h_ntx = p.ntx[hl], h_nty = p.nty[hl];
    const uint32_t h_slabb = p.slab_bytes[hl];
    const float h_scale = __int_as_float((127 - hl) << 23);
    float ncx = 0.f, ncy = 0.f;
    int nown_si = 0, nown_fp = 0;
    bool nown_ok = false;
    int it_gp = b0 * TL_UNITS + hv;
    int it_f = it_gp / p.P, it_pix = it_gp - it_f * p.P;
    int it_batch = b0;
    auto fetch = [&]() {
        nown_ok = hlane && it_batch < b1 && it_gp < p.total;
        ncx = ncy = 0.f;
        nown_si = 0; nown_fp = 0;
        if (nown_ok) {
            nown_si = it_f >= p.n_frames;
            nown_fp = it_gp - nown_si * p.per_set;
            const int f = it_f - nown_si * p.n_frames;
            const float* c = p.set[nown_si].coords + ((uint64_t)(uint32_t)f * (uint32_t)(2 * p.P) + (uint32_t)it_pix);
            ncx = __ldg(c);
            ncy = __ldg(c + p.P);
        }
        it_gp += TL_UNITS; it_pix += TL_UNITS; ++it_batch;
        if (it_pix >= p.P) { it_pix -= p.P; ++it_f; }
    };
    // writes the 12 unit headers of the prefetched batch into header slot `slot`, prefetches the next
    auto headers = [&](int slot) {
        const float cx = ncx, cy = ncy;
        const int own_si = nown_si, own_fp = nown_fp;
        const bool own_ok = nown_ok;
        fetch();
        if (hlane) {
            // coords / 2^l (corr.py:88).  Beyond [-64, 30000] every tap is out of range anyway.
            float sx = fminf(fmaxf(cx * h_scale, -64.0f), 30000.0f);
            float sy = fminf(fmaxf(cy * h_scale, -64.0f), 30000.0f);
            if (!(sx == sx)) sx = -64.0f;      // NaN -> all taps out of range -> zeros
            if (!(sy == sy)) sy = -64.0f;
            const float flx = floorf(sx), fly = floorf(sy);
            const int x0 = (int)flx - 4, y0 = (int)fly - 4;
            const int tx0 = x0 >> 2, ty0 = y0 >> 2, ox = x0 & 3, oy = y0 & 3;
            // tile rows a in [alo, ahi] of the 4 x 4 tile footprint are inside the slab AND needed
            const int alo = max(0, -tx0), ahi = min((ox + 9) >> 2, h_ntx - 1 - tx0);
            const int blo = max(0, -ty0), bhi = min((oy + 9) >> 2, h_nty - 1 - ty0);
            uint32_t ma = (alo <= ahi) ? ((2u << ahi) - (1u << alo)) : 0u;
            const uint32_t mb = (blo <= bhi) ? ((2u << bhi) - (1u << blo)) : 0u;
            if (!own_ok) ma = 0u;
            // bit (a*4 + b) of mask16 = tile (a, b) is to be fetched
            const uint32_t spread = (ma & 1u) | ((ma & 2u) << 3) | ((ma & 4u) << 6) | ((ma & 8u) << 9);
            const uint32_t mask16 = spread * mb;
            const uint8_t* slab = p.set[own_si].lvl[hl] + (uint64_t)(uint32_t)own_fp * h_slabb;
            const __half* orow = p.set[own_si].out + (uint64_t)(uint32_t)own_fp * (uint32_t)p.out_pitch;
            uint4 h0, h1;
            h0.x = (uint32_t)(uintptr_t)slab;
            h0.y = (uint32_t)((uintptr_t)slab >> 32);
            h0.z = (uint32_t)((tx0 * h_nty + ty0) * 32);
            h0.w = ox | (oy << 2) | (own_ok ? 16 : 0) | (mask16 << 16);
            h1.x = __float_as_uint(sx - flx);
            h1.y = __float_as_uint(sy - fly);
            h1.z = (uint32_t)(uintptr_t)orow;
            h1.w = (uint32_t)((uintptr_t)orow >> 32);
            const uint32_t a = hring + slot * TL_BATCH_HDR + (hl * TL_UNITS + hv) * TL_HDR_BYTES;
            sts128(a, h0);
            sts128(a + 16, h1);
        }
        __syncwarp();
    };
    fetch();

    // ---- issue side: lane -> half tile (a, b, half) of the 4 x 4 tile footprint, for each of the 3 pixels
    const int ta = lane >> 3, tb = (lane >> 1) & 3, th = lane & 1;
    const uint32_t tile_dst = ta * TL_AROW + tb * 32 + th * 16;
    const int lane_lin = tb * 32 + th * 16;
    const int tbit = 16 + (lane >> 1);
    // loads of one round (level l of the batch in header slot `slot`): 3 x 16 bytes per lane, into registers.
    // Plain sector-granular loads: cp.async.cg was measured to pull whole 128-byte lines from L2/DRAM.
    auto load_round = [&](const int l, const int slot, uint4 (&r)[TL_UNITS]) {
        const uint32_t hdr = hring + slot * TL_BATCH_HDR + l * (TL_UNITS * TL_HDR_BYTES);
        const int lane_off = ta * p.nty[l] * 32 + lane_lin;
        uint4 hh[TL_UNITS];
#pragma unroll
        for (int v = 0; v < TL_UNITS; ++v) hh[v] = lds128(hdr + v * TL_HDR_BYTES);
#pragma unroll
        for (int v = 0; v < TL_UNITS; ++v) {
            r[v] = make_uint4(0u, 0u, 0u, 0u);          // outside the level (zeros padding) or not needed
            if ((hh[v].w >> tbit) & 1u) {
                const uint8_t* src = reinterpret_cast<const uint8_t*>((uintptr_t)hh[v].x | ((uintptr_t)hh[v].y << 32)) +
                                     ((int)hh[v].z + lane_off);
                r[v] = ldg_stream128(src);
            }
        }
    };
    auto store_round = [&](const int stage, const uint4 (&r)[TL_UNITS]) {
#pragma unroll
        for (int v = 0; v < TL_UNITS; ++v) sts128(ring + stage * TL_STAGE_BYTES + v * TL_UNIT_BYTES + tile_dst, r[v]);
    };

    // ---- process side: lane = (pixel v, x column i)
    const int pv = lane / 10, pi = lane - pv * 10;
    auto process = [&](const int l, const int slot) {
        const uint32_t sbase = ring + (l & 1) * TL_STAGE_BYTES;
        const uint32_t hdr = hring + slot * TL_BATCH_HDR + (l * TL_UNITS + (pv < TL_UNITS ? pv : 0)) * TL_HDR_BYTES;
        const uint32_t flags = lds32(hdr + 12);
        const uint4 h1 = lds128(hdr + 16);
        const int ox = flags & 3, oy = (flags >> 2) & 3;
        const float wx = __uint_as_float(h1.x), wy = __uint_as_float(h1.y);
        const int x = ox + pi;
        const uint32_t base = sbase + (pv < TL_UNITS ? pv : 0) * TL_UNIT_BYTES + (x >> 2) * TL_AROW + (x & 3) * 8;
        const bool e2 = (oy & 2) != 0;
        const uint32_t baseA = base + (e2 ? 4 : 0), baseB = base + (e2 ? 32 : 4);
        uint32_t q[6];
        q[0] = lds32(baseA); q[1] = lds32(baseB);
        q[2] = lds32(baseA + 32); q[3] = lds32(baseB + 32);
        q[4] = lds32(baseA + 64); q[5] = lds32(baseB + 64);
        const uint32_t sh = (oy & 1) * 16;
        uint32_t o[5];
        if (F32MATH) {
            float U[10];
#pragma unroll
            for (int k = 0; k < 5; ++k) {
                const float2 t = __half22float2(u2h2(__funnelshift_r(q[k], q[k + 1], sh)));
                U[2 * k] = t.x; U[2 * k + 1] = t.y;
            }
            float pcol[9];
#pragma unroll
            for (int j = 0; j < 9; ++j) pcol[j] = fmaf(U[j + 1] - U[j], wy, U[j]);
#pragma unroll
            for (int j = 0; j < 9; ++j) {
                const float nb = __shfl_down_sync(0xffffffffu, pcol[j], 1);
                pcol[j] = fmaf(nb - pcol[j], wx, pcol[j]);
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) o[k] = h22u(__floats2half2_rn(pcol[2 * k], pcol[2 * k + 1]));
            o[4] = h22u(__floats2half2_rn(pcol[8], 0.0f));
        } else {
            const __half2 wy2 = __float2half2_rn(wy), wx2 = __float2half2_rn(wx);
            __half2 pc[5];
#pragma unroll
            for (int k = 0; k < 5; ++k) {
                const __half2 u = u2h2(__funnelshift_r(q[k], q[k + 1], sh));
                const __half2 s = u2h2(__funnelshift_rc(q[k], q[k + 1], sh + 16));
                pc[k] = __hfma2(__hsub2(s, u), wy2, u);
            }
#pragma unroll
            for (int k = 0; k < 5; ++k) {
                const __half2 nb = u2h2(__shfl_down_sync(0xffffffffu, h22u(pc[k]), 1));
                o[k] = h22u(__hfma2(__hsub2(nb, pc[k]), wx2, pc[k]));
            }
        }
        if (pv < TL_UNITS && pi < 9 && (flags & 16)) {
            __half* orow = reinterpret_cast<__half*>((uintptr_t)h1.z | ((uintptr_t)h1.w << 32));
            *reinterpret_cast<uint4*>(orow + (l * 9 + pi) * 8) = make_uint4(o[0], o[1], o[2], o[3]);
            reinterpret_cast<unsigned short*>(orow)[288 + l * 9 + pi] = (unsigned short)(o[4] & 0xffff);
        }
    };

    // Round g = (batch k, level l); its data sits in smem stage l & 1 while it is processed.  At iteration g
    // the loads of round g+2 are issued into the register set g & 1 (just emptied), round g is processed,
    // and round g+1 moves from registers to the other stage.
    const int nb = b1 - b0;
    uint4 ra[TL_UNITS], rb[TL_UNITS];
    headers(0);
    load_round(0, 0, ra);
    load_round(1, 0, rb);
    store_round(0, ra);
#pragma unroll 1
    for (int kb = 0; kb < nb; ++kb) {
        const int slot = kb & 1;
        const bool more = kb + 1 < nb;
        // l = 0
        load_round(2, slot, ra);
        __syncwarp();
        process(0, slot);
        __syncwarp();
        store_round(1, rb);
        // l = 1
        load_round(3, slot, rb);
        __syncwarp();
        process(1, slot);
        __syncwarp();
        store_round(0, ra);
        // l = 2
        if (more) {
            headers(slot ^ 1);
            load_round(0, slot ^ 1, ra);
        }
        __syncwarp();
        process(2, slot);
        __syncwarp();
        store_round(1, rb);
        // l = 3
        if (more) load_round(1, slot ^ 1, rb);
        __syncwarp();
        process(3, slot);
        __syncwarp();
        if (more) store_round(0, ra);
    }
}

// ---------------------------------------------------------------------------------------------
// Pyramid builder: one CTA per source pixel.  Reads that pixel's row of the row-major level-0
// volume, pools it three times (avg_pool2d(2, 2), each level rounded to fp16 like the stored
// pyramid of the row-major path) and writes all levels in the tiled layout.
// ---------------------------------------------------------------------------------------------
struct TilePyrParams {
    const __half* vol0;       // [rows, h*w]
    uint8_t* lvl[TL_LEVELS];
    int lh[TL_LEVELS], lw[TL_LEVELS], ntx[TL_LEVELS], nty[TL_LEVELS];
    int soff[TL_LEVELS];      // smem offsets (elements)
    int num_levels;
    int first;                // 0: start from the row-major level 0 (vol0); 1: levels 0 and 1 exist (corr_fused.cu), start from the TILED level 1
    int64_t rows;
};

__global__ void __launch_bounds__(256)
corr_tile_pyramid_kernel(const TilePyrParams p) {
    extern __shared__ __align__(16) __half tp_s[];
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int P0 = p.lh[0] * p.lw[0];
    for (int64_t row = blockIdx.x; row < p.rows; row += gridDim.x) {
        if (p.first) {
            // un-tile level 1: thread = (tile, x column of the tile); the 4 texels of a column are 8 contiguous bytes
            const int lw = p.lw[1], lh = p.lh[1], nty = p.nty[1], nt = p.ntx[1] * nty;
            const uint2* src = reinterpret_cast<const uint2*>(p.lvl[1] + row * ((int64_t)nt * 32));
            __half* out = tp_s + p.soff[1];
            for (int i = tid; i < nt * 4; i += nthr) {
                const int t = i >> 2, xr = i & 3, tx = t / nty, ty = t - tx * nty, x = tx * 4 + xr;
                const uint2 q = __ldg(src + i);
                const __half* hv = reinterpret_cast<const __half*>(&q);
                if (x < lw)
                    for (int yy = 0; yy < 4; ++yy)
                        if (ty * 4 + yy < lh) out[(ty * 4 + yy) * lw + x] = hv[yy];
            }
        } else {
            const __half* src = p.vol0 + row * P0;
            if ((P0 & 7) == 0) {
                const uint4* s4 = reinterpret_cast<const uint4*>(src);
                uint4* d4 = reinterpret_cast<uint4*>(tp_s);
                for (int i = tid; i < P0 / 8; i += nthr) d4[i] = __ldg(s4 + i);
            } else {
                for (int i = tid; i < P0; i += nthr) tp_s[i] = src[i];
            }
        }
        __syncthreads();
        for (int l = p.first + 1; l < p.num_levels; ++l) {
            const int wi = p.lw[l - 1], wo = p.lw[l], ho = p.lh[l];
            const __half* in = tp_s + p.soff[l - 1];
            __half* out = tp_s + p.soff[l];
            for (int i = tid; i < wo * ho; i += nthr) {
                const int y = i / wo, x = i - y * wo;
                const __half* s = in + (2 * y) * wi + 2 * x;
                const float a = ((__half2float(s[0]) + __half2float(s[1])) + __half2float(s[wi])) + __half2float(s[wi + 1]);
                out[i] = __float2half_rn(a * 0.25f);
            }
            __syncthreads();
        }
        for (int l = p.first ? 2 : 0; l < p.num_levels; ++l) {
            const int lw = p.lw[l], lh = p.lh[l], nty = p.nty[l], nt = p.ntx[l] * nty;
            const __half* in = tp_s + p.soff[l];
            uint4* dst = reinterpret_cast<uint4*>(p.lvl[l] + row * ((int64_t)nt * 32));
            for (int t = tid; t < nt; t += nthr) {
                const int tx = t / nty, ty = t - tx * nty;
                __align__(16) __half v[16];
#pragma unroll
                for (int xr = 0; xr < 4; ++xr)
#pragma unroll
                    for (int yy = 0; yy < 4; ++yy) {
                        const int x = tx * 4 + xr, y = ty * 4 + yy;
                        v[xr * 4 + yy] = (x < lw && y < lh) ? in[y * lw + x] : __float2half_rn(0.0f);
                    }
                dst[2 * t] = *reinterpret_cast<const uint4*>(v);
                dst[2 * t + 1] = *reinterpret_cast<const uint4*>(v + 8);
            }
        }
        __syncthreads();
    }
}

static void level_dims(int h, int w, int num_levels, int* lh, int* lw, int* ntx, int* nty) {
    for (int l = 0; l < num_levels; ++l) {
        lh[l] = h; lw[l] = w;
        ntx[l] = (w + 3) / 4; nty[l] = (h + 3) / 4;
        h /= 2; w /= 2;
    }
}

}  // namespace vfml

using namespace vfml;

extern "C" size_t vfml_corr_tiled_slab_bytes(int h, int w, int level) {
    if (h <= 0 || w <= 0 || level < 0 || level >= TL_LEVELS) return 0;
    h >>= level; w >>= level;
    return (size_t)((w + 3) / 4) * ((h + 3) / 4) * 32;
}

// first = 0: all levels from the row-major level 0; first = 1: levels 2.. from the tiled level 1 that is already in levels[1]
static int tile_pyramid_launch(const void* vol0, int64_t rows, int h, int w, void* const* levels, int num_levels, int first, cudaStream_t st) {
    TilePyrParams p{};
    level_dims(h, w, num_levels, p.lh, p.lw, p.ntx, p.nty);
    int off = 0;
    for (int l = 0; l < num_levels; ++l) {
        VFML_REQUIRE(levels[l], "corr_tile_pyramid: level %d is null", l);
        VFML_REQUIRE(p.lh[l] > 0 && p.lw[l] > 0, "corr_tile_pyramid: level %d is empty (h=%d, w=%d)", l, h, w);
        p.lvl[l] = reinterpret_cast<uint8_t*>(levels[l]);
        p.soff[l] = off;
        if (l >= first) off += (p.lh[l] * p.lw[l] + 7) & ~7;
    }
    p.vol0 = reinterpret_cast<const __half*>(vol0);
    p.num_levels = num_levels;
    p.first = first;
    p.rows = rows;
    const size_t smem = (size_t)off * sizeof(__half);
    VFML_REQUIRE(smem <= 200 * 1024, "corr_tile_pyramid: %d x %d does not fit shared memory", h, w);
    static PerDevice<size_t> attr;
    if (smem > 48 * 1024 && smem > attr.get()) {
        VFML_CUDA(cudaFuncSetAttribute(corr_tile_pyramid_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr.get() = smem;
    }
    const int threads = first ? 128 : 256;         // the level-1 slab is a quarter of level 0: smaller CTAs, more rows in flight
    const int blocks = (int)std::min<int64_t>(rows, (int64_t)device_sms() * 16);
    corr_tile_pyramid_kernel<<<blocks, threads, smem, st>>>(p);
    VFML_LAUNCH_CHECK("corr_tile_pyramid");
    return VFML_OK;
}

extern "C" int vfml_corr_tile_pyramid(const void* vol0, int64_t rows, int h, int w, void* const* levels, int num_levels,
                                      vfml_stream_t s) {
    VFML_REQUIRE(rows >= 0 && h > 0 && w > 0, "corr_tile_pyramid: bad shape");
    VFML_REQUIRE(num_levels >= 1 && num_levels <= TL_LEVELS, "corr_tile_pyramid: num_levels %d not in 1..4", num_levels);
    if (rows == 0) return VFML_OK;
    VFML_REQUIRE(vol0 && levels, "corr_tile_pyramid: null pointer");
    return tile_pyramid_launch(vol0, rows, h, w, levels, num_levels, 0, to_stream(s));
}

extern "C" size_t vfml_corr_volume_tiled_workspace(int n_pairs, int C, int h, int w) {
    const size_t P = (size_t)h * w;
    return (size_t)2 * n_pairs * P * C * sizeof(__half) + (size_t)n_pairs * P * P * sizeof(__half) + 512;
}

extern "C" int vfml_corr_volume_pyramid_tiled(const float* fmap1, const float* fmap2, int n_pairs, int C, int h, int w,
                                              void* const* levels, int num_levels, void* workspace,
                                              size_t workspace_bytes, vfml_stream_t s) {
    VFML_REQUIRE(n_pairs >= 0 && C > 0 && h > 0 && w > 0, "corr_volume_pyramid_tiled: bad shape");
    VFML_REQUIRE((C % 8) == 0, "corr_volume_pyramid_tiled: C=%d must be a multiple of 8", C);
    if (n_pairs == 0) return VFML_OK;
    VFML_REQUIRE(fmap1 && fmap2 && levels, "corr_volume_pyramid_tiled: null pointer");
    const size_t need = vfml_corr_volume_tiled_workspace(n_pairs, C, h, w);
    if (!workspace || workspace_bytes < need) {
        set_error("corr_volume_pyramid_tiled: workspace of %zu bytes needed, %zu given", need, workspace_bytes);
        return VFML_ERR_WORKSPACE;
    }
    const int P = h * w;
    __half* r1 = reinterpret_cast<__half*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
    __half* r2 = r1 + (size_t)n_pairs * P * C;
    __half* vol0 = r2 + (size_t)n_pairs * P * C;
    int rc = vfml_nchw_to_rows(fmap1, r1, n_pairs, C, P, C, 0, VFML_F16, s);
    if (rc) return rc;
    rc = vfml_nchw_to_rows(fmap2, r2, n_pairs, C, P, C, 0, VFML_F16, s);
    if (rc) return rc;
    if (corr_fused_supported(C, h, w, num_levels)) {
        // levels 0 and 1 straight from the accumulators (corr_fused.cu), levels 2 and 3 from the tiled level 1
        for (int l = 0; l < num_levels; ++l) VFML_REQUIRE(levels[l], "corr_volume_pyramid_tiled: level %d is null", l);
        rc = corr_volume_fused(r1, r2, n_pairs, C, h, w, levels, to_stream(s));
        if (rc) return rc;
        return tile_pyramid_launch(nullptr, (int64_t)n_pairs * P, h, w, levels, num_levels, 1, to_stream(s));
    }
    vfml_gemm_desc d{};
    d.A = r1; d.B = r2; d.D = vol0;
    d.lda = C; d.ldb = C; d.ldd = P;
    d.M = P; d.N = P; d.K = C; d.batch = n_pairs;
    d.strideA = (int64_t)P * C; d.strideB = (int64_t)P * C; d.strideD = (int64_t)P * P;
    d.alpha = 1.0f / sqrtf((float)C);
    d.in_dtype = VFML_F16; d.out_dtype = VFML_F16;
    rc = vfml_gemm(&d, s);
    if (rc) return rc;
    return vfml_corr_tile_pyramid(vol0, (int64_t)n_pairs * P, h, w, levels, num_levels, s);
}

extern "C" int vfml_corr_lookup_tiled(const void* const* levels_a, const void* const* levels_b, int n_frames, int h, int w,
                                      const float* coords_a, const float* coords_b, void* out_a, void* out_b,
                                      int out_pitch, int out_col, int math_f32, vfml_stream_t s) {
    VFML_REQUIRE(n_frames >= 0 && h > 0 && w > 0, "corr_lookup_tiled: bad shape");
    VFML_REQUIRE((h >> 3) > 0 && (w >> 3) > 0, "corr_lookup_tiled: %d x %d is too small for 4 levels", h, w);
    VFML_REQUIRE(w <= 16384 && h <= 16384, "corr_lookup_tiled: level larger than 16384");
    VFML_REQUIRE((out_pitch % 8) == 0 && (out_col % 8) == 0 && out_col >= 0 && out_pitch >= out_col + 324,
                 "corr_lookup_tiled: pitch %d / col %d must be multiples of 8 with room for 324 channels", out_pitch, out_col);
    if (n_frames == 0) return VFML_OK;
    VFML_REQUIRE(levels_a && coords_a && out_a, "corr_lookup_tiled: null pointer");
    VFML_REQUIRE((reinterpret_cast<uintptr_t>(out_a) & 15) == 0 && (reinterpret_cast<uintptr_t>(out_b) & 15) == 0,
                 "corr_lookup_tiled: output rows must be 16-byte aligned");
    const int n_sets = levels_b ? 2 : 1;
    if (levels_b) VFML_REQUIRE(coords_b && out_b, "corr_lookup_tiled: null pointer (second set)");
    const int64_t total = (int64_t)n_frames * h * w * n_sets;
    VFML_REQUIRE(total < (1ll << 30), "corr_lookup_tiled: too many pixels");
    TiledParams p{};
    int lh[TL_LEVELS], lw[TL_LEVELS];
    level_dims(h, w, TL_LEVELS, lh, lw, p.ntx, p.nty);
    for (int l = 0; l < TL_LEVELS; ++l) {
        VFML_REQUIRE(levels_a[l] && (!levels_b || levels_b[l]), "corr_lookup_tiled: level %d is null", l);
        p.set[0].lvl[l] = reinterpret_cast<const uint8_t*>(levels_a[l]);
        p.set[1].lvl[l] = levels_b ? reinterpret_cast<const uint8_t*>(levels_b[l]) : nullptr;
    }
    p.set[0].coords = coords_a; p.set[0].out = reinterpret_cast<__half*>(out_a) + out_col;
    p.set[1].coords = coords_b; p.set[1].out = reinterpret_cast<__half*>(out_b) + out_col;
    p.h = h; p.w = w; p.P = h * w;
    p.per_set = n_frames * h * w;
    p.n_frames = n_frames;
    for (int l = 0; l < TL_LEVELS; ++l) p.slab_bytes[l] = (uint32_t)p.ntx[l] * p.nty[l] * 32;
    p.total = (int)total;
    p.n_batches = (p.total + TL_UNITS - 1) / TL_UNITS;
    p.out_pitch = out_pitch;
    static PerDevice<bool> attr_done;
    if (!attr_done.get()) {
        VFML_CUDA(cudaFuncSetAttribute(corr_lookup_tiled_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, TL_SMEM));
        VFML_CUDA(cudaFuncSetAttribute(corr_lookup_tiled_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TL_SMEM));
        attr_done.get() = true;
    }
    const int blocks = std::min(device_sms() * TL_CTAS_PER_SM, (p.n_batches + TL_WARPS - 1) / TL_WARPS);
    if (math_f32) corr_lookup_tiled_kernel<true><<<blocks, TL_WARPS * 32, TL_SMEM, to_stream(s)>>>(p);
    else corr_lookup_tiled_kernel<false><<<blocks, TL_WARPS * 32, TL_SMEM, to_stream(s)>>>(p);
    VFML_LAUNCH_CHECK("corr_lookup_tiled");
    return VFML_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// Allocation-free variant of vfml_corr_volume_pyramid_tiled (no counterpart in the reference; added
// for callers that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that
// forbid cudaMalloc).  The workspace (fp16 rows copies of both feature maps + the level-0 volume) lives in a file-scope
// device array instead of a caller buffer.  Every GPU that loads the library holds these 16 MiB until the process exits.
//   * the array's name is only a host-side stand-in here: the device address comes from cudaGetSymbolAddress, once per
//     device with that device current, and that pointer is what the kernels get;
//   * nothing in it has to start at zero: the row copies and the volume GEMM overwrite every byte that is read later;
//   * every call on a device shares the one array: calls that overlap on the GPU (different streams, no dependency) race.
constexpr size_t STATIC_WS_BYTES = (size_t)16 << 20;
static __device__ __align__(256) unsigned char g_static_workspace[STATIC_WS_BYTES];

extern "C" size_t vfml_corr_volume_pyramid_tiled_static_max_bytes(void) { return STATIC_WS_BYTES; }

extern "C" int vfml_corr_volume_pyramid_tiled_static(const float* fmap1, const float* fmap2, int n_pairs, int C, int h, int w,
                                                     void* const* levels, int num_levels, vfml_stream_t s) {
    VFML_REQUIRE(n_pairs >= 0 && C > 0 && h > 0 && w > 0, "corr_volume_pyramid_tiled_static: bad shape");
    const size_t need = vfml_corr_volume_tiled_workspace(n_pairs, C, h, w);
    if (need > STATIC_WS_BYTES) {
        set_error("corr_volume_pyramid_tiled_static: n_pairs=%d C=%d h=%d w=%d needs %zu bytes of scratch, the built-in array "
                  "holds %zu; use vfml_corr_volume_pyramid_tiled with a caller workspace", n_pairs, C, h, w, need, STATIC_WS_BYTES);
        return VFML_ERR_WORKSPACE;
    }
    if (getenv("VFML_GEMM_TRACE")) {       // the debug trace of the GEMM allocates its counter buffer
        set_error("corr_volume_pyramid_tiled_static: not available with VFML_GEMM_TRACE set (the trace buffer is allocated)");
        return VFML_ERR_UNSUPPORTED;
    }
    static PerDevice<void*> addr;
    void*& ws = addr.get();
    if (!ws) {
        void* p = nullptr;
        VFML_CUDA(cudaGetSymbolAddress(&p, g_static_workspace));
        ws = p;
    }
    return vfml_corr_volume_pyramid_tiled(fmap1, fmap2, n_pairs, C, h, w, levels, num_levels, ws, STATIC_WS_BYTES, s);
}

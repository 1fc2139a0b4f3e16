// Weight-stationary cluster GEMM for the update block's pointwise layers (fp16 tier):
//     D[m, n] = epi(sum_k A[m, k] * B[n, k]),  M = 21 120 .. 42 240 rows, N, K = 128 .. 768, B = a conv weight.
// Same contract as gemm_tc.cu (which stays the general kernel: batches, split-K, fp32 output, ReLU, alpha); this kernel takes
// the layers that dominate the step.  What bounds them was measured in round 2 (scripts/micro/store_probe.cu, mma_loop.cu,
// pair_probe.cu, profiles/r2_gemm_ws_*.txt):
//   - not the tensor pipe (removing the MMAs from gemm_tc changes nothing), but bytes: each SM exchanges ~33 B/clk with L2
//     when all 148 are busy, and its shared memory moves ~105 B/clk in total (TMA writes + MMA operand reads + epilogue);
//   - gemm_tc re-reads the activation tile A once per column tile (3 x 27 MB for a 42240 x 486 x 324 layer) and its
//     lane-per-row global stores / residual loads run at 16 B/clk/SM, serialised with the operand stream;
//   - a single-CTA tcgen05.mma with N <= 128 is bound by its 128 x 16 A-operand read from shared memory (72 cycles
//     whatever N is); the CTA-PAIR form (cta_group::2, M = 256) runs at N/2 cycles down to N = 128 and reads half as
//     many operand bytes per SM and FLOP.
// Hence:
//   * a thread-block CLUSTER owns one row tile.  Pair mode: CTAs 2i, 2i+1 form a tcgen05 CTA pair that computes
//     256 rows x BN columns (column tile i) -- CTA c of the pair holds rows 128c.. of A and HALF of the resident weight
//     slice, the leader issues M = 256 MMAs.  Single mode: CTA r computes 128 rows x column tile r.
//   * the CTAs that need the same A rows (same c, all column tiles) each fetch a share of every A k-block and TMA-MULTICAST
//     it to the others: A leaves L2 once per row tile.  The slot-free signal is the mirror image: tcgen05.commit with
//     a cluster mask arrives on the `empty` barrier of every CTA of the cluster.
//   * the weight slice of the column tile stays in shared memory for the whole kernel.
//   * the epilogue never touches global memory with ld / st: the residual tile arrives by TMA (32 x 32 boxes, 64-byte
//     swizzle, one unit ahead, three boxes per warp), is updated IN PLACE in shared memory (a lane owns a row: its
//     16-byte chunks are conflict-free under the swizzle) and leaves by TMA store.
//   * up to four TMEM accumulators (BN <= 128) decouple the MMA warp from the epilogue warps.
// Pair-mode plumbing: every TMA load of either CTA of a pair is issued in its .cta_group::2 form with the mbarrier address
// of the pair's LEADER (peer bit cleared), so the leader's `full` barrier collects both CTAs' bytes -- also for the multicast
// shares that arrive from other pairs (measured: a software relay through the peer's own barrier costs ~1.3 K cycles per
// slot, profiles/r2_gemm_ws_timeline.txt); the peer's epilogue warps hand the accumulator back with remote arrives on
// the leader's `acc_empty`.
//
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (idle in the peer CTA of a pair), warps 2.. = epilogue (warp w reads TMEM
// lanes 32 (w % 4) ..; the warps of one lane quarter split the tile's 32-column boxes round-robin).
#include <cuda.h>
#include <stdlib.h>

#include <map>
#include <mutex>
#include <tuple>

#include "common.cuh"
#include "gelu_poly.cuh"
#include "tcgen05.cuh"

namespace vfml {

constexpr int WS_BM = 128, WS_BK = 64;
#ifndef VFML_WS_EPI_WARPS
#define VFML_WS_EPI_WARPS 12         // measured (profiles/r2_gemm_ws_variants.txt): 12 warps / 2 boxes beat 8 / 3 on the residual layers
#endif
constexpr int WS_EPI_WARPS = VFML_WS_EPI_WARPS;           // multiple of 4
constexpr int WS_NH = WS_EPI_WARPS / 4;                   // warps per TMEM lane quarter
constexpr int WS_THREADS = 64 + WS_EPI_WARPS * 32;
constexpr int WS_MAX_STAGES = 8, WS_MAX_CS = 8, WS_MAX_ACC = 4;
#ifndef VFML_WS_NBUF
#define VFML_WS_NBUF 2
#endif
constexpr int WS_BOX_COLS = 32, WS_BOX_BYTES = 32 * WS_BOX_COLS * 2, WS_NBUF = VFML_WS_NBUF;
constexpr uint32_t WS_A_BYTES = WS_BM * WS_BK * 2;
// A ring slot ("stage") = sk (1 or 2) k-blocks behind ONE full / empty barrier pair.  Single mode uses 2: the issuing thread
// pays its barrier wait, fence and commit (~85 cycles, scripts/micro/mma_loop.cu mode 39 vs 32) once per 8 MMAs.  Pair
// mode uses 1: its MMAs take >= 64 cycles each and hide the loop, and the finer ring keeps more loads in flight.
constexpr int WS_MAX_SK = 2;
#ifndef VFML_WS_SPIN
#define VFML_WS_SPIN 1
#endif
__device__ __forceinline__ void ws_wait(uint64_t* bar, uint32_t parity) { if (VFML_WS_SPIN) mbar_spin(bar, parity); else mbar_wait(bar, parity); }
__device__ __forceinline__ void ws_wait_cluster(uint64_t* bar, uint32_t parity) { if (VFML_WS_SPIN) mbar_spin_cluster(bar, parity); else mbar_wait_cluster(bar, parity); }

struct WsMaps {
    CUtensorMap A;                 // [M, K] fp16, box 64 x (128 / sharers) rows, 128-byte swizzle
    CUtensorMap B;                 // [N, K] fp16, box 64 x (BN or BN / 2) rows, 128-byte swizzle
    CUtensorMap R;                 // residual [M, N] fp16, box 32 x 32, 64-byte swizzle
    CUtensorMap D[WS_MAX_CS];      // output columns of column tile t: base D + t BN, width min(BN, N - t BN) & ~7: stores clip there
};

struct WsParams {
    const float* bias;
    const float* dw_w;
    const float* dw_b;
    __half* D;                     // for the ragged column tail only (N % 8 != 0): the bulk stores clip at 16-byte granularity
    int64_t ldd;
    int M, N, K;
    int BN, stages, sk, row_tiles, cs, n_clusters, nacc;
    uint32_t epi;
    int b_is_const;
    long long* trace;
};

template <int SPEC, bool PAIR>
__global__ void __launch_bounds__(WS_THREADS, 1)
gemm_ws_kernel(const __grid_constant__ WsMaps maps, const WsParams p) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int BN = p.BN, stages = p.stages, CS = p.cs, SK = p.sk;
    const int BH = PAIR ? BN / 2 : BN;                           // weight rows this CTA keeps
    const uint32_t b_bytes = (uint32_t)BH * WS_BK * 2, b_stride = (b_bytes + 1023) & ~1023u;
    const uint32_t stage_bytes = (uint32_t)SK * WS_A_BYTES;
    const int nkb = (p.K + WS_BK - 1) / WS_BK;
    uint8_t* sB = smem;                                         // [nkb] resident weight blocks
    uint8_t* sA = sB + (size_t)nkb * b_stride;                  // [stages] activation ring
    uint8_t* sBox = sA + (size_t)stages * stage_bytes;          // [epilogue warp][WS_NBUF] 32 x 32 fp16 boxes
    float* s_vec = reinterpret_cast<float*>(sBox + (size_t)WS_EPI_WARPS * WS_NBUF * WS_BOX_BYTES);   // [3][BN]: bias, dw_w, dw_b of MY columns
    __shared__ uint64_t full_bar[WS_MAX_STAGES], empty_bar[WS_MAX_STAGES];
    __shared__ uint64_t acc_full[WS_MAX_ACC], acc_empty[WS_MAX_ACC], b_full;
    __shared__ uint64_t res_full[WS_EPI_WARPS * WS_NBUF];
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();                    // 1-D cluster along x: blockIdx.x % CS
    const int cluster_id = (int)blockIdx.x / CS;
    const int c = PAIR ? (rank & 1) : 0;                        // which 128 rows of the row tile
    const int ct = PAIR ? (rank >> 1) : rank;                   // column tile
    const int sharers = PAIR ? CS / 2 : CS;                     // CTAs that need the same A rows
    const int leader = rank & ~1;
    const int n0 = ct * BN;
    const int cols_valid = min(BN, p.N - n0);                   // > 0: the host never creates an empty column tile
    const int nbx = (cols_valid + WS_BOX_COLS - 1) / WS_BOX_COLS;
    const int epi_active = 4 * min(WS_NH, nbx);                 // epilogue warps that have boxes
    const int TM = PAIR ? 2 * WS_BM : WS_BM;                    // rows per cluster tile
    const int t_first = cluster_id, t_step = p.n_clusters, t_end = p.row_tiles;
    const uint16_t mask_all = (uint16_t)((1u << CS) - 1u);
    uint16_t mask_a = mask_all;                                 // the CTAs that receive my A shares
    if (PAIR) {
        mask_a = 0;
        for (int i = 0; i < sharers; ++i) mask_a |= (uint16_t)(1u << (2 * i + c));
    }
    const uint16_t mask_pair = (uint16_t)(3u << leader);
    const int NACC = p.nacc;
    const uint32_t acc_cols = 512u / (uint32_t)NACC;

    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], (uint32_t)sharers);       // one commit from the MMA issuer of every column tile
        }
        for (int b = 0; b < WS_MAX_ACC; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], (uint32_t)(PAIR ? 2 * epi_active : epi_active));
        }
        mbar_init(&b_full, 1);
        for (int i = 0; i < WS_EPI_WARPS * WS_NBUF; ++i) mbar_init(&res_full[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.A) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.B) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.D[ct]) : "memory");
        if ((SPEC & VFML_EPI_RES) != 0) asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.R) : "memory");
    }
    if (warp == 1) {
        if (PAIR) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512u) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512u) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    if (!p.b_is_const) asm volatile("griddepcontrol.wait;" ::: "memory");     // bias / dw vectors may be activations
    if (warp >= 2) {
        for (int i = threadIdx.x - 64; i < BN; i += WS_EPI_WARPS * 32) {
            const bool ok = n0 + i < p.N;
            s_vec[i] = (ok && (SPEC & VFML_EPI_BIAS)) ? p.bias[n0 + i] : 0.0f;
            s_vec[BN + i] = (ok && (SPEC & VFML_EPI_DW1)) ? p.dw_w[n0 + i] : 0.0f;
            s_vec[2 * BN + i] = (ok && (SPEC & VFML_EPI_DW1)) ? p.dw_b[n0 + i] : 0.0f;
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    cluster_sync_all();        // every CTA of the cluster has initialised the barriers its peers will signal
    const uint32_t tmem_base = tmem_base_s;

    // the resident weight slice depends on constants only: fetched before the preceding kernel has finished
    // (pair mode: every load of either CTA signals the LEADER's barrier, which expects both CTAs' bytes)
    auto load_weights = [&]() {
        if (!PAIR || c == 0) mbar_expect_tx(&b_full, (uint32_t)nkb * b_bytes * (PAIR ? 2u : 1u));
        for (int kb = 0; kb < nkb; ++kb) {
            if (PAIR) tma_load_3d_2sm(&maps.B, leader_bar_addr(&b_full), sB + (size_t)kb * b_stride, kb * WS_BK, n0 + c * BH, 0);
            else tma_load_3d(&maps.B, &b_full, sB + (size_t)kb * b_stride, kb * WS_BK, n0 + c * BH, 0);
        }
    };
    if (warp == 0 && p.b_is_const && t_first < t_end && elect_one()) load_weights();
    asm volatile("griddepcontrol.wait;" ::: "memory");

    if (warp == 0) {
        // ---- producer: my share of every A k-block of my 128 rows, multicast to the CTAs that need the same rows
        int s = 0;
        uint32_t ph = 0;
        const bool tr = p.trace != nullptr;
        long long t_wait = 0, t0 = tr ? clock64() : 0;
        if (!p.b_is_const && t_first < t_end && elect_one()) load_weights();
        __syncwarp();
        // the A block is fetched in 2^k equal shares (swizzle atoms are 8 rows) by the first 2^k sharers; with 3 or 5..7 column
        // tiles the remaining sharers only receive
        int loaders = 1;
        while (loaders * 2 <= sharers) loaders *= 2;
        const bool i_load = ct < loaders;
        const int slice_rows = WS_BM / loaders;
        const uint32_t slice_off = (uint32_t)(ct * slice_rows) * 128u;
        for (int tile = t_first; tile < t_end; tile += t_step) {
            const int m0 = tile * TM + c * WS_BM;
            for (int kb = 0; kb < nkb; kb += SK) {
                const long long w0 = tr ? clock64() : 0;
                ws_wait(&empty_bar[s], ph ^ 1u);                 // every consumer of this slot, in every CTA I write to, has released it
                if (tr) t_wait += clock64() - w0;
                if (tr && tile == t_first + 2 * t_step && kb < 8 && lane == 0) {
                    p.trace[4096 * 16 + blockIdx.x * 32 + kb] = w0;              // started waiting for the slot
                    p.trace[4096 * 16 + blockIdx.x * 32 + 8 + kb] = clock64();   // slot free: issuing the loads
                }
                if (elect_one()) {
                    const int nblk = min(SK, nkb - kb);
                    // whole blocks: my shares + the other sharers' (pair mode: of both CTAs, on the leader's barrier)
                    if (!PAIR || c == 0) mbar_expect_tx(&full_bar[s], (uint32_t)nblk * WS_A_BYTES * (PAIR ? 2u : 1u));
#pragma unroll
                    for (int i = 0; i < WS_MAX_SK; ++i) {
                        if (i < nblk && i_load) {
                            uint8_t* dst = sA + (size_t)s * stage_bytes + (size_t)i * WS_A_BYTES + slice_off;
                            if (PAIR) {
                                if (sharers > 1) tma_load_3d_2sm_multicast(&maps.A, leader_bar_addr(&full_bar[s]), dst, (kb + i) * WS_BK, m0 + ct * slice_rows, 0, mask_a);
                                else tma_load_3d_2sm(&maps.A, leader_bar_addr(&full_bar[s]), dst, (kb + i) * WS_BK, m0, 0);
                            } else {
                                if (sharers > 1) tma_load_3d_multicast(&maps.A, &full_bar[s], dst, (kb + i) * WS_BK, m0 + ct * slice_rows, 0, mask_a);
                                else tma_load_3d(&maps.A, &full_bar[s], dst, (kb + i) * WS_BK, m0, 0);
                            }
                        }
                    }
                }
                __syncwarp();
                if (++s == stages) { s = 0; ph ^= 1u; }
            }
        }
        if (tr && lane == 0) { p.trace[blockIdx.x * 16 + 0] = t_wait; p.trace[blockIdx.x * 16 + 1] = clock64() - t0; }
    } else if (warp == 1 && PAIR && c == 1) {
        // (the peer CTA of a pair issues no MMAs: the leader's instructions drive both tensor cores)
    } else if (warp == 1) {
        // ---- MMA issuer (the only CTA in single mode, the leader of the pair in pair mode)
        const uint32_t idesc = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
        const bool tr = p.trace != nullptr;
        long long t_we = 0, t_wf = 0, t0 = tr ? clock64() : 0;
        if (t_first < t_end) {
            mbar_wait(&b_full, 0);
        }
        int s = 0;
        uint32_t ph = 0, buf = 0, aph = 0;
        const uint64_t adesc0 = make_sw128_desc(smem_u32(sA)), bdesc0 = make_sw128_desc(smem_u32(sB));
        const uint64_t a_step = (uint64_t)(WS_A_BYTES >> 4), b_step = (uint64_t)(b_stride >> 4), s_step = (uint64_t)(stage_bytes >> 4);
        const int k16_total = (p.K + 15) >> 4;                  // K = 16 steps that hold data: the zero-filled tail of the last block is skipped
        uint64_t adesc = adesc0;
        for (int tile = t_first; tile < t_end; tile += t_step) {
            const long long w0 = tr ? clock64() : 0;
            if (PAIR) ws_wait_cluster(&acc_empty[buf], aph ^ 1u);      // both CTAs' epilogues have drained this accumulator
            else ws_wait(&acc_empty[buf], aph ^ 1u);
            if (tr) t_we += clock64() - w0;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t tmem_d = tmem_base + buf * acc_cols;
            uint64_t bdesc = bdesc0;
            for (int kb = 0; kb < nkb; kb += SK) {
                const long long w1 = tr ? clock64() : 0;
                ws_wait(&full_bar[s], ph);
                if (tr) t_wf += clock64() - w1;
                if (tr && tile == t_first + 2 * t_step && kb < 8 && lane == 0) {
                    p.trace[4096 * 16 + blockIdx.x * 32 + 16 + kb] = w1;          // MMA warp started waiting for the block
                    p.trace[4096 * 16 + blockIdx.x * 32 + 24 + kb] = clock64();   // block (both CTAs') ready
                }
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                if (elect_one()) {
                    const int nblk = min(SK, nkb - kb);
#pragma unroll
                    for (int i = 0; i < WS_MAX_SK; ++i) {
#pragma unroll
                        for (int k = 0; k < WS_BK / 16; ++k) {
                            if (i < nblk && (kb + i) * (WS_BK / 16) + k < k16_total) {
                                const uint64_t ad = adesc + (uint64_t)i * a_step + (uint64_t)(k * 2), bd = bdesc + (uint64_t)i * b_step + (uint64_t)(k * 2);
                                if (PAIR) umma2_f16(tmem_d, ad, bd, idesc, (kb | i | k) ? 1u : 0u);
                                else umma_f16(tmem_d, ad, bd, idesc, (kb | i | k) ? 1u : 0u);
                            }
                        }
                    }
                    if (PAIR) umma2_commit_multicast(&empty_bar[s], mask_all);
                    else if (CS > 1) umma_commit_multicast(&empty_bar[s], mask_all);
                    else umma_commit(&empty_bar[s]);
                }
                __syncwarp();
                bdesc += (uint64_t)SK * b_step;
                adesc += s_step;
                if (++s == stages) { s = 0; ph ^= 1u; adesc = adesc0; }
            }
            if (elect_one()) {
                if (PAIR) umma2_commit_multicast(&acc_full[buf], mask_pair);
                else umma_commit(&acc_full[buf]);
            }
            __syncwarp();
            if (++buf == (uint32_t)NACC) { buf = 0; aph ^= 1u; }
        }
        if (tr && lane == 0) { p.trace[blockIdx.x * 16 + 2] = t_we; p.trace[blockIdx.x * 16 + 3] = t_wf; p.trace[blockIdx.x * 16 + 4] = clock64() - t0; }
    } else {
        // ---- epilogue.  Unit of work = 32 rows (my TMEM lane quarter) x 32 columns = one 2 KB box; the WS_NH warps of a
        // quarter take the boxes of a tile round-robin.  Per unit: (residual box arrives by TMA, requested one unit ahead)
        // -> two 16-column accumulator chunks through the fused math -> fp16 into the box, in place -> TMA store.
        constexpr bool has_bias = (SPEC & VFML_EPI_BIAS) != 0, has_res = (SPEC & VFML_EPI_RES) != 0, has_dw = (SPEC & VFML_EPI_DW1) != 0;
        constexpr bool do_gelu = (SPEC & VFML_EPI_GELU) != 0;
        const int ew = warp - 2;
        const int q = warp & 3;                    // TMEM lane quarter this warp may access
        const int h = ew >> 2;                     // position among the warps of the quarter
        long long e_wait = 0, e_res = 0, e_t0 = clock64();
        const bool tr = p.trace != nullptr;
        if (h < nbx) {                             // (inactive warps are not counted in acc_empty)
            uint8_t* my_boxes = sBox + (size_t)ew * WS_NBUF * WS_BOX_BYTES;
            uint64_t* my_bars = res_full + ew * WS_NBUF;
            const CUtensorMap* mapD = &maps.D[ct];
            // lane's row inside a box: 64 bytes, 16-byte chunk k at ((k ^ ((row >> 1) & 3)) << 4)   (CU_TENSOR_MAP_SWIZZLE_64B)
            const uint32_t row_off = (uint32_t)lane * 64u, sw = ((uint32_t)lane >> 1) & 3u;
            uint32_t u = 0;                        // units done by this warp: box u % 3, residual barrier parity (u / 3) & 1
            if (has_res && t_first < t_end && lane == 0) {
                mbar_expect_tx(&my_bars[0], WS_BOX_BYTES);
                tma_load_2d(&maps.R, &my_bars[0], my_boxes, n0 + h * WS_BOX_COLS, t_first * TM + c * WS_BM + q * 32);
            }
            uint32_t buf = 0, aph = 0;
            for (int tile = t_first; tile < t_end; tile += t_step) {
                const int m_row = tile * TM + c * WS_BM + q * 32;
                const uint32_t trow = tmem_base + buf * acc_cols + ((uint32_t)(q * 32) << 16);
                for (int j = h; j < nbx; j += WS_NH, ++u) {
                    const uint32_t b = u % WS_NBUF;
                    uint8_t* box = my_boxes + b * WS_BOX_BYTES;
                    if (has_res) {
                        // request the NEXT unit's residual: its box was the source of the store issued two units ago
                        int nj = j + WS_NH, ntile = tile;
                        if (nj >= nbx) { nj = h; ntile += t_step; }
                        if (ntile < t_end && lane == 0) {
                            bulk_wait_read<WS_NBUF - 2>();
                            const uint32_t nb = (u + 1) % WS_NBUF;
                            mbar_expect_tx(&my_bars[nb], WS_BOX_BYTES);
                            tma_load_2d(&maps.R, &my_bars[nb], my_boxes + nb * WS_BOX_BYTES, n0 + nj * WS_BOX_COLS, ntile * TM + c * WS_BM + q * 32);
                        }
                    } else {
                        if (lane == 0) bulk_wait_read<WS_NBUF - 1>();      // the store that last used this box has read it
                        __syncwarp();
                    }
                    if (j == h) {
                        const long long w0 = tr ? clock64() : 0;
                        mbar_wait(&acc_full[buf], aph);
                        if (tr) e_wait += clock64() - w0;
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    }
                    if (has_res) {
                        const long long w0 = tr ? clock64() : 0;
                        mbar_wait(&my_bars[b], (u / WS_NBUF) & 1u);
                        if (tr) e_res += clock64() - w0;
                    }
#pragma unroll
                    for (int cc = 0; cc < 2; ++cc) {
                        const int c = j * WS_BOX_COLS + cc * 16;            // column inside my BN-wide slice
                        if (c >= cols_valid) break;                         // warp-uniform; the box columns beyond are clipped by the store
#ifdef VFML_EXP_NOEPI
                        if (p.M != 12345) break;
#endif
                        uint32_t tc[16];
                        tmem_ld16_issue(trow + (uint32_t)c, tc);
                        uint4* p0 = reinterpret_cast<uint4*>(box + row_off + (((uint32_t)(2 * cc) ^ sw) << 4));
                        uint4* p1 = reinterpret_cast<uint4*>(box + row_off + (((uint32_t)(2 * cc + 1) ^ sw) << 4));
                        uint4 rc0 = make_uint4(0, 0, 0, 0), rc1 = rc0;
                        if (has_res) { rc0 = *p0; rc1 = *p1; }
                        tmem_ld_wait(tc);
                        float2 x[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) x[i] = make_float2(__uint_as_float(tc[2 * i]), __uint_as_float(tc[2 * i + 1]));
                        if (has_bias) {
                            const float4* b4 = reinterpret_cast<const float4*>(s_vec + c);
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const float4 bb = b4[i];
                                x[2 * i] = add2(x[2 * i], make_float2(bb.x, bb.y));
                                x[2 * i + 1] = add2(x[2 * i + 1], make_float2(bb.z, bb.w));
                            }
                        }
                        if (has_res) {
                            const __half2* h0 = reinterpret_cast<const __half2*>(&rc0);
                            const __half2* h1 = reinterpret_cast<const __half2*>(&rc1);
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                x[i] = add2(x[i], __half22float2(h0[i]));
                                x[4 + i] = add2(x[4 + i], __half22float2(h1[i]));
                            }
                        }
                        if (do_gelu) {
#pragma unroll
                            for (int i = 0; i < 8; ++i) x[i] = gelu_fast2(x[i]);
                        }
                        if (has_dw) {   // k = 1 depthwise stage: gelu(v + (w * v + b))
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const float2 w2 = *reinterpret_cast<const float2*>(s_vec + BN + c + 2 * i);
                                const float2 b2 = *reinterpret_cast<const float2*>(s_vec + 2 * BN + c + 2 * i);
                                x[i] = gelu_fast2(add2(x[i], fma2(w2, x[i], b2)));
                            }
                        }
                        uint4 o0, o1;
                        __half2* g0 = reinterpret_cast<__half2*>(&o0);
                        __half2* g1 = reinterpret_cast<__half2*>(&o1);
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            g0[i] = __float22half2_rn(x[i]);
                            g1[i] = __float22half2_rn(x[4 + i]);
                        }
                        *p0 = o0;
                        *p1 = o1;
                        // measured: a bulk tensor store clips at 16-byte granularity, so the map of this rank ends at the last
                        // multiple of 8 columns and the (< 8) columns after it are stored from registers
                        if ((cols_valid & 7) && c + 16 > (cols_valid & ~7)) {
                            const int m = m_row + lane;
                            if (m < p.M) {
                                __half* drow = p.D + (int64_t)m * p.ldd + n0;
                                const __half* o = (cols_valid & ~7) >= c + 8 ? reinterpret_cast<const __half*>(&o1) : reinterpret_cast<const __half*>(&o0);
                                const int t0c = cols_valid & ~7;         // first tail column (inside this chunk), a multiple of 8
                                for (int e = 0; e < (cols_valid & 7); ++e) drow[t0c + e] = o[e];
                            }
                        }
                    }
                    if (j + WS_NH >= nbx) {        // my last box of this tile: the accumulator goes back to the MMA warp
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                        __syncwarp();
                        if (lane == 0) {
                            if (PAIR && c == 1) mbar_arrive_remote(&acc_empty[buf], (uint32_t)leader);
                            else mbar_arrive(&acc_empty[buf]);
                        }
                    }
                    fence_proxy_async_smem();      // my shared-memory writes, visible to the bulk-copy engine
                    __syncwarp();
                    if (lane == 0) {
#ifdef VFML_EXP_NOSTORE
                        if (p.M == 12345)
#endif
                        if (j * WS_BOX_COLS < (cols_valid & ~7)) tma_store_2d(mapD, box, j * WS_BOX_COLS, m_row);
                        else asm volatile("cp.async.bulk.commit_group;" ::: "memory");       // keeps the group count per unit
                    }
                }
                if (++buf == (uint32_t)NACC) { buf = 0; aph ^= 1u; }
            }
            if (lane == 0) bulk_wait_all();
        }
        if (tr && warp == 2 && lane == 0) {
            p.trace[blockIdx.x * 16 + 5] = e_wait; p.trace[blockIdx.x * 16 + 6] = clock64() - e_t0; p.trace[blockIdx.x * 16 + 9] = e_res;
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync_all();        // no CTA leaves while a peer may still multicast into it or signal its barriers
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}


// ------------------------------------------------------------------------------------------ host
static int ws_map(CUtensorMap* map, const void* base, int rank, const cuuint64_t* dims, const cuuint64_t* strides, const cuuint32_t* box,
                  CUtensorMapSwizzle swz, const char* what) {
    tensor_map_encode_fn encode = tensor_map_encoder();
    if (!encode) {
        set_error("gemm_ws: cuTensorMapEncodeTiled is not available from the driver");
        return VFML_ERR_CUDA;
    }
    const cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, (cuuint32_t)rank, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("gemm_ws: cuTensorMapEncodeTiled(%s) failed: CUresult %d", what, (int)r);
        return VFML_ERR_CUDA;
    }
    return VFML_OK;
}

struct WsKernel { int spec; const void* fn[2]; };      // [single, pair]
#define WS_K(S) {S, {reinterpret_cast<const void*>(&gemm_ws_kernel<S, false>), reinterpret_cast<const void*>(&gemm_ws_kernel<S, true>)}}
static const WsKernel ws_kernels[] = {
    WS_K(0), WS_K(VFML_EPI_BIAS), WS_K(VFML_EPI_BIAS | VFML_EPI_GELU), WS_K(VFML_EPI_BIAS | VFML_EPI_RES),
    WS_K(VFML_EPI_BIAS | VFML_EPI_RES | VFML_EPI_GELU), WS_K(VFML_EPI_BIAS | VFML_EPI_RES | VFML_EPI_GELU | VFML_EPI_DW1)};
#undef WS_K
constexpr int WS_NUM_KERNELS = sizeof(ws_kernels) / sizeof(ws_kernels[0]);
constexpr size_t WS_SMEM_MAX = 227 * 1024 - 512;

static int ws_kernel_index(const vfml_gemm_desc& d) {
    for (int i = 0; i < WS_NUM_KERNELS; ++i)
        if ((d.epi & ~VFML_EPI_WEIGHTS) == (uint32_t)ws_kernels[i].spec) return i;
    return -1;
}

// does this descriptor fit the kernel's contract?  (fp16 in and out, one batch, alpha 1, one of the compiled epilogues,
// TMA-compatible bases and pitches for A, B, D and the residual)
bool gemm_ws_supported(const vfml_gemm_desc& d) {
    static const bool off = getenv("VFML_GEMM_NO_WS") != nullptr;
    if (off) return false;
    if (d.in_dtype != VFML_F16 || d.out_dtype != VFML_F16 || d.batch != 1 || d.alpha != 1.0f) return false;
    if (ws_kernel_index(d) < 0) return false;
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    if (!al16(d.A) || !al16(d.B) || !al16(d.D) || (d.lda % 8) || (d.ldb % 8) || (d.ldd % 8)) return false;
    if ((d.epi & VFML_EPI_RES) && (!d.res || !al16(d.res) || (d.ldr % 8))) return false;
    if ((d.epi & VFML_EPI_BIAS) && !d.bias) return false;
    if ((d.epi & VFML_EPI_DW1) && (!d.dw_w || !d.dw_b)) return false;
    return d.K >= 1 && d.M >= 4 * WS_BM && d.N >= 16;
}

struct WsConfig { int pair, cs, bn, stages, sk, clusters; size_t smem; double cost; };

// > 0: declined (no configuration whose weight slice fits shared memory): the caller falls back to gemm_tc
int gemm_ws(const vfml_gemm_desc& d, cudaStream_t st, int mode, int cluster) {      // mode: -1 = cost model (or VFML_WS_PAIR), 0 = single CTAs, 1 = CTA pairs; cluster: 0 = cost model (or VFML_WS_CS)
    const int ki = ws_kernel_index(d);
    static PerDevice<bool> attr_done;
    if (!attr_done.get()) {
        for (int i = 0; i < WS_NUM_KERNELS; ++i)
            for (int m = 0; m < 2; ++m)
                VFML_CUDA(cudaFuncSetAttribute(ws_kernels[i].fn[m], cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WS_SMEM_MAX));
        attr_done.get() = true;
    }
    const int nkb = cdiv(d.K, WS_BK), nk16 = cdiv(d.K, 16);
    const bool has_res = (d.epi & VFML_EPI_RES) != 0;
    // experiments / tests: pin the mode and the cluster size (read per call)
    const char* e_cs = getenv("VFML_WS_CS");
    const char* e_pair = getenv("VFML_WS_PAIR");
    const int force_cs = cluster > 0 ? cluster : e_cs ? atoi(e_cs) : 0, force_pair = mode >= 0 ? mode : e_pair ? atoi(e_pair) : -1;
    // co-resident clusters per (kernel, mode, cluster size): GPC boundaries leave some SMs out for the larger sizes
    static PerDevice<int[WS_NUM_KERNELS][2][WS_MAX_CS + 1]> occ_cache;
    WsConfig best{};
    best.cost = 1e30;
    for (int pair = 0; pair < 2; ++pair) {
        if (force_pair >= 0 && pair != force_pair) continue;
        for (int cs = pair ? 2 : 1; cs <= WS_MAX_CS; cs += pair ? 2 : 1) {
            if (force_cs && cs != force_cs) continue;
            const int n_ct = pair ? cs / 2 : cs;                                 // column tiles = CTAs sharing A rows
            const int bn = (cdiv(d.N, n_ct) + 15) / 16 * 16;
            if (bn > 256 || (n_ct - 1) * bn >= d.N) continue;
            const int bh = pair ? bn / 2 : bn;
            const int sk = pair ? 1 : 2;
            const size_t b_stride = ((size_t)bh * WS_BK * 2 + 1023) & ~(size_t)1023, stage_bytes = (size_t)sk * WS_A_BYTES;
            const size_t fixed = 1024 + nkb * b_stride + (size_t)WS_EPI_WARPS * WS_NBUF * WS_BOX_BYTES + (size_t)3 * bn * sizeof(float);
            if (fixed + (pair ? 3 : 2) * stage_bytes > WS_SMEM_MAX) continue;
            const int stages = (int)std::min<size_t>(std::min<size_t>(WS_MAX_STAGES, (size_t)cdiv(nkb, sk) * 2), (WS_SMEM_MAX - fixed) / stage_bytes);
            const size_t smem = fixed + (size_t)stages * stage_bytes;
            int& occ = occ_cache.get()[ki][pair][cs];
            if (occ == 0) {
                cudaLaunchConfig_t q{};
                q.gridDim = dim3(cs * 64); q.blockDim = dim3(WS_THREADS); q.dynamicSmemBytes = WS_SMEM_MAX;
                cudaLaunchAttribute qa[1];
                qa[0].id = cudaLaunchAttributeClusterDimension;
                qa[0].val.clusterDim.x = cs; qa[0].val.clusterDim.y = 1; qa[0].val.clusterDim.z = 1;
                q.attrs = qa; q.numAttrs = 1;
                int n = 0;
                if (cudaOccupancyMaxActiveClusters(&n, ws_kernels[ki].fn[pair], &q) != cudaSuccess || n <= 0) {
                    cudaGetLastError();
                    occ = -1;                      // this cluster size cannot be scheduled on this device
                } else {
                    occ = std::max(1, std::min(n, device_sms() / cs));
                }
            }
            if (occ < 0) continue;
            const int row_tiles = cdiv(d.M, pair ? 2 * WS_BM : WS_BM);
            const int clusters = std::min(occ, row_tiles);
            // cycles per CTA and tile: the slowest of the tensor pipe, the shared-memory ports (~100 B/clk) and the SM's L2 link
            // (~30 B/clk with every SM busy); see the header
            const double kp = nkb * 64.0;
            const double mma = nk16 * std::max(bn / 2.0, pair ? 43.0 : 76.0) + 300.0;
            const double smem_bytes = 2.0 * 128 * kp * 2 + (double)bh * kp * 2 + 128.0 * bn * 2 * (has_res ? 4 : 2);
            const double l2_bytes = 128.0 * kp * 2 / n_ct + 128.0 * bn * 2 * (has_res ? 2 : 1);
            const double tile = std::max(mma, std::max(smem_bytes / 100.0, l2_bytes / 30.0));
            const double cost = (double)cdiv(row_tiles, clusters) * tile;
            if (cost < best.cost) best = WsConfig{pair, cs, bn, stages, sk, clusters, smem, cost};
        }
    }
    if (!best.cs) return 1;
    const int CS = best.cs, BN = best.bn, PAIRM = best.pair;
    const int n_ct = PAIRM ? CS / 2 : CS;
    const void* fn = ws_kernels[ki].fn[PAIRM];
    WsMaps maps;
    {
        const cuuint64_t dims[3] = {(cuuint64_t)d.K, (cuuint64_t)d.M, 1}, strides[2] = {(cuuint64_t)d.lda * 2, (cuuint64_t)d.lda * d.M * 2};
        int loaders = 1;
        while (loaders * 2 <= n_ct) loaders *= 2;
        const cuuint32_t box[3] = {(cuuint32_t)WS_BK, (cuuint32_t)(WS_BM / loaders), 1};
        int rc = ws_map(&maps.A, d.A, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B, "A");
        if (rc) return rc;
    }
    {
        const cuuint64_t dims[3] = {(cuuint64_t)d.K, (cuuint64_t)d.N, 1}, strides[2] = {(cuuint64_t)d.ldb * 2, (cuuint64_t)d.ldb * d.N * 2};
        const cuuint32_t box[3] = {(cuuint32_t)WS_BK, (cuuint32_t)(PAIRM ? BN / 2 : BN), 1};
        int rc = ws_map(&maps.B, d.B, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B, "B");
        if (rc) return rc;
    }
    const cuuint32_t box2[2] = {(cuuint32_t)WS_BOX_COLS, 32};
    if (has_res) {
        const cuuint64_t dims[2] = {(cuuint64_t)d.N, (cuuint64_t)d.M}, strides[1] = {(cuuint64_t)d.ldr * 2};
        int rc = ws_map(&maps.R, d.res, 2, dims, strides, box2, CU_TENSOR_MAP_SWIZZLE_64B, "residual");
        if (rc) return rc;
    } else {
        maps.R = maps.A;
    }
    for (int t = 0; t < WS_MAX_CS; ++t) {
        maps.D[t] = maps.A;                                     // never used beyond the column tiles (and not for widths < 8)
        if (t >= n_ct) continue;
        const int width = std::min(BN, d.N - t * BN) & ~7;      // whole 16-byte units only; the kernel stores the ragged tail itself
        if (width == 0) continue;
        const cuuint64_t dims[2] = {(cuuint64_t)width, (cuuint64_t)d.M}, strides[1] = {(cuuint64_t)d.ldd * 2};
        int rc = ws_map(&maps.D[t], reinterpret_cast<const __half*>(d.D) + (size_t)t * BN, 2, dims, strides, box2, CU_TENSOR_MAP_SWIZZLE_64B, "D");
        if (rc) return rc;
    }
    WsParams p{};
    p.bias = d.bias; p.dw_w = d.dw_w; p.dw_b = d.dw_b;
    p.D = reinterpret_cast<__half*>(d.D); p.ldd = d.ldd;
    p.M = d.M; p.N = d.N; p.K = d.K;
    p.BN = BN; p.stages = best.stages; p.sk = best.sk; p.row_tiles = cdiv(d.M, PAIRM ? 2 * WS_BM : WS_BM); p.cs = CS; p.n_clusters = best.clusters;
    p.nacc = BN <= 128 ? 4 : 2;
    p.epi = d.epi;
    p.b_is_const = (d.epi & VFML_EPI_WEIGHTS) != 0;
    static const bool trace = getenv("VFML_GEMM_TRACE") != nullptr;
    constexpr int TRACE_MAX_CTAS = 1024;
    static PerDevice<long long*> trace_bufs;
    long long*& trace_buf = trace_bufs.get();
    const int grid = best.clusters * CS;
    if (trace) {
        if (!trace_buf) VFML_CUDA(cudaMalloc(&trace_buf, (4096 * 16 + TRACE_MAX_CTAS * 32) * sizeof(long long)));
        VFML_CUDA(cudaMemsetAsync(trace_buf, 0, (4096 * 16 + TRACE_MAX_CTAS * 32) * sizeof(long long), st));
        p.trace = trace_buf;
    }
    static const bool no_pdl = getenv("VFML_NO_PDL") != nullptr;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(WS_THREADS);
    cfg.dynamicSmemBytes = best.smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CS; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = no_pdl ? 1 : 2;
    void* args[2] = {&maps, &p};
    cudaError_t le = cudaLaunchKernelExC(&cfg, fn, args);
    if (le != cudaSuccess) {
        set_error("gemm_ws: launch failed: %s (pair=%d CS=%d BN=%d stages=%d smem=%zu)", cudaGetErrorString(le), PAIRM, CS, BN, best.stages, best.smem);
        cudaGetLastError();
        return VFML_ERR_CUDA;
    }
    VFML_LAUNCH_CHECK("gemm_ws");
    if (trace) {      // debug only: synchronises
        static long long hbuf[TRACE_MAX_CTAS * 16];
        VFML_CUDA(cudaMemcpyAsync(hbuf, trace_buf, (size_t)grid * 16 * sizeof(long long), cudaMemcpyDeviceToHost, st));
        VFML_CUDA(cudaStreamSynchronize(st));
        // producer and epilogue counters: every CTA; MMA counters: the issuing CTAs only
        double a[16] = {0};
        int n_mma = 0;
        for (int cta = 0; cta < grid; ++cta) {
            for (int i = 0; i < 16; ++i)
                if (i < 2 || i > 4) a[i] += (double)hbuf[cta * 16 + i] / grid;
            if (hbuf[cta * 16 + 4]) { ++n_mma; for (int i = 2; i <= 4; ++i) a[i] += (double)hbuf[cta * 16 + i]; }
        }
        for (int i = 2; i <= 4; ++i) a[i] /= std::max(1, n_mma);
        if (getenv("VFML_WS_TIMELINE")) {
            static long long tl[TRACE_MAX_CTAS * 32];
            VFML_CUDA(cudaMemcpy(tl, trace_buf + 4096 * 16, (size_t)grid * 32 * sizeof(long long), cudaMemcpyDeviceToHost));
            for (int cta = 0; cta < std::min(grid, 2 * CS); cta += 1) {
                const long long* t = tl + cta * 32;
                const long long base = t[0];
                fprintf(stderr, "  timeline CTA %d (3rd tile), cycles from the producer's first slot wait:", cta);
                for (int kb = 0; kb < std::min(8, nkb); ++kb)
                    fprintf(stderr, "  kb%d[slot wait %lld free %lld | mma wait %lld ready %lld]", kb, t[kb] - base, t[8 + kb] - base, t[16 + kb] ? t[16 + kb] - base : -1,
                            t[24 + kb] ? t[24 + kb] - base : -1);
                fprintf(stderr, "\n");
            }
        }
        fprintf(stderr, "[gemm_ws trace] M=%d N=%d K=%d pair=%d CS=%d BN=%d stages=%dx%d clusters=%d nacc=%d epi=%u | producer wait_empty %.0f of %.0f | mma wait_acc_empty %.0f wait_full %.0f of %.0f | epilogue(warp2) wait_acc_full %.0f wait_res %.0f of %.0f cycles (avg per CTA)\n",
                d.M, d.N, d.K, PAIRM, CS, BN, best.stages, best.sk, best.clusters, p.nacc, d.epi, a[0], a[1], a[2], a[3], a[4], a[5], a[9], a[6]);
    }
    return VFML_OK;
}

// ---- which kernel runs a layer: measured, not modelled.  vfml_gemm_autotune() times the general kernel (gemm_tc) and every
// feasible (mode, cluster size) of this one on the caller's descriptor and remembers the winner per (device, M, N, K,
// epilogue); vfml_gemm() looks it up.  Layers that were never tuned run on gemm_tc, as before.
using WsKey = std::tuple<int, int, int, int, uint32_t>;
static std::mutex ws_mu;
static std::map<WsKey, int> ws_choice;            // 0 = gemm_tc, else 1 + 16 * pair + cluster size
static WsKey ws_key(const vfml_gemm_desc& d) { return WsKey(device_index(), d.M, d.N, d.K, d.epi & ~VFML_EPI_WEIGHTS); }

int gemm_ws_choice(const vfml_gemm_desc& d) {
    if (getenv("VFML_WS_PAIR") || getenv("VFML_WS_CS")) return -1;       // experiments / tests pin the kernel through the environment
    std::lock_guard<std::mutex> lk(ws_mu);
    auto it = ws_choice.find(ws_key(d));
    return it == ws_choice.end() ? 0 : it->second;
}

int gemm_ws_run_choice(const vfml_gemm_desc& d, cudaStream_t st, int choice) {
    if (choice < 0) return gemm_ws(d, st, -1, 0);
    return gemm_ws(d, st, ((choice - 1) >> 4) & 1, (choice - 1) & 15);
}

int gemm_tc(const vfml_gemm_desc& d, cudaStream_t st);
bool gemm_tc_supported(const vfml_gemm_desc& d);

int gemm_autotune(const vfml_gemm_desc& d, cudaStream_t st) {
    if (!gemm_ws_supported(d) || !gemm_tc_supported(d)) return 0;
    {
        std::lock_guard<std::mutex> lk(ws_mu);
        auto it = ws_choice.find(ws_key(d));
        if (it != ws_choice.end()) return it->second;
    }
    // the candidates are run repeatedly on the caller's buffers: only safe when the output does not feed back into the inputs
    auto overlaps = [](const void* a, size_t na, const void* b, size_t nb) {
        const uintptr_t x = reinterpret_cast<uintptr_t>(a), y = reinterpret_cast<uintptr_t>(b);
        return x < y + nb && y < x + na;
    };
    const size_t d_bytes = (size_t)d.M * d.ldd * 2;
    if (overlaps(d.D, d_bytes, d.A, (size_t)d.M * d.lda * 2) || (d.res && overlaps(d.D, d_bytes, d.res, (size_t)d.M * d.ldr * 2))) return 0;
    cudaEvent_t e0, e1;
    VFML_CUDA(cudaEventCreate(&e0));
    VFML_CUDA(cudaEventCreate(&e1));
    float best_ms = 1e30f, tc_ms = 1e30f;
    int best = 0;
    for (int cand = 0; cand <= 32; ++cand) {       // 0: gemm_tc; 1 + 16 pair + cs
        const int pair = cand ? ((cand - 1) >> 4) & 1 : 0, cs = cand ? (cand - 1) & 15 : 0;
        if (cand && (cs < 1 || cs > WS_MAX_CS || (pair && (cs & 1)))) continue;
        auto run = [&]() { return cand == 0 ? gemm_tc(d, st) : gemm_ws(d, st, pair, cs); };
        int rc = 0;
        for (int i = 0; i < 2 && rc == 0; ++i) rc = run();
        if (rc > 0) continue;                      // this mode has no configuration for the shape
        if (rc < 0) { cudaEventDestroy(e0); cudaEventDestroy(e1); return rc; }
        float ms_min = 1e30f;
        for (int rep = 0; rep < 3; ++rep) {        // best of three groups of five back-to-back launches
            VFML_CUDA(cudaEventRecord(e0, st));
            for (int i = 0; i < 5; ++i) run();
            VFML_CUDA(cudaEventRecord(e1, st));
            VFML_CUDA(cudaEventSynchronize(e1));
            float ms = 0;
            VFML_CUDA(cudaEventElapsedTime(&ms, e0, e1));
            ms_min = std::min(ms_min, ms);
        }
        if (cand == 0) tc_ms = ms_min;
        if (ms_min < best_ms) { best_ms = ms_min; best = cand; }
    }
    if (best != 0 && best_ms > 0.97f * tc_ms) best = 0;       // switch only for a clear win
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (getenv("VFML_GEMM_AUTOTUNE_LOG")) {
        if (best) fprintf(stderr, "[gemm autotune] M=%d N=%d K=%d epi=%u: gemm_tc %.1f us, best gemm_ws %s cluster %d: %.1f us\n", d.M, d.N, d.K, d.epi, tc_ms * 200,
                          ((best - 1) >> 4) & 1 ? "pair" : "single", (best - 1) & 15, best_ms * 200);
        else fprintf(stderr, "[gemm autotune] M=%d N=%d K=%d epi=%u: gemm_tc %.1f us stays\n", d.M, d.N, d.K, d.epi, tc_ms * 200);
    }
    std::lock_guard<std::mutex> lk(ws_mu);
    ws_choice[ws_key(d)] = best;
    return best;
}

}  // namespace vfml

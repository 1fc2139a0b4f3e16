// Micro-benchmark: how fast can one SM write (and read-modify-write) a GEMM output tile, per access pattern?
// 148 persistent CTAs x 384 threads walk 128 x 192 fp16 tiles of a [M, 488] matrix (the 42240 x 486 layer of config C1):
//   mode 0  lane-per-row:   each lane stores 32 B of ITS row per instruction (the V1 GEMM epilogue: 32 lines per instruction)
//   mode 1  row-segments:   a warp instruction covers 4 rows x 128 B, 16 B per lane (the V2 epilogue's idea, without its shuffles)
//   mode 2  TMA store:      warps write their 32 x 64 box into 128B-swizzled shared memory, one lane issues cp.async.bulk.tensor
//   mode 3..5               the same three with a residual READ of the tile first (ld / ld / TMA load into the same box)
// Prints bytes per clock per SM of tile payload (for 3..5: read + written).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/store_probe scripts/micro/store_probe.cu
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

constexpr int BM = 128, BN = 192, NW = 12, THREADS = NW * 32;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void st256(void* ptr, uint4 a, uint4 b) {
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(ptr), "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w) : "memory");
}
__device__ __forceinline__ void ld256(const void* ptr, uint4& a, uint4& b) {
    asm volatile("ld.global.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w) : "l"(ptr));
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(THREADS, 1)
probe(const __grid_constant__ CUtensorMap mapD, const __grid_constant__ CUtensorMap mapR, __half* D, const __half* R, int M, int N, int ld,
      int passes, long long* cycles) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bars[NW][2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3, hsel = warp >> 2;                   // row quarter, 64-column share
    constexpr bool RES = MODE >= 3;
    constexpr int PAT = MODE % 3;
    const int m_tiles = M / BM, n_tiles = (N + BN - 1) / BN;
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[warp][0])));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[warp][1])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint8_t* box[2] = {smem + (size_t)warp * 8192, smem + (size_t)warp * 8192 + 4096};
    const long long t0 = clock64();
    uint32_t it = 0;
    uint32_t acc = 0;
    for (int pass = 0; pass < passes; ++pass)
        for (int tile = blockIdx.x; tile < m_tiles * n_tiles; tile += gridDim.x) {
            const int m0 = (tile / n_tiles) * BM + q * 32, n0 = (tile % n_tiles) * BN + hsel * 64;
            if (n0 >= N) continue;
            const uint32_t b = it & 1u;
            if (PAT == 0) {
                __half* row = D + (size_t)(m0 + lane) * ld + n0;
                const __half* rrow = R + (size_t)(m0 + lane) * ld + n0;
#pragma unroll
                for (int c = 0; c < 64; c += 16) {
                    if (n0 + c + 16 > N) break;
                    uint4 a = make_uint4(tile, c, lane, pass), bb = a;
                    const bool al = (reinterpret_cast<uintptr_t>(row + c) & 31) == 0;      // every other row with pitch 976 B
                    if (RES) {
                        uint4 x, y;
                        if (al) ld256(rrow + c, x, y);
                        else { x = *reinterpret_cast<const uint4*>(rrow + c); y = *reinterpret_cast<const uint4*>(rrow + c + 8); }
                        a.x += x.x ^ y.w; bb.y += x.y ^ y.z;
                    }
                    if (al) st256(row + c, a, bb);
                    else { *reinterpret_cast<uint4*>(row + c) = a; *reinterpret_cast<uint4*>(row + c + 8) = bb; }
                }
            } else if (PAT == 1) {
                // instruction i covers rows 4i .. 4i+3: 8 lanes x 16 B per 128-byte row segment
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int r = m0 + 4 * i + (lane >> 3), c = (lane & 7) * 8;
                    if (n0 + c + 8 > N) continue;
                    uint4 a = make_uint4(tile, c, lane, pass);
                    if (RES) { const uint4 x = *reinterpret_cast<const uint4*>(R + (size_t)r * ld + n0 + c); a.x += x.x ^ x.w; }
                    *reinterpret_cast<uint4*>(D + (size_t)r * ld + n0 + c) = a;
                }
            } else {
                // box b was last stored from two tiles ago: wait until all but the most recent store have read their source
                if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                __syncwarp();
                if (RES) {
                    if (lane == 0) {
                        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bars[warp][b])), "r"(4096u) : "memory");
                        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                                     ::"r"(smem_u32(box[b])), "l"(&mapR), "r"(smem_u32(&bars[warp][b])), "r"(n0), "r"(m0) : "memory");
                    }
                    mbar_wait(&bars[warp][b], (it >> 1) & 1u);
                }
#pragma unroll
                for (int c = 0; c < 4; ++c) {       // 16 columns = two 16-byte chunks of this lane's row, swizzled
                    uint4 a = make_uint4(tile, c, lane, pass), bb = a;
                    uint4* p0 = reinterpret_cast<uint4*>(box[b] + lane * 128 + (((2 * c) ^ (lane & 7)) << 4));
                    uint4* p1 = reinterpret_cast<uint4*>(box[b] + lane * 128 + (((2 * c + 1) ^ (lane & 7)) << 4));
                    if (RES) { const uint4 x = *p0, y = *p1; a.x += x.x ^ y.w; bb.y += x.y ^ y.z; }
                    *p0 = a; *p1 = bb;
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) {
                    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                                 ::"l"(&mapD), "r"(smem_u32(box[b])), "r"(n0), "r"(m0) : "memory");
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                }
            }
            ++it;
        }
    if (PAT == 2 && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0 + (acc == 0x12345 ? 1 : 0);
}

// mode 6: TMA loads only, each warp keeps two 32 x 64 boxes in flight (96 KB per SM): the L2 -> SM read ceiling of this access shape.
// mode 7: TMA load one tile ahead (three boxes per warp) / in-place update / TMA store: read and write streams both kept busy.
template <int MODE>
__global__ void __launch_bounds__(THREADS, 1)
probe_pipe(const __grid_constant__ CUtensorMap mapD, const __grid_constant__ CUtensorMap mapR, int M, int N, int passes, long long* cycles) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bars[NW][3];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3, hsel = warp >> 2;
    constexpr int NB = MODE == 6 ? 2 : 3;
    const int m_tiles = M / BM, n_tiles = (N + BN - 1) / BN;
    if (lane == 0) {
        for (int i = 0; i < 3; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[warp][i])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint8_t* base = smem + (size_t)warp * 3 * 4096;
    const int total = m_tiles * n_tiles;
    // this warp's list of work: (pass, tile) pairs with a valid column share, walked with a running index
    auto coords = [&](long long idx, int& m0, int& n0) -> bool {      // idx-th tile of this CTA over all passes
        const long long per_pass = (total - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
        if (idx >= per_pass * passes) return false;
        const int tile = (int)blockIdx.x + (int)(idx % per_pass) * (int)gridDim.x;
        m0 = (tile / n_tiles) * BM + q * 32; n0 = (tile % n_tiles) * BN + hsel * 64;
        return true;
    };
    auto issue_load = [&](long long idx) {
        int m0, n0;
        if (!coords(idx, m0, n0)) return;
        const int b = (int)(idx % NB);
        if (lane == 0) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bars[warp][b])), "r"(4096u) : "memory");
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                         ::"r"(smem_u32(base + b * 4096)), "l"(&mapR), "r"(smem_u32(&bars[warp][b])), "r"(n0 < N ? n0 : 0), "r"(m0) : "memory");
        }
    };
    const long long t0 = clock64();
    uint32_t acc = 0;
    if (MODE == 6) { issue_load(0); issue_load(1); } else issue_load(0);
    for (long long idx = 0;; ++idx) {
        int m0, n0;
        if (!coords(idx, m0, n0)) break;
        const int b = (int)(idx % NB);
        if (MODE == 7) {
            // box (idx + 1) % 3 was the source of the store issued two tiles ago
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            __syncwarp();
            issue_load(idx + 1);
        }
        mbar_wait(&bars[warp][b], (uint32_t)((idx / NB) & 1));
        uint8_t* box = base + b * 4096;
        if (MODE == 6) {
            acc += *reinterpret_cast<const uint32_t*>(box + lane * 128);
            __syncwarp();
            issue_load(idx + 2);
        } else {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint4* p0 = reinterpret_cast<uint4*>(box + lane * 128 + (((2 * c) ^ (lane & 7)) << 4));
                uint4* p1 = reinterpret_cast<uint4*>(box + lane * 128 + (((2 * c + 1) ^ (lane & 7)) << 4));
                uint4 x = *p0, y = *p1;
                x.x += y.w; y.y ^= x.z;
                *p0 = x; *p1 = y;
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0 && n0 < N) {
                asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                             ::"l"(&mapD), "r"(smem_u32(box)), "r"(n0), "r"(m0) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0 + (acc == 0x12345 ? 1 : 0);
}

static CUtensorMap make_map(void* base, int N, int M, int ld) {
    typedef CUresult (*fn_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                             const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    CUtensorMap m;
    cuuint64_t dims[2] = {(cuuint64_t)N, (cuuint64_t)M}, strides[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {64, 32}, es[2] = {1, 1};
    CUresult r = reinterpret_cast<fn_t>(fn)(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(1); }
    return m;
}

template <int MODE>
static void run(const char* name, __half* D, __half* R, int M, int N, int ld, int sms, long long* cyc) {
    const CUtensorMap mD = make_map(D, N, M, ld), mR = make_map(R, N, M, ld);
    const int smem = NW * 8192 + 1024, passes = 8;
    cudaFuncSetAttribute(probe<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<MODE><<<sms, THREADS, smem>>>(mD, mR, D, R, M, N, ld, 1, cyc);
    cudaEventRecord(e0);
    probe<MODE><<<sms, THREADS, smem>>>(mD, mR, D, R, M, N, ld, passes, cyc);
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(err)); exit(1); }
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    long long* h = (long long*)malloc(sms * sizeof(long long));
    cudaMemcpy(h, cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost);
    double mx = 0;
    for (int i = 0; i < sms; ++i) mx = h[i] > mx ? h[i] : mx;
    const double bytes = (double)passes * M * N * 2 * (MODE >= 3 ? 2 : 1);
    printf("%-44s %8.1f us/pass  %6.1f B/clk/SM  %6.2f TB/s\n", name, ms * 1e3 / passes, bytes / sms / mx, bytes / (ms * 1e-3) / 1e12);
    free(h);
}

template <int MODE>
static void run_pipe(const char* name, __half* D, __half* R, int M, int N, int ld, int sms, long long* cyc) {
    const CUtensorMap mD = make_map(D, N, M, ld), mR = make_map(R, N, M, ld);
    const int smem = NW * 3 * 4096 + 1024, passes = 8;
    cudaFuncSetAttribute(probe_pipe<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe_pipe<MODE><<<sms, THREADS, smem>>>(mD, mR, M, N, 1, cyc);
    cudaEventRecord(e0);
    probe_pipe<MODE><<<sms, THREADS, smem>>>(mD, mR, M, N, passes, cyc);
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(err)); exit(1); }
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    long long* h = (long long*)malloc(sms * sizeof(long long));
    cudaMemcpy(h, cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost);
    double mx = 0;
    for (int i = 0; i < sms; ++i) mx = h[i] > mx ? h[i] : mx;
    // payload: every tile moves 128 x 192 x 2 B per direction (the third column tile's out-of-range share is still fetched as a box)
    const double tiles = (double)passes * (M / BM) * ((N + BN - 1) / BN);
    const double bytes = tiles * BM * BN * 2 * (MODE == 7 ? 2 : 1);
    printf("%-44s %8.1f us/pass  %6.1f B/clk/SM  %6.2f TB/s  (clock %.0f MHz)\n", name, ms * 1e3 / passes, bytes / sms / mx, bytes / (ms * 1e-3) / 1e12,
           mx / (ms * 1e3));
    free(h);
}

int main(int argc, char** argv) {
    const int M = 42240, N = 486, ld = 488;
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    if (argc > 1) sms = atoi(argv[1]);          // fewer CTAs: is the limit per SM or chip-wide?
    __half *D, *R;
    long long* cyc;
    cudaMalloc(&D, (size_t)M * ld * 2); cudaMalloc(&R, (size_t)M * ld * 2); cudaMalloc(&cyc, 1024 * 8);
    cudaMemset(D, 0, (size_t)M * ld * 2); cudaMemset(R, 0, (size_t)M * ld * 2);
    printf("tile walk over %d x %d fp16 (pitch %d), %d CTAs x %d threads\n", M, N, ld, sms, THREADS);
    run<0>("store, lane-per-row 32 B (V1 epilogue)", D, R, M, N, ld, sms, cyc);
    run<1>("store, 4 rows x 128 B per instruction", D, R, M, N, ld, sms, cyc);
    run<2>("store, smem box + TMA store", D, R, M, N, ld, sms, cyc);
    run<3>("read + store, lane-per-row 32 B", D, R, M, N, ld, sms, cyc);
    run<4>("read + store, 4 rows x 128 B", D, R, M, N, ld, sms, cyc);
    run<5>("read + store, TMA load / in place / TMA store", D, R, M, N, ld, sms, cyc);
    run_pipe<6>("TMA loads only, 2 boxes in flight per warp", D, R, M, N, ld, sms, cyc);
    run_pipe<7>("TMA load 1 tile ahead / in place / TMA store", D, R, M, N, ld, sms, cyc);
    return 0;
}

// How fast can the GEMM's A operand be streamed from HBM, and does its layout matter?
// The GMA aggregate reads one 7040 x 7040 fp16 attention matrix per centre frame and iteration (99 MB, three of them: they
// do not fit L2) through TMA boxes of 128 rows x 64 columns: 128 separate 128-byte pieces, 14 KB apart.  This probe streams
// the same bytes (3 matrices) with 148 persistent CTAs, 6 boxes of 16 KB in flight each, two ways:
//   mode 0  row-major matrix, box = 64 columns x 128 rows (the GEMM's access pattern)
//   mode 1  the same matrix stored as contiguous 16 KB blocks [row tile][k block][128][64], box = one block
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/hbm_stream_probe scripts/micro/hbm_stream_probe.cu
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

constexpr int STAGES = 6, BOX_BYTES = 128 * 64 * 2;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

// tiles walked like the split-K aggregate: CTA = (row tile, K split); consecutive boxes of a CTA = consecutive k blocks
__global__ void __launch_bounds__(64, 1) probe(const __grid_constant__ CUtensorMap map, const __grid_constant__ CUtensorMap mapB, int mode, int with_b, int row_tiles, int kblocks, int splits, uint32_t* sink) {
    extern __shared__ __align__(1024) uint8_t raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t full[STAGES];
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&full[s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    const int units = row_tiles * splits;
    const int kb_per = (kblocks + splits - 1) / splits;
    uint32_t acc = 0;
    long long issued = 0, done = 0;
    // flattened list of this CTA's boxes
    auto box_of = [&](long long i, int& rt, int& kb) -> bool {
        const long long per_unit = kb_per;
        const long long u = i / per_unit;
        const int unit = (int)blockIdx.x + (int)u * (int)gridDim.x;
        if (unit >= units) return false;
        rt = unit / splits;
        kb = (unit % splits) * kb_per + (int)(i % per_unit);
        return true;
    };
    auto issue = [&](long long i) {
        int rt, kb;
        if (!box_of(i, rt, kb)) return false;
        if (kb >= kblocks) kb = kblocks - 1;
        const int s = (int)(i % STAGES);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&full[s])), "r"((uint32_t)BOX_BYTES * (with_b ? 2u : 1u)) : "memory");
        if (with_b)      // the GEMM's other operand: the same 128 x 64 block of a small (L2-resident) matrix for every row tile
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                         ::"r"(smem_u32(smem + (STAGES + s) * BOX_BYTES)), "l"(&mapB), "r"(smem_u32(&full[s])), "r"(kb * 64), "r"(0) : "memory");
        if (mode == 0)
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                         ::"r"(smem_u32(smem + s * BOX_BYTES)), "l"(&map), "r"(smem_u32(&full[s])), "r"(kb * 64), "r"(rt * 128) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                         ::"r"(smem_u32(smem + s * BOX_BYTES)), "l"(&map), "r"(smem_u32(&full[s])), "r"(0), "r"((rt * kblocks + kb) * 128) : "memory");
        return true;
    };
    for (int i = 0; i < STAGES; ++i) if (issue(issued)) ++issued;
    while (done < issued) {
        const int s = (int)(done % STAGES);
        mbar_wait(&full[s], (uint32_t)((done / STAGES) & 1));
        acc += *reinterpret_cast<volatile uint32_t*>(smem + s * BOX_BYTES);
        ++done;
        if (issue(issued)) ++issued;
    }
    if (acc == 0x12345u) sink[0] = acc;
}

static CUtensorMap make_map(void* base, uint64_t inner, uint64_t rows, uint64_t pitch_bytes, uint32_t box_inner, uint32_t box_rows) {
    typedef CUresult (*fn_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                             const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    CUtensorMap m;
    cuuint64_t dims[2] = {inner, rows}, strides[1] = {pitch_bytes};
    cuuint32_t box[2] = {box_inner, box_rows}, es[2] = {1, 1};
    CUresult r = reinterpret_cast<fn_t>(fn)(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(1); }
    return m;
}

int main() {
    const int P = 7040, NM = 3, row_tiles = NM * (P / 128), kblocks = P / 64;      // 3 matrices stacked: 165 row tiles x 110 k blocks
    const size_t bytes = (size_t)NM * P * P * 2;
    void* buf;
    uint32_t* sink;
    cudaMalloc(&buf, bytes);
    cudaMalloc(&sink, 4);
    cudaMemset(buf, 1, bytes);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * STAGES * BOX_BYTES + 1024);
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const CUtensorMap m_rows = make_map(buf, P, (uint64_t)NM * P, (uint64_t)P * 2, 64, 128);
    const CUtensorMap m_blk = make_map(buf, 64, (uint64_t)row_tiles * kblocks * 128, 128, 64, 128);
    void* bufB;
    cudaMalloc(&bufB, (size_t)128 * P * 2);
    cudaMemset(bufB, 1, (size_t)128 * P * 2);
    const CUtensorMap m_b = make_map(bufB, P, 128, (uint64_t)P * 2, 64, 128);
    printf("%d matrices of %d x %d fp16 = %.0f MB, %d SMs, %d x 16 KB boxes in flight per CTA\n", NM, P, P, bytes / 1e6, sms, STAGES);
    for (int splits : {2, 5})
        for (int with_b = 0; with_b < 2; ++with_b) {
            const int mode = 0;
            cudaEvent_t e0, e1;
            cudaEventCreate(&e0); cudaEventCreate(&e1);
            const int grid = sms;
            probe<<<grid, 64, 2 * STAGES * BOX_BYTES + 1024>>>(mode ? m_blk : m_rows, m_b, mode, with_b, row_tiles, kblocks, splits, sink);
            cudaEventRecord(e0);
            for (int r = 0; r < 5; ++r) probe<<<grid, 64, 2 * STAGES * BOX_BYTES + 1024>>>(mode ? m_blk : m_rows, m_b, mode, with_b, row_tiles, kblocks, splits, sink);
            cudaEventRecord(e1);
            cudaError_t e = cudaDeviceSynchronize();
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            printf("K splits %d, %-44s %7.1f us per pass  %5.2f TB/s of A  %s\n", splits, with_b ? "A from HBM + the B block from L2 per k block" : "A only",
                   ms * 1e3 / 5, bytes / (ms / 5 * 1e-3) / 1e12, e == cudaSuccess ? "" : cudaGetErrorString(e));
        }
    return 0;
}

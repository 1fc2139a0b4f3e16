// Fused softmax attention for the Twins-SVT encoder stages (fp16 tier): head dimension 32, any number of queries and
// keys, one pass over the keys with the running-max / running-sum ("online") softmax, so the [queries x keys] score
// matrix never exists in memory.
//
//   LocallyGroupedAttn   twins_ft.py:233-261   49 queries x 49 keys per 7x7 window and head
//   GlobalSubSampleAttn  twins_ft.py:291-312   every token against the keys of the stride-sr sub-sampled map
//   (scale = head_dim^-0.5, softmax over keys, no mask, no dropout at inference)
//
// CTA = 4 warps = 64 queries of one (batch, head); the keys stream through shared memory in blocks of 64.  Per warp and
// key block: S = Q K^T as 16 mma.m16n8k16 (Q fragments live in registers for the whole kernel), the online-softmax update in
// fp32 (exp2f on pre-scaled logits), P V as 16 more mma with the probabilities re-used as A fragments straight from the
// accumulator registers and V read through ldmatrix.trans.  Shared rows are padded to 80 bytes: both the K fragment loads
// (32-bit) and the V ldmatrix rows then hit 32 distinct banks.  Legacy mma.sync is the right tool for these tiles: a
// 49 x 49 x 32 window is far below a tcgen05 tile (M = 128) and this stage is < 3 % of a window.
#include "common.cuh"

namespace vfml {

constexpr int AT_D = 32, AT_KB = 64, AT_PITCH = 40;     // head dim, keys per block, smem row pitch in halves (80 B)

struct AttnParams {
    const __half* q; const __half* k; const __half* v; __half* o;
    int64_t q_sb, q_sn, k_sb, k_sn, v_sb, v_sn, o_sb, o_sn;    // element strides: batch, token (head h sits at + 32 h)
    int nq, nk;
    float scale_log2;       // head_dim^-0.5 * log2(e)
};

__device__ __forceinline__ void mma_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ uint32_t pack_half2(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
}

__global__ void __launch_bounds__(128) attention_hd32_kernel(const AttnParams p) {
    __shared__ __align__(16) __half sK[AT_KB * AT_PITCH];
    __shared__ __align__(16) __half sV[AT_KB * AT_PITCH];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const int head = blockIdx.y, b = blockIdx.x;       // batch on x: windows x frames can exceed the 65535 limit of y / z
    const int q0 = blockIdx.z * 64 + warp * 16;
    const __half* qb = p.q + b * p.q_sb + head * AT_D;
    const __half* kb = p.k + b * p.k_sb + head * AT_D;
    const __half* vb = p.v + b * p.v_sb + head * AT_D;

    // Q fragments (m16k16 x 2): a[kk][0] = (row g, k = 16 kk + 2t, +1), [1] = row g + 8, [2] = (row g, k + 8), [3] = (row g + 8, k + 8)
    uint32_t qa[2][4];
#pragma unroll
    for (int kk = 0; kk < 2; ++kk)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int row = q0 + g + (i & 1) * 8, col = kk * 16 + 2 * t + (i >> 1) * 8;
            qa[kk][i] = row < p.nq ? *reinterpret_cast<const uint32_t*>(qb + (int64_t)row * p.q_sn + col) : 0u;
        }
    float o[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) o[i][j] = 0.0f;
    float m_lo = -INFINITY, m_hi = -INFINITY, l_lo = 0.0f, l_hi = 0.0f;      // rows g and g + 8

    for (int k0 = 0; k0 < p.nk; k0 += AT_KB) {
        __syncthreads();                       // the previous block's fragments have been read
#pragma unroll
        for (int i = 0; i < 2; ++i) {          // 64 rows x 4 chunks of 16 bytes per matrix, 128 threads
            const int chunk = threadIdx.x + i * 128, row = chunk >> 2, c16 = chunk & 3;
            uint4 kq = make_uint4(0, 0, 0, 0), vq = kq;
            if (k0 + row < p.nk) {
                kq = *reinterpret_cast<const uint4*>(kb + (int64_t)(k0 + row) * p.k_sn + c16 * 8);
                vq = *reinterpret_cast<const uint4*>(vb + (int64_t)(k0 + row) * p.v_sn + c16 * 8);
            }
            *reinterpret_cast<uint4*>(sK + row * AT_PITCH + c16 * 8) = kq;
            *reinterpret_cast<uint4*>(sV + row * AT_PITCH + c16 * 8) = vq;
        }
        __syncthreads();
        // S = Q K^T for 64 keys: s[nb] = rows (g, g + 8) x keys 8 nb + 2t, + 1
        float s[8][4];
#pragma unroll
        for (int nb = 0; nb < 8; ++nb) {
#pragma unroll
            for (int j = 0; j < 4; ++j) s[nb][j] = 0.0f;
#pragma unroll
            for (int kk = 0; kk < 2; ++kk) {
                const __half* kr = sK + (nb * 8 + g) * AT_PITCH + kk * 16 + 2 * t;
                mma_16816(s[nb], qa[kk], *reinterpret_cast<const uint32_t*>(kr), *reinterpret_cast<const uint32_t*>(kr + 8));
            }
        }
        // online softmax on logits pre-scaled by scale * log2(e); keys past nk are masked out
        float bm_lo = -INFINITY, bm_hi = -INFINITY;
#pragma unroll
        for (int nb = 0; nb < 8; ++nb)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int key = k0 + nb * 8 + 2 * t + (j & 1);
                const float v = key < p.nk ? s[nb][j] * p.scale_log2 : -INFINITY;
                s[nb][j] = v;
                if (j < 2) bm_lo = fmaxf(bm_lo, v); else bm_hi = fmaxf(bm_hi, v);
            }
#pragma unroll
        for (int x = 1; x <= 2; x <<= 1) {
            bm_lo = fmaxf(bm_lo, __shfl_xor_sync(0xffffffffu, bm_lo, x));
            bm_hi = fmaxf(bm_hi, __shfl_xor_sync(0xffffffffu, bm_hi, x));
        }
        const float mn_lo = fmaxf(m_lo, bm_lo), mn_hi = fmaxf(m_hi, bm_hi);      // finite: every block holds >= 1 valid key
        const float c_lo = exp2f(m_lo - mn_lo), c_hi = exp2f(m_hi - mn_hi);      // exp2f(-inf) = 0 on the first block
        m_lo = mn_lo; m_hi = mn_hi;
        float rs_lo = 0.0f, rs_hi = 0.0f;
        uint32_t pa[4][4];                     // P as A fragments: k16 step j <- s[2j] (keys + 0..7), s[2j + 1] (keys + 8..15)
#pragma unroll
        for (int nb = 0; nb < 8; ++nb) {
            const float p0 = exp2f(s[nb][0] - mn_lo), p1 = exp2f(s[nb][1] - mn_lo);
            const float p2 = exp2f(s[nb][2] - mn_hi), p3 = exp2f(s[nb][3] - mn_hi);
            rs_lo += p0 + p1; rs_hi += p2 + p3;
            pa[nb >> 1][(nb & 1) * 2 + 0] = pack_half2(p0, p1);
            pa[nb >> 1][(nb & 1) * 2 + 1] = pack_half2(p2, p3);
        }
        l_lo = l_lo * c_lo + rs_lo; l_hi = l_hi * c_hi + rs_hi;        // per-thread partial sums; reduced over the quad at the end
#pragma unroll
        for (int nb = 0; nb < 4; ++nb) {
            o[nb][0] *= c_lo; o[nb][1] *= c_lo; o[nb][2] *= c_hi; o[nb][3] *= c_hi;
        }
        // O += P V: B fragments of V (keys x dims) through ldmatrix.trans; one x4 = 16 keys x 16 dims
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int dn = 0; dn < 2; ++dn) {
                // lane -> row address: matrices 0 / 1 = keys 16j + 0..7 / 8..15 at dims 16 dn, matrices 2 / 3 the same keys at dims 16 dn + 8
                const int mrow = j * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, mcol = dn * 16 + (lane >> 4) * 8;
                uint32_t vf[4];
                ldsm_x4_trans(vf, (uint32_t)__cvta_generic_to_shared(sV + mrow * AT_PITCH + mcol));
                mma_16816(o[dn * 2], pa[j], vf[0], vf[1]);
                mma_16816(o[dn * 2 + 1], pa[j], vf[2], vf[3]);
            }
    }
#pragma unroll
    for (int x = 1; x <= 2; x <<= 1) {
        l_lo += __shfl_xor_sync(0xffffffffu, l_lo, x);
        l_hi += __shfl_xor_sync(0xffffffffu, l_hi, x);
    }
    const float i_lo = 1.0f / l_lo, i_hi = 1.0f / l_hi;
    __half* ob = p.o + b * p.o_sb + head * AT_D;
#pragma unroll
    for (int nb = 0; nb < 4; ++nb) {
        const int col = nb * 8 + 2 * t;
        if (q0 + g < p.nq) *reinterpret_cast<uint32_t*>(ob + (int64_t)(q0 + g) * p.o_sn + col) = pack_half2(o[nb][0] * i_lo, o[nb][1] * i_lo);
        if (q0 + g + 8 < p.nq) *reinterpret_cast<uint32_t*>(ob + (int64_t)(q0 + g + 8) * p.o_sn + col) = pack_half2(o[nb][2] * i_hi, o[nb][3] * i_hi);
    }
}

}  // namespace vfml

using namespace vfml;

extern "C" int vfml_attention_hd32(const void* q, const void* k, const void* v, void* o, int n_batch, int n_heads, int n_q, int n_k,
                                   int64_t q_sb, int64_t q_sn, int64_t k_sb, int64_t k_sn, int64_t v_sb, int64_t v_sn, int64_t o_sb,
                                   int64_t o_sn, float scale, vfml_stream_t s) {
    VFML_REQUIRE(n_batch >= 0 && n_heads >= 1 && n_q >= 0 && n_k >= 1, "attention_hd32: bad shape");
    if (n_batch == 0 || n_q == 0) return VFML_OK;
    VFML_REQUIRE(q && k && v && o, "attention_hd32: null pointer");
    VFML_REQUIRE(n_heads <= 65535 && cdiv(n_q, 64) <= 65535, "attention_hd32: more than 65535 heads or query blocks (grid y / z)");
    auto al = [](const void* p_, int a) { return (reinterpret_cast<uintptr_t>(p_) & (a - 1)) == 0; };
    VFML_REQUIRE(al(q, 4) && al(o, 4) && al(k, 16) && al(v, 16), "attention_hd32: q / o must be 4-byte, k / v 16-byte aligned");
    VFML_REQUIRE((q_sn % 2) == 0 && (q_sb % 2) == 0 && (o_sn % 2) == 0 && (o_sb % 2) == 0 && (k_sn % 8) == 0 && (k_sb % 8) == 0 &&
                 (v_sn % 8) == 0 && (v_sb % 8) == 0, "attention_hd32: strides must keep half2 (q, o) / 16-byte (k, v) alignment");
    AttnParams p{static_cast<const __half*>(q), static_cast<const __half*>(k), static_cast<const __half*>(v), static_cast<__half*>(o),
                 q_sb, q_sn, k_sb, k_sn, v_sb, v_sn, o_sb, o_sn, n_q, n_k, scale * 1.4426950408889634f};
    attention_hd32_kernel<<<dim3(n_batch, n_heads, cdiv(n_q, 64)), 128, 0, to_stream(s)>>>(p);
    VFML_LAUNCH_CHECK("attention_hd32");
    return VFML_OK;
}

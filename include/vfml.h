/*
 * vfml.h -- C ABI of libvfml.so, the B200 (sm_100a) kernels behind the video-flow-ml
 * optical-flow hot path.
 *
 * The reference has no FFI of its own on this path except the pybind module
 * alt_cuda_corr (VideoFlow/alt_cuda_corr/correlation.cpp:51-54); everything else is
 * PyTorch library calls made from Python.  Each entry point below cites the
 * reference code (paths relative to the reference checkout) whose work it replaces.
 *
 * Rules of the boundary
 *   - plain pointers and sizes only; all pointers are DEVICE pointers unless a
 *     parameter is documented as host memory;
 *   - nothing here allocates device memory or synchronises: the caller owns every
 *     buffer (including workspaces) and the stream;
 *   - every function returns 0 on success or a negative vfml_status; the text of
 *     the last error of the calling thread is available from vfml_last_error();
 *   - one host thread per stream.
 *
 * "rows" layout: activations of the update step are kept pixel-major
 * ([n_rows, channels] with a row pitch in elements), i.e. NHWC.  The reference's
 * NCHW tensors are converted at the boundary with vfml_nchw_to_rows /
 * vfml_rows_to_nchw.
 */
#ifndef VFML_H
#define VFML_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VFML_VERSION 100

typedef void* vfml_stream_t; /* a cudaStream_t */

enum vfml_status {
    VFML_OK = 0,
    VFML_ERR_ARG = -1,     /* invalid argument (null pointer, bad size, bad enum) */
    VFML_ERR_CUDA = -2,    /* a CUDA runtime call failed; see vfml_last_error() */
    VFML_ERR_UNSUPPORTED = -3,
    VFML_ERR_WORKSPACE = -4 /* caller-provided workspace too small */
};

enum vfml_dtype { VFML_F32 = 0, VFML_F16 = 1 };

enum vfml_layout { VFML_NCHW = 0, VFML_ROWS = 1 };

int vfml_version(void);
const char* vfml_last_error(void);
/* Number of kernels this library has launched since it was loaded (all threads). */
uint64_t vfml_launch_count(void);
/* SM count of the calling thread's current CUDA device (grids are sized from it; 148 on B200). */
int vfml_device_sms(void);

/* ---------------------------------------------------------------------------------
 * Layout conversion.  src [n, c, hw] fp32 planar  <->  rows [n*hw, pitch] of `dtype`.
 * Replaces the permute/contiguous/cat glue of MOFNetStack/network.py:160-170.
 * ------------------------------------------------------------------------------- */
int vfml_nchw_to_rows(const float* src, void* dst, int n, int c, int hw, int dst_pitch,
                      int dst_col, int dtype, vfml_stream_t s);
int vfml_rows_to_nchw(const void* src, float* dst, int n, int c, int hw, int src_pitch,
                      int src_col, int dtype, vfml_stream_t s);

/* ---------------------------------------------------------------------------------
 * All-pairs correlation volume + average-pool pyramid.
 * Replaces CorrBlock.__init__ / CorrBlock.corr, MOFNetStack/corr.py:60-74,99-107
 * (torch.matmul / sqrt(C) followed by 3x F.avg_pool2d(2, stride=2)).
 *
 *   fmap1, fmap2 : [n_pairs, C, h, w] fp32 contiguous
 *   levels[l]    : [n_pairs*h*w, h_l, w_l] of vol_dtype, h_0=h, h_{l+1} = h_l/2 (floor)
 *   workspace    : vfml_corr_volume_workspace(...) bytes
 * ------------------------------------------------------------------------------- */
size_t vfml_corr_volume_workspace(int n_pairs, int C, int h, int w, int vol_dtype);
int vfml_corr_volume_pyramid(const float* fmap1, const float* fmap2, int n_pairs, int C, int h,
                             int w, void* const* levels, int num_levels, int vol_dtype,
                             void* workspace, size_t workspace_bytes, vfml_stream_t s);

/* ---------------------------------------------------------------------------------
 * Multi-level (2r+1)^2 bilinear correlation lookup.
 * Replaces CorrBlock.__call__, MOFNetStack/corr.py:76-97 + bilinear_sampler
 * (core/utils/utils.py:62-76: grid_sample, align_corners=True, zeros padding), and the
 * forward of alt_cuda_corr (alt_cuda_corr/correlation_kernel.cu:18-119) as a lookup.
 *
 *   levels[l] : as written by vfml_corr_volume_pyramid
 *   coords    : [n_frames, 2, h, w] fp32 (channel 0 = x, 1 = y), full-resolution pixels
 *   out       : layout VFML_NCHW -> [n_frames, L*(2r+1)^2, h, w]           (out_pitch ignored)
 *               layout VFML_ROWS -> [n_frames*h*w, out_pitch], columns out_col..out_col+L*(2r+1)^2
 *   channel l*(2r+1)^2 + i*(2r+1) + j samples level l at (x/2^l + i - r, y/2^l + j - r).
 * ------------------------------------------------------------------------------- */
int vfml_corr_lookup(const void* const* levels, int num_levels, int vol_dtype, int n_frames, int h,
                     int w, int radius, const float* coords, void* out, int out_layout,
                     int out_pitch, int out_col, int out_dtype, vfml_stream_t s);

/* Forward and backward lookups of one update iteration (network.py:185-186) as ONE launch, radius 4,
 * rows layout: out_fwd / out_bwd are [n_frames*h*w, out_pitch] buffers. */
int vfml_corr_lookup_pair(const void* const* levels_fwd, const void* const* levels_bwd, int num_levels,
                          int vol_dtype, int n_frames, int h, int w, const float* coords_fwd,
                          const float* coords_bwd, void* out_fwd, void* out_bwd, int out_pitch,
                          int out_col, int out_dtype, vfml_stream_t s);

/* ---------------------------------------------------------------------------------
 * Volume-free lookup.  Replaces alt_cuda_corr.forward -- the reference's one native FFI on this path
 * (VideoFlow/alt_cuda_corr/correlation.cpp:23-33,51-54; kernel correlation_kernel.cu:18-119), called per
 * pyramid level by AlternateCorrBlock.__call__ (MOFNetStack/corr.py:160-177).
 *   fmap1  : [B, H1, W1, C] NHWC, fmap_dtype      fmap2 : [B, H2, W2, C] NHWC, fmap_dtype
 *   coords : [B, N, H1, W1, 2] fp32 (x, y) in fmap2 pixels
 *   corr   : [B, N, (2r+1)^2, H1, W1] fp32; channel ix*(2r+1) + iy = bilinear sample (zeros outside) of
 *            <fmap1[b,h1,w1,:], fmap2[b,.,.,:]> at (x - r + ix, y - r + iy).  No 1/sqrt(C) factor, like the
 *            reference (the Python caller divides).  Fully overwritten (the reference accumulates into zeros).
 *   C must be 128, 256 or 512 (fp16 also 1024); radius 1..4.
 * ------------------------------------------------------------------------------- */
int vfml_alt_corr_forward(const void* fmap1, const void* fmap2, const float* coords, float* corr, int B, int N,
                          int H1, int W1, int H2, int W2, int C, int radius, int fmap_dtype, vfml_stream_t s);

/* ---------------------------------------------------------------------------------
 * Gather-friendly ("tiled") fp16 pyramid + lookup: the engine's layout of the same data.
 * Replaces the same reference code as vfml_corr_volume_pyramid / vfml_corr_lookup
 * (MOFNetStack/corr.py:59-97, core/utils/utils.py:62-76).
 *
 * Level l of source pixel q is a slab of vfml_corr_tiled_slab_bytes(h, w, l) bytes:
 * NTX x NTY tiles (NTX = ceil(w_l/4), NTY = ceil(h_l/4)), tile (tx, ty) at byte
 * (tx*NTY + ty)*32; a tile holds 4 x 4 fp16 texels x-major: element [xr][yy] is texel
 * (4tx+xr, 4ty+yy) of the level, zero beyond the level.  levels[l] is [n_rows] slabs.
 *
 * vfml_corr_tile_pyramid  : vol0 [rows, h*w] fp16 (row-major level 0) -> the tiled levels
 *                           (levels 1.. are 2x2 averages of the previous fp16 level).
 * vfml_corr_volume_pyramid_tiled : fmaps [n_pairs, C, h, w] fp32 -> tiled levels, C % 8 == 0,
 *                           workspace of vfml_corr_volume_tiled_workspace() bytes.  With 4 levels and
 *                           C = 64 or 128 the pyramid is written by the volume GEMM's epilogue
 *                           (corr.py:68-74 fused into :59-66: no row-major volume is materialised;
 *                           environment VFML_CORR_FUSED=0 selects the GEMM + vfml_corr_tile_pyramid
 *                           sequence, same bytes out).
 * vfml_corr_lookup_tiled  : radius 4, 4 levels, one or two sets (levels_b may be NULL) in one
 *                           launch; out_* are fp16 rows buffers [n_frames*h*w, out_pitch],
 *                           out_pitch % 8 == 0, out_col % 8 == 0, 16-byte aligned.  Channel
 *                           order is the PERMUTED one:  tap (level l, x offset i, y offset j)
 *                           of the reference's channel l*81 + i*9 + j is written to
 *                           column (l*9+i)*8 + j for j < 8 and 288 + l*9 + i for j == 8.
 *                           math_f32 != 0 interpolates in fp32 (one rounding), else in fp16x2.
 * ------------------------------------------------------------------------------- */
size_t vfml_corr_tiled_slab_bytes(int h, int w, int level);
int vfml_corr_tile_pyramid(const void* vol0, int64_t rows, int h, int w, void* const* levels,
                           int num_levels, vfml_stream_t s);
size_t vfml_corr_volume_tiled_workspace(int n_pairs, int C, int h, int w);
int vfml_corr_volume_pyramid_tiled(const float* fmap1, const float* fmap2, int n_pairs, int C, int h,
                                   int w, void* const* levels, int num_levels, void* workspace,
                                   size_t workspace_bytes, vfml_stream_t s);
int vfml_corr_lookup_tiled(const void* const* levels_a, const void* const* levels_b, int n_frames,
                           int h, int w, const float* coords_a, const float* coords_b, void* out_a,
                           void* out_b, int out_pitch, int out_col, int math_f32, vfml_stream_t s);

/* ---------------------------------------------------------------------------------
 * vfml_corr_volume_pyramid_tiled_static: NO counterpart in the reference; added
 * for callers that cannot allocate device memory (CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc).  Same inputs, outputs and results as
 * vfml_corr_volume_pyramid_tiled, but the workspace is a 16 MiB array built into the library
 * (vfml_corr_volume_pyramid_tiled_static_max_bytes()); every GPU that loads the library holds
 * it until the process exits.  The call allocates nothing (no cudaMalloc / cudaFree, no
 * allocating library routine).  Shapes with vfml_corr_volume_tiled_workspace(n_pairs, C, h, w)
 * above the maximum return VFML_ERR_WORKSPACE (e.g. C = 256, n_pairs = 1: h*w <= 2651, a
 * 320 x 512 pixel tile).  All calls on a device share the array: calls that overlap on the GPU
 * race; order them on one stream or with events.  VFML_ERR_UNSUPPORTED while VFML_GEMM_TRACE is set.
 * ------------------------------------------------------------------------------- */
size_t vfml_corr_volume_pyramid_tiled_static_max_bytes(void);
int vfml_corr_volume_pyramid_tiled_static(const float* fmap1, const float* fmap2, int n_pairs, int C,
                                          int h, int w, void* const* levels, int num_levels,
                                          vfml_stream_t s);

/* ---------------------------------------------------------------------------------
 * Convex 8x upsampling (softmax over 9 taps fused).
 * Replaces MOFNet.upsample_flow, MOFNetStack/network.py:97-108.
 *   flow : [n, 2, h, w] fp32;  mask: [n, 576, h, w] fp32 (VFML_NCHW) or rows
 *          [n*h*w, mask_pitch] starting at column mask_col (VFML_ROWS, mask_dtype)
 *   out  : fp32, element (i, c, Y, X) at out + i*out_stride_n + c*out_stride_c + Y*out_stride_y + X
 *          for Y in [crop_top, crop_top+out_h), X in [crop_left, crop_left+out_w) of the 8h x 8w
 *          result, written at (Y-crop_top, X-crop_left): unpad + tile paste fused.
 *   flow_is_coords: `flow` holds coords1; the pixel grid is subtracted on the fly (network.py:206-207).
 * ------------------------------------------------------------------------------- */
int vfml_convex_upsample8(const float* flow, const void* mask, int mask_layout, int mask_pitch,
                          int mask_col, int mask_dtype, int n, int h, int w, float* out,
                          int64_t out_stride_n, int64_t out_stride_c, int64_t out_stride_y,
                          int crop_top, int crop_left, int out_h, int out_w, int flow_is_coords,
                          vfml_stream_t s);

/* ---------------------------------------------------------------------------------
 * Flow encodings (bit-exact with the numpy reference given the same fp32 flow).
 * encode_gamedev: GamedevFlowEncoder.encode, encoding/flow_encoders.py:90-117
 * encode_mv_rg8 : MotionVectorsRG8FlowEncoder.encode, encoding/flow_encoders.py:131-153
 *   flow [height, width, 2] fp32 -> rgb [height, width, 3] u8
 *   norm_width/norm_height: the width/height arguments of the reference call.
 * ------------------------------------------------------------------------------- */
int vfml_encode_gamedev(const float* flow, uint8_t* rgb, int height, int width, float norm_width,
                        float norm_height, float scale, float clamp, vfml_stream_t s);
int vfml_encode_mv_rg8(const float* flow, uint8_t* rgb, int height, int width, float clamp,
                       vfml_stream_t s);
/* MotionVectorsRG8FlowEncoder.decode, flow_encoders.py:155-177. rgb u8 -> flow fp32 */
int vfml_decode_mv_rg8(const uint8_t* rgb, float* flow, int height, int width, float clamp,
                       vfml_stream_t s);

/* ---------------------------------------------------------------------------------
 * Remaining flow encoders / decoders, bit-exact with the reference's numpy / torchvision /
 * OpenCV arithmetic given the same fp32 flow.  flow [h, w, 2] fp32, rgb [h, w, 3] u8.
 *   vfml_encode_mv_rgb8   MotionVectorsRGB8FlowEncoder.encode, 'rgb+' method (encoding/flow_encoders.py:10,242-262,292-295)
 *   vfml_decode_mv_rgb8   MotionVectorsRGB8FlowEncoder.decode, 'rgb+' (:308,340-347,362-369)
 *   vfml_encode_hsv       HSVFlowEncoder.encode (:38-67) incl. cv2.cvtColor(HSV2RGB)
 *   vfml_encode_torchvision  TorchvisionFlowEncoder.encode (:399-427): torchvision.utils.flow_to_image;
 *                         ref_wrap = 1 also applies the reference's `(img * 255).astype(uint8)` on the uint8 image (:425)
 * scratch4: 4 bytes of device memory (the max-magnitude reduction of the two normalising encoders).
 * ------------------------------------------------------------------------------- */
int vfml_encode_mv_rgb8(const float* flow, uint8_t* rgb, int height, int width, float clamp, vfml_stream_t s);
int vfml_decode_mv_rgb8(const uint8_t* rgb, float* flow, int height, int width, float clamp, vfml_stream_t s);
int vfml_encode_hsv(const float* flow, uint8_t* rgb, int height, int width, void* scratch4, vfml_stream_t s);
int vfml_encode_torchvision(const float* flow, uint8_t* rgb, int height, int width, int ref_wrap, void* scratch4,
                            vfml_stream_t s);

/* ---------------------------------------------------------------------------------
 * LOD pyramid of a flow field, LODGenerator.generate_lods (storage/cache_manager.py:77-161): level l+1 [ceil(h/2),
 * ceil(w/2), 2] = 0.5 * mean over the existing cells of each 2x2 block of level l.  One launch writes num_levels
 * (1..4) reduced levels into levels[0..num_levels-1] (HOST array of device pointers); chain calls for more.
 * Bit-identical with the reference's per-pixel Python loop (same summation order).
 * ------------------------------------------------------------------------------- */
int vfml_lod_pyramid(const float* flow, int height, int width, float* const* levels, int num_levels, vfml_stream_t s);

/* ---------------------------------------------------------------------------------
 * Flow quality map, generate_quality_frame_gpu (correction_worker.py:175-208): frame1 / frame2 [h, w, 3] u8, flow
 * [flow_h, flow_w, 2] fp32 (resized bilinearly, align_corners=False, and rescaled when its size differs, :182-185),
 * out [h, w, 3] u8: green = similarity of frame1[p] and frame2[p - flow], red = dissimilar / out of bounds.
 * ------------------------------------------------------------------------------- */
int vfml_quality_frame(const uint8_t* frame1, const uint8_t* frame2, const float* flow, int flow_h, int flow_w,
                       uint8_t* out, int height, int width, float threshold, vfml_stream_t s);

/* ---------------------------------------------------------------------------------
 * Temporal anti-aliasing, effects/taa_processor.py:41-262.
 *   cur      : [h, w, 3] current frame, u8 (cur_f32 = 0) or fp32 (cur_f32 = 1: the reference converts with
 *              current_frame.astype(np.float32), :64, so float frames keep their fractions)
 *   flow     : [h, w, 2] fp32 (previous frame's flow, flow_processor.py:1018-1029) or NULL
 *   history  : [h, w, 3] f64 previous output; out: [h, w, 3] f64 (may not alias history)
 *   mode     : 0 = simple EMA (:78-80), 1 = flow + bilinear (:218-262), 2 = flow + bilateral (:149-216)
 *   hist_f32 : 1 when the reference would hold the history as float32 at this point (frame 1 of
 *              a sequence, and always in modes 0/1): colour weights / blends are then evaluated in
 *              fp32 exactly as numpy's dtype promotion does; values are still stored f64.
 * ------------------------------------------------------------------------------- */
int vfml_taa(const void* cur, int cur_f32, const float* flow, const double* history, double* out, int h, int w,
             double alpha, double sigma_color, int mode, int hist_f32, vfml_stream_t s);
/* first frame of a sequence: out = double(cur) (taa_processor.py:70-74) */
int vfml_taa_init(const void* cur, int cur_f32, double* out, int h, int w, vfml_stream_t s);

/* ---------------------------------------------------------------------------------
 * Building blocks of the update step (MOFNetStack/stack.py, gma.py).
 * ------------------------------------------------------------------------------- */

/* epilogue flags of vfml_gemm */
#define VFML_EPI_BIAS 1u   /* + bias[n] */
#define VFML_EPI_RES 2u    /* + res[m, n] */
#define VFML_EPI_GELU 4u   /* erf GELU */
#define VFML_EPI_RELU 8u
#define VFML_EPI_DW1 16u   /* then v = gelu(v + dw_w[n]*v + dw_b[n]): the k=1 depthwise stage of a PCBlock */
#define VFML_EPI_ADDTO 32u /* out += v (fp32 out only) */
#define VFML_EPI_WEIGHTS 64u /* B, bias, dw_w, dw_b are constants (layer weights), not outputs of preceding kernels on the
                                stream: the kernel may fetch them while the preceding kernel is still draining
                                (programmatic dependent launch).  Never changes results. */

typedef struct vfml_gemm_desc {
    /* D[m, n] = epi(alpha * sum_k A[m, k] * B[n, k]),  m < M, n < N, k < K */
    const void* A;   /* [M, lda] rows, K contiguous */
    const void* B;   /* [N, ldb] rows, K contiguous (a conv weight [Cout, Cin]) */
    void* D;         /* [M, ldd] */
    const float* bias;  /* [N] fp32 or NULL */
    const void* res;    /* [M, ldr] same dtype as A, or NULL */
    const float* dw_w;  /* [N] fp32 (VFML_EPI_DW1) */
    const float* dw_b;  /* [N] fp32 */
    int64_t lda, ldb, ldd, ldr;
    int M, N, K;
    int batch;                          /* >= 1; strides below in elements */
    int64_t strideA, strideB, strideD, strideR;
    float alpha;
    uint32_t epi;
    int in_dtype;   /* dtype of A, B, res */
    int out_dtype;  /* dtype of D */
} vfml_gemm_desc;

/* Pointwise (1x1) convolution / matmul with fused epilogue.  Replaces the nn.Conv2d(k=1) +
 * GELU + residual chains of PCBlock4_Deep_nopool_res (stack.py:8-32), Attention.to_qk / q.k^T
 * (gma.py:54-74) and Aggregate's attn.v (gma.py:102-115). */
int vfml_gemm(const vfml_gemm_desc* d, vfml_stream_t s);
/* Optional, outside CUDA graph capture (it SYNCHRONISES the stream): times the kernels that can run this descriptor -- the
 * general tcgen05 kernel and the weight-stationary cluster kernel in its single-CTA and CTA-pair modes -- on the caller's
 * buffers, and makes later vfml_gemm calls of the same (device, M, N, K, epilogue) use the fastest.  Never changes results
 * beyond fp16 rounding of identical math.  Returns 0 (general kernel) or 1 + 16 * pair + cluster_size of the chosen cluster
 * configuration, or a negative vfml_status; descriptors whose output overlaps an input are left on the general kernel.  Same reference code as vfml_gemm. */
int vfml_gemm_autotune(const vfml_gemm_desc* d, vfml_stream_t s);
/* 1 if vfml_gemm would run this descriptor on the tcgen05 tensor-core kernel (fp16 operands, 16-byte
 * aligned bases and pitches), 0 if on the exact fp32 FFMA kernel. */
int vfml_gemm_uses_tensor_cores(const vfml_gemm_desc* d);

/* Tensor-core path for fp32 operands (the exact tier's pointwise layers): x = hi + lo with hi = fp16(x), lo = fp16(x - hi);
 * A B^T ~= Ahi Bhi^T + Ahi Blo^T + Alo Bhi^T as ONE fp16 tcgen05 GEMM over a tripled K with fp32 accumulation (relative error
 * ~2^-21 per product), then the fp32 epilogue of vfml_gemm (alpha, bias, residual, erf-GELU / ReLU, dw1) in a second pass.
 * Same reference code as vfml_gemm.  d: fp32-in descriptor, batch 1, no VFML_EPI_ADDTO; d->B is ignored in favour of
 * b_split = the weight as written by vfml_gemm_split_weight ([N, 3 * roundup8(K)] fp16, vfml_gemm_split_weight_bytes);
 * workspace: vfml_gemm_split_workspace(M, N, K) bytes.  |x| must stay below the fp16 range (65504). */
size_t vfml_gemm_split_weight_bytes(int N, int K);
int vfml_gemm_split_weight(const float* w, int N, int K, int64_t ldb, void* out, vfml_stream_t s);
size_t vfml_gemm_split_workspace(int M, int N, int K);
int vfml_gemm_f32_split(const vfml_gemm_desc* d, const void* b_split, void* workspace, size_t workspace_bytes,
                        vfml_stream_t s);

/* Depthwise k x k convolution over rows layout with the PCBlock residual + GELU fused:
 *   y = gelu(x + dwconv_k(x) + b)     stack.py:25-26
 * x, y: [n_img*h*w, pitch] (`dtype`), C channels from column 0; weight fp32 TAP-MAJOR [k*k, C]
 * (prepacked from the checkpoint's [C, 1, k, k]), bias [C]. */
int vfml_dwconv_res_gelu(const void* x, void* y, int n_img, int h, int w, int C, int64_t pitch_x,
                         int64_t pitch_y, const float* weight, const float* bias, int k, int dtype,
                         vfml_stream_t s);

/* The same operation on the tensor cores (fp16 rows in and out, fp32 accumulation; k = 7 or 15).
 * The depthwise weights are prepacked once into a banded-Toeplitz fragment table (fp16):
 *   packed = vfml_dwconv_packed_bytes(C, k) bytes, 16-byte aligned, written by vfml_dwconv_pack_weights
 *   from the same fp32 tap-major [k*k, C] weight vfml_dwconv_res_gelu takes.
 * x != y; pitches are multiples of 8 elements and >= C rounded up to 8 (the padding columns are read
 * and written: they must belong to the row). */
size_t vfml_dwconv_packed_bytes(int C, int k);
int vfml_dwconv_pack_weights(const float* weight, int C, int k, void* packed, vfml_stream_t s);
int vfml_dwconv_tc_res_gelu(const void* x, void* y, int n_img, int h, int w, int C, int64_t pitch_x,
                            int64_t pitch_y, const void* packed, const float* bias, int k,
                            vfml_stream_t s);

/* out[r, :n] = res[r, :n] + sum_s partial[s, r, :n]: combines split-K partial products (fp32, [n_splits, rows, n]
 * contiguous) of the GMA aggregate `attn @ v` (gma.py:113) / MemFlow read-out (memory_manager_skflow.py:88-89) with the
 * motion features (`fmap + gamma * out`, gma.py:115; gamma rides in the GEMM's alpha).  res may be NULL; n % 4 == 0. */
int vfml_splitk_reduce(const float* partial, int n_splits, int64_t rows, int n, const void* res, int64_t ld_res,
                       void* out, int64_t ld_out, int dtype, vfml_stream_t s);

/* y[r, :C] = LayerNorm(x[r, :C]) * gamma + beta over rows buffers (x, y, gamma, beta of `dtype`, fp32 statistics).
 * Replaces nn.LayerNorm of the Twins blocks (VideoFlow/core/Networks/twins_ft.py norm1 / norm2 of Block, eps 1e-6),
 * the first native piece of SURVEY 8(f) n1.  C a multiple of 8 (fp16) / 4 (fp32), C <= 1024 (fp16) / 512 (fp32). */
int vfml_layer_norm_rows(const void* x, void* y, int64_t rows, int C, int64_t ld_x, int64_t ld_y, const void* gamma,
                         const void* beta, float eps, int dtype, vfml_stream_t s);

/* Row softmax in place-capable form: out[r, :] = softmax(in[r, :n]) (gma.py:74). in fp32. */
int vfml_softmax_rows(const float* in, void* out, int64_t rows, int n, int64_t ld_in, int64_t ld_out,
                      int out_dtype, vfml_stream_t s);

/* Motion-encoder glue (stack.py:69-103), one launch:
 *   flow4 = [fcoords - grid, bcoords - grid]                                  (network.py:183-184)
 *   x400[:, 256:304] = sample(mhs of frame t+1, fcoords)  (zero state past the last frame)
 *   x400[:, 304:352] = sample(mhs of frame t-1, bcoords)  (zero state before the first frame)
 *   f128 = convf1_(flow4)  (1x1 conv, 4 -> 128)            written to f128 rows
 *   mf[:, 124:128] = flow4
 * mhs lives in x400[:, 352:400].
 * rot (0 <= rot < n_frames): the frames are stored in RING order -- slot f of every buffer holds the window's frame
 * (f - rot) mod n_frames -- so "frame t+1" of slot f is slot (f + 1) mod n_frames, and the window's last / first frame
 * (zero state) sits in slot (rot - 1) mod n_frames / rot.  The engine keeps a sliding window's per-frame state (pyramids,
 * attention) in place and only rotates; rot = 0 is the reference's order. */
int vfml_motion_glue(const float* fcoords, const float* bcoords, void* x400, int64_t pitch400,
                     void* f128, int64_t pitch128, void* mf, int64_t pitch_mf, int mf_col,
                     const float* convf1_w, const float* convf1_b, int n_frames, int h, int w,
                     int rot, int dtype, vfml_stream_t s);

/* Fused softmax attention of the Twins-SVT stages, fp16, head dimension 32 (LocallyGroupedAttn twins_ft.py:233-261,
 * GlobalSubSampleAttn :291-312): o[b, i, h, :] = softmax_j(scale * q[b, i, h, :] . k[b, j, h, :]) v[b, j, h, :].
 * Tensors are addressed as base + b * *_sb + token * *_sn + 32 * h (element strides), so the q / k / v slices of a fused
 * qkv or kv projection are used in place.  The score matrix is never materialised (online softmax over key blocks). */
int vfml_attention_hd32(const void* q, const void* k, const void* v, void* o, int n_batch, int n_heads, int n_q, int n_k,
                        int64_t q_sb, int64_t q_sn, int64_t k_sb, int64_t k_sn, int64_t v_sb, int64_t v_sn, int64_t o_sb,
                        int64_t o_sn, float scale, vfml_stream_t s);

/* Twins PEG (PosConv, VideoFlow/core/Networks/twins_ft.py:341-355): y = x + dwconv3x3(x) + b on token-major rows
 * [n_img*h*w, C] (fp32 or fp16; C multiple of 4 / 8), weight tap-major [9, C] and bias [C] in the same dtype. */
int vfml_peg3x3(const void* x, void* y, int n_img, int h, int w, int C, int64_t ld_x, int64_t ld_y, const void* weight,
                const void* bias, int dtype, vfml_stream_t s);

/* coords[n,2,h,w] (+)= delta rows: fcoords += delta[:, 0:2]; bcoords += delta[:, 2:4] (network.py:196-197) */
int vfml_coords_add(float* fcoords, float* bcoords, const float* delta, int64_t ld_delta, int n_frames,
                    int hw, vfml_stream_t s);

/* coords_grid, core/utils/utils.py:97-100 */
int vfml_coords_grid(float* coords, int n, int h, int w, vfml_stream_t s);

/* fill rows[:, col:col+C] with a per-channel constant (init_hidden_state broadcast, stack.py:72-73) */
int vfml_rows_fill(void* rows, int64_t pitch, int col, int C, int64_t n_rows, const float* values,
                   int dtype, vfml_stream_t s);

/* im2col for the 3x3 mask-head convolution (stack.py:121-124): x rows [n*h*w, pitch] C channels ->
 * out rows [n*h*w, 9*C], column (ky*3+kx)*C + c, zero padding. */
int vfml_im2col3x3(const void* x, void* out, int n_img, int h, int w, int C, int64_t pitch_x,
                   int64_t pitch_out, int dtype, vfml_stream_t s);

#ifdef __cplusplus
}
#endif
#endif /* VFML_H */

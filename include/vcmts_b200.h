/*
 * include/vcmts_b200.h -- C-ABI of libvcmts_b200.so (hand-written sm_100a CUDA kernels +
 * host entropy coder) for the DCVC-HEM P-/I-frame encode/decode path that vcm-ts wraps.
 *
 * Conventions
 *   - every entry returns 0 on success, non-zero on error; vcm_last_error() gives the text
 *     (thread-local).  No exceptions cross this boundary, nothing is allocated on the device,
 *     the caller owns every buffer.
 *   - device tensors are NHWC ("channels-last") *views*: base pointer already offset to the
 *     first channel of the slice, `cstride` = channels per pixel of the underlying buffer.
 *     This is how torch.cat / chunk / split of the reference become free.
 *   - `dtype`: VCM_F32 (exact mode) or VCM_BF16 (tensor-core mode) is the storage type of
 *     activations; weights/bias/tables are fp32 unless stated.
 *   - `stream` is a cudaStream_t passed as void*.
 *
 * Each entry cites the reference interface (under /root/reference/) it replaces.
 */
#ifndef VCMTS_B200_H
#define VCMTS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VCM_F32 0
#define VCM_BF16 1

/* conv flags */
#define VCM_CONV_PRE_ACT 1        /* leaky_relu(pre_slope) applied to the input while loading  */
#define VCM_CONV_POST_ACT 2       /* leaky_relu(post_slope) after bias (slope 0 == ReLU)       */
#define VCM_CONV_PIXEL_SHUFFLE 4  /* store through nn.PixelShuffle(2)                          */
#define VCM_CONV_RES 8            /* out = res + out            (after the activation)         */
#define VCM_CONV_RES_SCALE 16     /* out = out + res*res_scale[n][c]   (SELayer + up_dim add)  */
#define VCM_CONV_RES2 32          /* out = res2 + out           (after RES)                    */
#define VCM_CONV_OUT2 64          /* also store leaky_relu(out, out2_slope) to y2: the activated
                                     twin a following ResBlock's first conv reads, so that the
                                     tensor-core engine never has to modify TMA-loaded tiles     */
#define VCM_CONV_SPLIT3 128       /* parity mode of the tensor-core engine: x holds THREE bf16 planes
                                     x0 | x1 | x2 with x0 + x1 + x2 == the fp32 activation exactly
                                     (vcm_split_bf16x3), plane p at channel offset p*in_pstride, and
                                     w_umma holds the three bf16 parts of the fp32 weights
                                     (vcm_pack_conv_weight_umma_split3).  The engine issues the six
                                     products x_i*w_j with i + j <= 2 (smallest first) into the fp32
                                     TMEM accumulator: fp32-class accuracy on tcgen05.               */

#define VCM_CONV_PDL 256          /* tensor-core engine: launch with programmatic dependent launch, i.e. the
                                     kernel's prologue (TMEM allocation, barrier init, WEIGHT loads) may
                                     overlap the tail of the previous kernel in the stream; activations,
                                     residuals and outputs are only touched after the predecessor finished.
                                     Only set it when the packed weights were written by an EARLIER,
                                     completed launch (not by the kernel just before this one).           */

#define VCM_CONV_SE_SUM 512       /* tensor-core engine, bf16: the SELayer mean (video_net.py:149-162 `avg_pool`)
                                     fused into the conv that produces its input: y2 is then a float array
                                     [vcm_conv2d_se_sum_rows(d)][Cout] that receives per-warp channel sums of
                                     the bf16 outputs; vcm_se_scale_from_sums turns them into the scales.
                                     Only where vcm_conv2d_se_sum_rows(d) > 0 (N = 1, whole tiles, Cout 64 / 32,
                                     no OUT2); otherwise the launch fails.                               */

typedef struct {
  int32_t N, H, W;        /* input batch / height / width                                   */
  int32_t Cin, Cout;      /* logical channels (Cout before pixel shuffle)                   */
  int32_t K, stride;      /* square kernel K in {1,3,7}, stride in {1,2}, padding = K/2      */
  int32_t in_cstride, out_cstride, res_cstride, res2_cstride, out2_cstride;
  int32_t flags;
  float pre_slope, post_slope, out2_slope;
  int32_t in_dtype;       /* VCM_F32 | VCM_BF16 : storage of x                               */
  int32_t out_dtype;      /* storage of y, res and res2 (mixed in/out serves the fp32<->bf16
                             boundaries: flow fields and frames stay fp32)                    */
  int32_t in_pstride;     /* VCM_CONV_SPLIT3: channels between consecutive bf16 planes of x   */
} vcm_conv_desc;

const char* vcm_last_error(void);
int vcm_version(void);
/* 1 if a CUDA device of compute capability 10.x is visible */
int vcm_device_ok(void);

/* ---- convolution stack ---------------------------------------------------------------
 * replaces every nn.Conv2d (+ LeakyReLU/ReLU, residual add, PixelShuffle, concat store) of
 * DCVC_HEM/src/models/video_model.py:17-224, video_net.py:74-276, layers/layers.py:18-127,
 * image_model.py:16-48.  Weights are pre-packed by vcm_pack_conv_weight.                  */

/* OIHW fp32 (host or device pointer on the device side: device) -> packed [K*K][Cin_p][Cout_p]
 * fp32 with optional input-channel gather `in_map` (Cin_p entries, -1 = zero row).        */
int vcm_pack_conv_weight(const float* w_oihw, float* w_packed, int32_t Cout, int32_t Cin,
                         int32_t K, int32_t Cin_p, const int32_t* in_map, void* stream);
/* exact-mode engine: fp32 FMA implicit GEMM (any shape)                                    */
int vcm_conv2d_fwd(const vcm_conv_desc* d, const void* x, const float* w_packed,
                   const float* bias, const void* res, const float* res_scale,
                   const void* res2, void* y, void* y2, void* stream);

/* tensor-core engine: tcgen05 bf16 implicit GEMM, TMA-fed, TMEM accumulators.
 * w_umma: weights pre-packed by vcm_pack_conv_weight_umma.  Supported: K in {1,3,7},
 * stride 1 or 2, Cin % 16 == 0, Cout % 16 == 0 (<= 256), in/out dtype VCM_BF16, no PRE_ACT.  Returns 2 when the shape
 * is not supported (caller then uses vcm_conv2d_fwd).                                      */
int vcm_conv2d_tc_supported(const vcm_conv_desc* d);
int64_t vcm_conv2d_tc_weight_bytes(int32_t Cout, int32_t Cin, int32_t K);
int vcm_pack_conv_weight_umma(const float* w_oihw, void* w_umma, int32_t Cout, int32_t Cin,
                              int32_t K, int32_t Cin_p, const int32_t* in_map, void* stream);
int vcm_conv2d_tc_fwd(const vcm_conv_desc* d, const void* x, const void* w_umma,
                      const float* bias, const void* res, const float* res_scale,
                      const void* res2, void* y, void* y2, void* stream);
/* parity mode (VCM_CONV_SPLIT3).  The reference computes every nn.Conv2d in fp32
 * (video_model.py:17-224 has no autocast); its quantised symbols (common_model.py:91-102) flip
 * under plain bf16 operands.  Splitting both operands into three bf16 terms keeps 24 significand
 * bits: src fp32 view [npix][C] (cstride src_cstride), optional leaky_relu(pre_slope) first, ->
 * dst bf16 planes p = 0..2 at dst + p*dst_pstride, each [npix][Cp] with cstride dst_cstride,
 * channels C..Cp-1 zero.  x0 = bf16(x), x1 = bf16(x - x0), x2 = bf16(x - x0 - x1).
 * Weights: 3 * vcm_conv2d_tc_weight_bytes bytes, layout [part][tap][chunk][Cout_p][CHUNK].      */
int vcm_split_bf16x3(const float* src, void* dst, int64_t npix, int32_t C, int32_t Cp,
                     int32_t src_cstride, int32_t dst_cstride, int32_t dst_pstride,
                     int32_t pre_act, float pre_slope, void* stream);
int vcm_pack_conv_weight_umma_split3(const float* w_oihw, void* w_umma, int32_t Cout, int32_t Cin,
                                     int32_t K, int32_t Cin_p, const int32_t* in_map, void* stream);
/* ---- layout / elementwise glue ------------------------------------------------------- */
/* NCHW fp32 -> NHWC view (dtype) ; callers' tensors are NCHW fp32 (video_coder.py:99-117)  */
int vcm_nchw_to_nhwc(const float* src, void* dst, int32_t N, int32_t C, int32_t H, int32_t W,
                     int32_t dst_cstride, int32_t dtype, void* stream);
/* frame I/O: uint8 HWC [N][H][W][3] (PNG decode output, DCVC_HEM/src/utils/png_reader.py:37-39) ->
 * float NCHW [N][3][Hp][Wp] = u8 / 255.f, zero-padded right / bottom (test_video.py:118-125,
 * video_coder.py:112); and the inverse of save_torch_image (stream_helper.py:151-154):
 * crop to H x W, clip(rint(v * 255), 0, 255) from a float NCHW [N][3][Hs][Ws] reconstruction.        */
int vcm_u8hwc_to_frame(const uint8_t* src, float* dst, int32_t N, int32_t H, int32_t W, int32_t Hp,
                       int32_t Wp, void* stream);
int vcm_frame_to_u8hwc(const float* src, uint8_t* dst, int32_t N, int32_t H, int32_t W, int32_t Hs,
                       int32_t Ws, void* stream);
/* ---- enhancement layer: uint8 HWC frames [N][H][W][C] on the device ----
 * boxes: int32 [total][4] = (x1, y1, x2, y2), already normalised to 0 <= x1 <= x2 <= W, 0 <= y1 <= y2 <= H;
 * box_start: int32 [N + 1], frame n owns boxes [box_start[n], box_start[n+1]) (<= 256 per frame, in the
 * order the reference assigns them: later boxes overwrite earlier ones); max_boxes: largest per-frame count.
 * vcm_enh_residual = compute_residuals' per-frame statements (video_coder.py:415-447):
 *   out = inside any box ? clip(src - enc + 128, 0, 255) : 0
 * vcm_enh_fuse = fuse_layers' per-frame statements (video_coder.py:570-600) with create_gradient_mask
 *   (:521-530): out = clip(base + mask * (enh - 128), 0, 255), mask = 0 outside the boxes and inside box b
 *   ring[ring_start[b] + min(d, len_b - 1)], d = distance to the box border, len_b = ring_start[b+1] -
 *   ring_start[b]; ring holds float32(1 - linspace(0.9, 0, border)) per box ({1.0} for border 0).        */
int vcm_enh_residual(const uint8_t* src, const uint8_t* enc, uint8_t* out, int32_t N, int32_t H, int32_t W,
                     int32_t C, const int32_t* boxes, const int32_t* box_start, int32_t max_boxes, void* stream);
int vcm_enh_fuse(const uint8_t* base, const uint8_t* enh, uint8_t* out, int32_t N, int32_t H, int32_t W, int32_t C,
                 const int32_t* boxes, const int32_t* box_start, const int32_t* ring_start, const float* ring,
                 int32_t max_boxes, void* stream);
/* NHWC view -> NCHW fp32, optional clamp to [lo,hi] (video_model.py:413 `recon.clamp(0,1)`)  */
int vcm_nhwc_to_nchw(const void* src, float* dst, int32_t N, int32_t C, int32_t H, int32_t W,
                     int32_t src_cstride, int32_t dtype, int32_t do_clamp, float lo, float hi,
                     void* stream);
/* channel-slice copy with optional fp32 <-> bf16 conversion */
int vcm_copy_channels(const void* src, void* dst, int64_t npix, int32_t C, int32_t src_cstride,
                      int32_t dst_cstride, int32_t src_dtype, int32_t dst_dtype, void* stream);
/* x[n,:,c] (op) vec[n,c]; op 0 = divide, 1 = multiply (video_model.py:280,297,306,326)      */
int vcm_scale_channels(const void* src, void* dst, const float* vec, int32_t N, int64_t hw,
                       int32_t C, int32_t src_cstride, int32_t dst_cstride, int32_t op,
                       int32_t dtype, void* stream);
/* F.max_pool2d(x,2,2)  video_net.py:185,207,210 */
int vcm_maxpool2(const void* src, void* dst, int32_t N, int32_t H, int32_t W, int32_t C,
                 int32_t src_cstride, int32_t dst_cstride, int32_t dtype, void* stream);
/* SELayer video_net.py:149-162: deterministic per-(n,c) mean, then sigmoid(W2 relu(W1 m)).
 * `partial` is scratch of vcm_se_partial_count(H*W)*N*C floats.                             */
int32_t vcm_se_partial_count(int64_t hw);
int vcm_se_scale(const void* x, float* partial, const float* w1, const float* w2, float* scale,
                 int32_t N, int64_t hw, int32_t C, int32_t Cmid, int32_t cstride, int32_t dtype,
                 void* stream);
/* the same from channel sums written by a VCM_CONV_SE_SUM conv (N = 1): rows of [C] floats, added in index
 * order; rows = vcm_conv2d_se_sum_rows(d) of that conv (0: the conv cannot produce them, use vcm_se_scale) */
int32_t vcm_conv2d_se_sum_rows(const vcm_conv_desc* d);
int vcm_se_scale_from_sums(const float* sums, int32_t rows, const float* w1, const float* w2, float* scale,
                           int64_t hw, int32_t C, int32_t Cmid, void* stream);

/* ---- motion glue (fp32 always: flow fields and SpyNet image pyramids) ------------------
 * F.avg_pool2d(x,2,2)            video_net.py:131-133
 * bilinearupsacling(x)*2         video_net.py:58-63,140
 * bilineardownsacling(mv)/2      video_net.py:66-71, video_model.py:237-238
 * torch_warp / flow_warp         video_net.py:32-55                                         */
int vcm_avgpool2(const float* src, float* dst, int32_t N, int32_t H, int32_t W, int32_t C,
                 int32_t src_cstride, int32_t dst_cstride, void* stream);
int vcm_flow_up2(const float* src, float* dst, int32_t N, int32_t H, int32_t W,
                 int32_t src_cstride, int32_t dst_cstride, float gain, void* stream);
int vcm_flow_down2(const float* src, float* dst, int32_t N, int32_t H, int32_t W,
                   int32_t src_cstride, int32_t dst_cstride, float gain, void* stream);
/* Fused SpyNet glue (ME_Spynet.forward, video_net.py:118-146), bit-identical to chains of the four entries above.
 * vcm_spynet_pyramid: NCHW fp32 frames x (im1) and ref (im2), H and W multiples of 8 -> all four pyramid levels
 *   of both (avg_pool2d three times, :131-133): im1 level k into channels 0..2 of the 8-channel fp32 level-input
 *   buffer lvl[k] [N, H>>k, W>>k, 8], im2 level k into im2[k] [N, H>>k, W>>k, 3].
 * vcm_spynet_level_input: one level's network input (:139-146): flow_up = bilinear_x2(flow_prev) * 2 into
 *   channels 6,7 of lvl, flow_warp(im2 level, flow_up) into channels 3..5 (0..2 hold im1 already); buf16 (may
 *   be NULL): the same eight channels as bf16 padded to 16 (tensor-core first layer).
 *   flow_prev: [N, h/2, w/2] 2-channel fp32 view (prev_cstride channels per pixel).                        */
int vcm_spynet_pyramid(const float* x, const float* ref, float* const* lvl, float* const* im2, int32_t N,
                       int32_t H, int32_t W, void* stream);
int vcm_spynet_level_input(const float* flow_prev, const float* im2, float* lvl, void* buf16, int32_t N,
                           int32_t h, int32_t w, int32_t prev_cstride, void* stream);
/* feature view (dtype) warped by a 2-channel fp32 flow view: bilinear, border, align_corners */
int vcm_flow_warp(const void* feat, const float* flow, void* out, int32_t N, int32_t H,
                  int32_t W, int32_t C, int32_t feat_cstride, int32_t flow_cstride,
                  int32_t out_cstride, int32_t dtype, void* stream);

/* ---- training path: backward kernels ------------------------------------
 * The reference trains DMC through torch autograd (core/model/dcvc_hem.py:254-322, loss.backward() in
 * core/engine/train_multi.py:203-268).  Behind this ABI: the weight gradient of a convolution
 *   dW[co][ci][ky][kx] = sum_pixels dY[pix][co] * X[pix*stride + tap - pad][ci]      (OIHW fp32 out)
 * with x described by d (in_*) and dy = the gradient of the conv output [N,Ho,Wo] (out_dtype, out_cstride);
 * `scratch`: vcm_conv2d_wgrad_scratch_floats(d) floats (split partial sums, reduced in fixed order);
 * and the backward of flow_warp (video_net.py:32-55): gfeat fp32 dense [N,H,W,C] and gflow fp32 dense
 * [N,H,W,2], both ZERO-INITIALISED by the caller (taps accumulate atomically).  The data gradient of a
 * convolution is vcm_conv2d_*_fwd on flipped, transposed weights (vcm_ts_b200/training.py).             */
int64_t vcm_conv2d_wgrad_scratch_floats(const vcm_conv_desc* d);
int vcm_conv2d_wgrad(const vcm_conv_desc* d, const void* x, const void* dy, float* dw_oihw,
                     float* scratch, void* stream);
int vcm_flow_warp_bwd(const void* feat, const float* flow, const void* gout, float* gfeat, float* gflow,
                      int32_t N, int32_t H, int32_t W, int32_t C, int32_t feat_cstride,
                      int32_t flow_cstride, int32_t gout_cstride, int32_t dtype, void* stream);

/* ---- entropy-parameter / quantise / likelihood epilogues -------------------------------
 * common_model.py:91-217 (process_with_mask, forward/compress/decompress_dual_prior),
 * entropy_models.py:264-268 (build_indexes).  One fused kernel per checkerboard pass.
 *
 *  y        : latent already divided by the outer q (video_model.py:280,306)  [N,h,w,C]
 *  fusion   : prior-fusion output, channel order (q_step, scales, means), 3C channels
 *  prior    : spatial-prior output, channel order (scales_0, means_0, scales_1, means_1), 2C
 *  params   : the 4C-channel spatial-prior input  (y_hat_0_0, y_hat_1_1, means, scales, q_step)
 *  yq       : scratch, y / max(q_step, .5)
 *  sym/idx  : [N, C/2, h, w] planes in the reference's coding order (int16 / uint8)
 *  y_q_full / scales_full (optional, may be NULL): y_q and scales_hat of forward_dual_prior,
 *             NHWC fp32 [N,h,w,C], for the likelihood kernel.
 *  mode: 0 = encode (quantise, emit sym+idx, reconstruct)
 *        1 = decoder index pass (emit idx only)
 *        2 = decoder reconstruct pass (consume sym)
 *  log_scale_min / log_scale_step : GaussianEncoder table constants (entropy_models.py:202-216)
 */
typedef struct {
  int32_t N, h, w, C;
  int32_t y_cstride, fusion_cstride, prior_cstride, params_cstride, yhat_cstride;
  int32_t dtype;   /* storage of y, fusion, prior, params, yhat */
  int32_t mode;
  float log_scale_min, log_scale_step;
} vcm_dp_desc;
int vcm_dual_prior_pass0(const vcm_dp_desc* d, const void* y, const void* fusion, void* params,
                         float* yq, int16_t* sym, uint8_t* idx, float* y_q_full,
                         float* scales_full, void* stream);
/* q_outer: [N,C] fp32, the outer quantiser the result is multiplied with (video_model.py:297,326) */
int vcm_dual_prior_pass1(const vcm_dp_desc* d, const float* yq, const void* fusion,
                         const void* prior, const void* params, const float* q_outer,
                         void* y_hat, int16_t* sym, uint8_t* idx, float* y_q_full,
                         float* scales_full, void* stream);
/* z_hat = round(z) (video_model.py:284,310): NHWC view -> z_hat NHWC view + int16 [N,C,h,w]  */
int vcm_round_latent(const void* z, void* z_hat, int16_t* sym, int32_t N, int32_t h, int32_t w,
                     int32_t C, int32_t z_cstride, int32_t zhat_cstride, int32_t dtype,
                     void* stream);
/* decoded int16 [N,C,h,w] -> NHWC view (float symbols, entropy_models.py:189-195)           */
int vcm_symbols_to_latent(const int16_t* sym, void* z_hat, int32_t N, int32_t h, int32_t w,
                          int32_t C, int32_t zhat_cstride, int32_t dtype, void* stream);

/* likelihood -> bits (common_model.py:51-73, entropy_models.py:68-73,109-117); deterministic
 * two-stage reduction, out_bits[n] = sum over the sample.  dist: 0 Laplace(sigma>=1e-5),
 * 1 Normal(sigma>=0.11).  `partial` scratch: vcm_bits_partial_count(per_sample)*N floats.    */
int32_t vcm_bits_partial_count(int64_t per_sample);
int vcm_latent_bits(const float* y_q, const float* scales, float* partial, float* out_bits,
                    int32_t N, int64_t per_sample, int32_t dist, void* stream);
/* factorized prior bits of z_hat (NHWC fp32 [N,h,w,C]); fparams: [11][C] = f1.h,f1.b,f1.a,
 * f2.h,f2.b,f2.a,f3.h,f3.b,f3.a,f4.h,f4.b                                                    */
int vcm_factorized_bits(const void* z_hat, const float* fparams, float* partial,
                        float* out_bits, int32_t N, int64_t hw, int32_t C, int32_t cstride,
                        int32_t dtype, void* stream);
/* sum((a-b)^2) per sample over NCHW fp32 tensors (video_model.py:544-545 mse / me_mse)      */
int vcm_sq_err_sum(const float* a, const float* b, float* partial, float* out, int32_t N,
                   int64_t per_sample, void* stream);
/* Allocation-free variant of vcm_sq_err_sum.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  Same kernels and summation order as
 * vcm_sq_err_sum (results are bit-identical), but the partial sums live in a 32 KiB
 * `static __device__` array inside the library (held by every GPU that loads it, until exit)
 * instead of caller scratch; the call itself allocates nothing.
 * Limit: N * ceil(per_sample / 8192) <= 8192 partial sums (64 Mi elements in all, e.g. ten padded
 * 1080p RGB frames); larger requests return 1 with a message in vcm_last_error().
 * All calls on one device share the array: calls that overlap on the GPU (different streams) race.
 * The first call on each device looks the array's address up (cudaGetSymbolAddress); make it
 * outside any stream capture.                                                                 */
int vcm_sq_err_sum_noalloc(const float* a, const float* b, float* out, int32_t N,
                           int64_t per_sample, void* stream);

/* ---- host entropy coder (CPU, GIL-free; bitstream identical to the reference) ----------
 * replaces MLCodec_rans BufferedRansEncoder / RansDecoder
 * (DCVC_HEM/src/cpp/rans/rans_interface.cpp:85-244, rans_interface.hpp:40-91) and
 * MLCodec_CXX.pmf_to_quantized_cdf (DCVC_HEM/src/cpp/ops/ops.cpp:24-82).                     */
typedef struct vcm_rans_enc vcm_rans_enc;
typedef struct vcm_rans_dec vcm_rans_dec;
/* a CDF table set: cdf [ntab][stride] int32, sizes[ntab], offsets[ntab] (copied)            */
typedef struct vcm_cdf_tables vcm_cdf_tables;
vcm_cdf_tables* vcm_cdf_tables_new(const int32_t* cdf, const int32_t* sizes,
                                   const int32_t* offsets, int32_t ntab, int32_t stride);
void vcm_cdf_tables_free(vcm_cdf_tables* t);
vcm_rans_enc* vcm_rans_enc_new(void);
void vcm_rans_enc_free(vcm_rans_enc* e);
void vcm_rans_enc_reset(vcm_rans_enc* e);
/* symbols int16, per-symbol table index uint8 (scale index).  The encoder keeps pointers into
   `t` (precomputed per-symbol records): `t` must stay alive until the next flush / reset.       */
int vcm_rans_enc_push(vcm_rans_enc* e, const int16_t* sym, const uint8_t* idx, int64_t n,
                      const vcm_cdf_tables* t);
/* symbols int16 laid out [C][hw]; table index = channel (BitEstimator, entropy_models.py:176-187) */
int vcm_rans_enc_push_channels(vcm_rans_enc* e, const int16_t* sym, int32_t C, int64_t hw,
                               const vcm_cdf_tables* t);
/* returns byte count; *bytes valid until the next reset/flush/free                          */
int64_t vcm_rans_enc_flush(vcm_rans_enc* e, const uint8_t** bytes);
vcm_rans_dec* vcm_rans_dec_new(void);
void vcm_rans_dec_free(vcm_rans_dec* d);
int vcm_rans_dec_set_stream(vcm_rans_dec* d, const uint8_t* bytes, int64_t nbytes);
int vcm_rans_dec_decode(vcm_rans_dec* d, const uint8_t* idx, int64_t n, const vcm_cdf_tables* t,
                        int16_t* out);
int vcm_rans_dec_decode_channels(vcm_rans_dec* d, int32_t C, int64_t hw,
                                 const vcm_cdf_tables* t, int16_t* out);
/* pmf[n] fp32 -> cdf[n+1] int32 ; 0 ok, 1 = no frequency donor                              */
int vcm_pmf_to_quantized_cdf(const float* pmf, int32_t n, int32_t precision, int32_t* cdf);

#ifdef __cplusplus
}
#endif
#endif /* VCMTS_B200_H */

// vcm_ts_b200/csrc/conv_common.cuh -- parameters + fused epilogue shared by the two conv engines
// (fp32 FMA implicit GEMM in conv_simt.cu, tcgen05 bf16 implicit GEMM in conv_tc.cu).
//
// The epilogue reproduces, in the reference's order of float operations,
//   bias -> LeakyReLU/ReLU -> (+ residual) -> (+ SE-scaled residual) -> (+ second residual)
//   -> PixelShuffle(2) / channel-slice store
// i.e. layers.py:42-127 (ResidualBlock*, subpel_conv), video_net.py:74-96 (ResBlock),
// video_net.py:165-179 (ConvBlockResidual: up_dim(x) + conv(x)*se), video_model.py:40-68.
#pragma once
#include <stdlib.h>

#include "common.cuh"

struct ConvParams {
  const void* x;
  const float* w;      // SIMT: [K*K][Cin][Cout] fp32
  const float* bias;
  const void* res;
  const float* res_scale;
  const void* res2;
  void* y;
  void* y2;
  int N, H, W, Ho, Wo, Cin, Cout, K, stride, pad;
  int in_cs, out_cs, res_cs, res2_cs, out2_cs;
  int flags;
  float pre_slope, post_slope, out2_slope;
  int cin_chunks, vecA, vecB, tiles_x, tiles_y;
};

template <typename T>
__device__ __forceinline__ void conv_epilogue_store(const ConvParams& P, int n, int oy, int ox,
                                                    int co, float v, float bias) {
  v += bias;
  if (P.flags & VCM_CONV_POST_ACT) v = leaky(v, P.post_slope);
  int oc = co, y2 = oy, x2 = ox, Hout = P.Ho, Wout = P.Wo, Cres = P.Cout;
  if (P.flags & VCM_CONV_PIXEL_SHUFFLE) {
    oc = co >> 2;
    y2 = 2 * oy + ((co >> 1) & 1);
    x2 = 2 * ox + (co & 1);
    Hout = 2 * P.Ho;
    Wout = 2 * P.Wo;
    Cres = P.Cout >> 2;
  }
  const size_t pix = ((size_t)n * Hout + y2) * Wout + x2;
  if (P.flags & VCM_CONV_RES)
    v = ld_as_float(reinterpret_cast<const T*>(P.res) + pix * P.res_cs + oc) + v;
  if (P.flags & VCM_CONV_RES_SCALE) {
    const float r = ld_as_float(reinterpret_cast<const T*>(P.res) + pix * P.res_cs + oc);
    v = v + __fmul_rn(r, __ldg(P.res_scale + (size_t)n * Cres + oc));
  }
  if (P.flags & VCM_CONV_RES2)
    v = ld_as_float(reinterpret_cast<const T*>(P.res2) + pix * P.res2_cs + oc) + v;
  st_from_float(reinterpret_cast<T*>(P.y) + pix * P.out_cs + oc, v);
  if (P.flags & VCM_CONV_OUT2)
    st_from_float(reinterpret_cast<T*>(P.y2) + pix * P.out2_cs + oc, leaky(v, P.out2_slope));
}

static inline int conv_fill_params(const vcm_conv_desc* d, ConvParams& P) {
  if (!(d->K == 1 || d->K == 3 || d->K == 7)) return 1;
  if (!(d->stride == 1 || d->stride == 2)) return 1;
  if (d->N <= 0 || d->H <= 0 || d->W <= 0 || d->Cin <= 0 || d->Cout <= 0) return 1;
  if ((d->flags & VCM_CONV_PIXEL_SHUFFLE) && (d->Cout % 4)) return 1;
  if ((d->flags & VCM_CONV_RES) && (d->flags & VCM_CONV_RES_SCALE)) return 1;
  P.N = d->N; P.H = d->H; P.W = d->W;
  P.Cin = d->Cin; P.Cout = d->Cout; P.K = d->K; P.stride = d->stride; P.pad = d->K / 2;
  P.Ho = (d->H + 2 * P.pad - d->K) / d->stride + 1;
  P.Wo = (d->W + 2 * P.pad - d->K) / d->stride + 1;
  P.in_cs = d->in_cstride; P.out_cs = d->out_cstride;
  P.res_cs = d->res_cstride; P.res2_cs = d->res2_cstride; P.out2_cs = d->out2_cstride;
  P.flags = d->flags; P.pre_slope = d->pre_slope; P.post_slope = d->post_slope;
  P.out2_slope = d->out2_slope;
  return 0;
}

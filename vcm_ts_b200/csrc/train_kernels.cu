// vcm_ts_b200/csrc/train_kernels.cu -- backward kernels of the training path (SURVEY 8(a) row A16, 8(f)
// rank 3): the reference trains DMC through torch autograd (core/model/dcvc_hem.py:254-322,
// core/engine/train_multi.py:203-268); here the convolution weight gradient and the backward of
// flow_warp (video_net.py:32-55, grid_sample bilinear / border / align_corners=True) are device kernels
// behind the C-ABI; the data gradient of a convolution is the forward engine run on flipped, transposed
// weights (vcm_ts_b200/training.py).
#include <stdlib.h>

#include "common.cuh"

namespace {

// ------------------------------------------------------------------------------------------ wgrad
// dW[co][ci][ky][kx] = sum over output pixels of dY[pix][co] * X[pix*stride + tap - pad][ci].
// One block = a 64 (co) x 64 (ci) tile of one tap over one slice of the pixels; 256 threads, 4 x 4
// accumulators each, operands staged through shared memory in fp32.  Partials [split][tap][co][ci] are
// summed in fixed order by wgrad_reduce_kernel (deterministic).
constexpr int WG_T = 64, WG_PB = 16;

template <typename TX, typename TY>
__global__ void __launch_bounds__(256)
wgrad_partial_kernel(const TX* __restrict__ x, const TY* __restrict__ dy, float* __restrict__ partial, int N, int H,
                     int W, int Ho, int Wo, int Cin, int Cout, int K, int stride, int pad, int x_cs, int dy_cs,
                     int tiles_ci, int tiles_co, int64_t pix_per_split) {
  __shared__ float sdy[WG_PB][WG_T + 4];
  __shared__ float sx[WG_PB][WG_T + 4];
  int b = blockIdx.x;
  const int tci = b % tiles_ci;
  b /= tiles_ci;
  const int tco = b % tiles_co;
  const int tap = b / tiles_co;
  const int ky = tap / K, kx = tap - ky * K;
  const int co0 = tco * WG_T, ci0 = tci * WG_T;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int64_t npix = (int64_t)N * Ho * Wo;
  const int64_t p_begin = (int64_t)blockIdx.y * pix_per_split;
  const int64_t p_end = p_begin + pix_per_split < npix ? p_begin + pix_per_split : npix;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int64_t p0 = p_begin; p0 < p_end; p0 += WG_PB) {
    // stage 16 pixels x 64 channels of both operands (4 elements per thread each)
    for (int e = threadIdx.x; e < WG_PB * WG_T; e += 256) {
      const int pp = e / WG_T, c = e - pp * WG_T;
      const int64_t p = p0 + pp;
      float vy = 0.f, vx = 0.f;
      if (p < p_end) {
        const int ox = (int)(p % Wo);
        const int64_t q = p / Wo;
        const int oy = (int)(q % Ho);
        const int n = (int)(q / Ho);
        if (co0 + c < Cout) vy = ld_as_float(dy + p * dy_cs + co0 + c);
        const int iy = oy * stride + ky - pad, ix = ox * stride + kx - pad;
        if (ci0 + c < Cin && iy >= 0 && iy < H && ix >= 0 && ix < W)
          vx = ld_as_float(x + (((int64_t)n * H + iy) * W + ix) * x_cs + ci0 + c);
      }
      sdy[pp][c] = vy;
      sx[pp][c] = vx;
    }
    __syncthreads();
#pragma unroll
    for (int pp = 0; pp < WG_PB; ++pp) {
      const float4 a = *reinterpret_cast<const float4*>(&sdy[pp][ty * 4]);
      const float4 bb = *reinterpret_cast<const float4*>(&sx[pp][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
  float* out = partial + ((int64_t)blockIdx.y * K * K + tap) * Cout * Cin;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int co = co0 + ty * 4 + i;
    if (co >= Cout) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int ci = ci0 + tx * 4 + j;
      if (ci < Cin) out[(int64_t)co * Cin + ci] = acc[i][j];
    }
  }
}

__global__ void wgrad_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dw, int nsplit, int taps,
                                    int Cout, int Cin) {
  const int64_t per = (int64_t)taps * Cout * Cin;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < per;
       i += (int64_t)gridDim.x * blockDim.x) {
    // i indexes the OIHW output: ((co * Cin + ci) * taps + tap)
    const int tap = (int)(i % taps);
    const int64_t r = i / taps;
    const int ci = (int)(r % Cin);
    const int co = (int)(r / Cin);
    const int64_t src = ((int64_t)tap * Cout + co) * Cin + ci;
    float s = 0.f;
    for (int k = 0; k < nsplit; ++k) s += partial[(int64_t)k * per + src];
    dw[i] = s;
  }
}

// ------------------------------------------------------------------------------------------ warp backward
__device__ __forceinline__ float lin_pm1(int i, int steps) {
  const float step = __fdiv_rn(2.f, (float)(steps - 1));
  return (i < steps / 2) ? __fadd_rn(-1.f, __fmul_rn(step, (float)i))
                         : __fsub_rn(1.f, __fmul_rn(step, (float)(steps - 1 - i)));
}

// one thread = one output pixel x 4 channels.  gfeat (fp32, zero-initialised) receives the four bilinear
// taps by atomicAdd; gflow (fp32 [N,H,W,2], zero-initialised) the per-thread partial of
// sum_c g_c * d out_c / d (x, y): d x / d u = 1 inside the image and 0 where the sample was clamped to the
// border (ATen clip_coordinates_set_grad), as flow is in pixels and align_corners=True.
template <typename T>
__global__ void flow_warp_bwd_kernel(const T* __restrict__ feat, const float* __restrict__ flow,
                                     const T* __restrict__ gout, float* __restrict__ gfeat, float* __restrict__ gflow,
                                     int N, int H, int W, int C, int feat_cs, int flow_cs, int gout_cs) {
  const int G = (C + 3) / 4;
  const int64_t total = (int64_t)N * H * W * G;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int g = (int)(i % G);
    const int64_t p = i / G;
    const int ox = (int)(p % W);
    const int64_t q = p / W;
    const int oy = (int)(q % H);
    const int n = (int)(q / H);
    const float u = __ldg(flow + p * flow_cs), v = __ldg(flow + p * flow_cs + 1);
    const float sxf = __fdiv_rn((float)W - 1.f, 2.f), syf = __fdiv_rn((float)H - 1.f, 2.f);
    const float gx = __fadd_rn(lin_pm1(ox, W), __fdiv_rn(u, sxf));
    const float gy = __fadd_rn(lin_pm1(oy, H), __fdiv_rn(v, syf));
    const float xr = __fmul_rn(__fadd_rn(gx, 1.f), sxf), yr = __fmul_rn(__fadd_rn(gy, 1.f), syf);
    const float mx = (xr > 0.f && xr < (float)(W - 1)) ? 1.f : 0.f;
    const float my = (yr > 0.f && yr < (float)(H - 1)) ? 1.f : 0.f;
    const float x = fminf((float)(W - 1), fmaxf(xr, 0.f)), y = fminf((float)(H - 1), fmaxf(yr, 0.f));
    const float xw = floorf(x), yn = floorf(y);
    const float w = x - xw, e = 1.f - w, nn = y - yn, s = 1.f - nn;
    const int x0 = (int)xw, y0 = (int)yn;
    const bool x1ok = x0 + 1 <= W - 1, y1ok = y0 + 1 <= H - 1;
    const int x1 = x1ok ? x0 + 1 : W - 1, y1 = y1ok ? y0 + 1 : H - 1;
    const int64_t base = (int64_t)n * H * W;
    const int64_t i00 = base + (int64_t)y0 * W + x0, i01 = base + (int64_t)y0 * W + x1;
    const int64_t i10 = base + (int64_t)y1 * W + x0, i11 = base + (int64_t)y1 * W + x1;
    float dgx = 0.f, dgy = 0.f;
    for (int k = 0; k < 4; ++k) {
      const int c = g * 4 + k;
      if (c >= C) break;
      const float go = ld_as_float(gout + p * gout_cs + c);
      const float a = ld_as_float(feat + i00 * feat_cs + c);
      const float b = x1ok ? ld_as_float(feat + i01 * feat_cs + c) : 0.f;
      const float cc = y1ok ? ld_as_float(feat + i10 * feat_cs + c) : 0.f;
      const float d = (x1ok && y1ok) ? ld_as_float(feat + i11 * feat_cs + c) : 0.f;
      atomicAdd(gfeat + i00 * C + c, go * s * e);
      if (x1ok) atomicAdd(gfeat + i01 * C + c, go * s * w);
      if (y1ok) atomicAdd(gfeat + i10 * C + c, go * nn * e);
      if (x1ok && y1ok) atomicAdd(gfeat + i11 * C + c, go * nn * w);
      dgx += go * (s * (b - a) + nn * (d - cc));
      dgy += go * (e * (cc - a) + w * (d - b));
    }
    atomicAdd(gflow + p * 2, dgx * mx);
    atomicAdd(gflow + p * 2 + 1, dgy * my);
  }
}

inline int grid_for(int64_t total) {
  int64_t g = (total + 255) / 256;
  return (int)(g < 148 * 32 ? (g < 1 ? 1 : g) : 148 * 32);
}

}  // namespace

// tensor-core path (wgrad_tc.cu)
int64_t vcm_wgrad_tc_scratch_floats(const vcm_conv_desc* d);
int vcm_wgrad_tc_launch(const vcm_conv_desc* d, const void* x, const void* dy, float* scratch, int* nsplit_out,
                        cudaStream_t stream);

static int64_t simt_scratch_floats(const vcm_conv_desc* d) {
  const int pad = d->K / 2;
  const int Ho = (d->H + 2 * pad - d->K) / d->stride + 1, Wo = (d->W + 2 * pad - d->K) / d->stride + 1;
  const int64_t npix = (int64_t)d->N * Ho * Wo;
  const int tiles = cdiv(d->Cout, WG_T) * cdiv(d->Cin, WG_T) * d->K * d->K;
  int nsplit = cdiv(148 * 4, tiles);
  const int max_split = cdiv(npix, 4 * WG_PB);
  if (nsplit > max_split) nsplit = max_split;
  if (nsplit < 1) nsplit = 1;
  return (int64_t)nsplit * d->K * d->K * d->Cout * d->Cin;
}

extern "C" int64_t vcm_conv2d_wgrad_scratch_floats(const vcm_conv_desc* d) {
  if (!d) return 0;
  const int64_t a = simt_scratch_floats(d), b = vcm_wgrad_tc_scratch_floats(d);
  return a > b ? a : b;
}

extern "C" int vcm_conv2d_wgrad(const vcm_conv_desc* d, const void* x, const void* dy, float* dw_oihw,
                                float* scratch, void* stream) {
  VCM_CHECK_ARG(d && x && dy && dw_oihw && scratch, "null argument");
  VCM_CHECK_ARG((d->K == 1 || d->K == 3 || d->K == 7) && (d->stride == 1 || d->stride == 2), "unsupported conv");
  const int pad = d->K / 2;
  const int Ho = (d->H + 2 * pad - d->K) / d->stride + 1, Wo = (d->W + 2 * pad - d->K) / d->stride + 1;
  const int64_t npix = (int64_t)d->N * Ho * Wo;
  const int taps = d->K * d->K;
  const int tiles_ci = cdiv(d->Cin, WG_T), tiles_co = cdiv(d->Cout, WG_T);
  const int64_t per = (int64_t)taps * d->Cout * d->Cin;
  cudaStream_t s = (cudaStream_t)stream;
  {   // bf16 operands with 16-aligned Cin: implicit GEMM over pixels on tcgen05 (wgrad_tc.cu)
    static const bool no_tc = getenv("VCM_WGRAD_SIMT") && getenv("VCM_WGRAD_SIMT")[0] == '1';
    int ns = 0;
    const int rc = no_tc ? 3 : vcm_wgrad_tc_launch(d, x, dy, scratch, &ns, s);
    if (rc == 1) return 1;
    if (rc == 0) {
      wgrad_reduce_kernel<<<grid_for(per), 256, 0, s>>>(scratch, dw_oihw, ns, taps, d->Cout, d->Cin);
      VCM_CHECK_LAUNCH();
      return 0;
    }
  }
  const int nsplit = (int)(simt_scratch_floats(d) / per);
  int64_t pps = (npix + nsplit - 1) / nsplit;
  pps = (pps + WG_PB - 1) / WG_PB * WG_PB;
  dim3 grid(tiles_ci * tiles_co * taps, nsplit);
#define WG_LAUNCH(TX, TY)                                                                                         \
  wgrad_partial_kernel<TX, TY><<<grid, 256, 0, s>>>((const TX*)x, (const TY*)dy, scratch, d->N, d->H, d->W, Ho, Wo, \
                                                    d->Cin, d->Cout, d->K, d->stride, pad, d->in_cstride,          \
                                                    d->out_cstride, tiles_ci, tiles_co, pps)
  if (d->in_dtype == VCM_F32 && d->out_dtype == VCM_F32) WG_LAUNCH(float, float);
  else if (d->in_dtype == VCM_BF16 && d->out_dtype == VCM_BF16) WG_LAUNCH(__nv_bfloat16, __nv_bfloat16);
  else if (d->in_dtype == VCM_F32 && d->out_dtype == VCM_BF16) WG_LAUNCH(float, __nv_bfloat16);
  else if (d->in_dtype == VCM_BF16 && d->out_dtype == VCM_F32) WG_LAUNCH(__nv_bfloat16, float);
  else {
    vcm_set_error("vcm_conv2d_wgrad: bad dtype");
    return 1;
  }
#undef WG_LAUNCH
  VCM_CHECK_LAUNCH();
  wgrad_reduce_kernel<<<grid_for(per), 256, 0, s>>>(scratch, dw_oihw, nsplit, taps, d->Cout, d->Cin);
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_flow_warp_bwd(const void* feat, const float* flow, const void* gout, float* gfeat, float* gflow,
                                 int32_t N, int32_t H, int32_t W, int32_t C, int32_t feat_cstride,
                                 int32_t flow_cstride, int32_t gout_cstride, int32_t dtype, void* stream) {
  VCM_CHECK_ARG(feat && flow && gout && gfeat && gflow && N > 0 && H > 1 && W > 1 && C > 0, "bad argument");
  const int64_t total = (int64_t)N * H * W * ((C + 3) / 4);
  cudaStream_t s = (cudaStream_t)stream;
  VCM_DISPATCH_DTYPE(dtype, {
    flow_warp_bwd_kernel<T><<<grid_for(total), 256, 0, s>>>((const T*)feat, flow, (const T*)gout, gfeat, gflow, N, H,
                                                           W, C, feat_cstride, flow_cstride, gout_cstride);
  });
  VCM_CHECK_LAUNCH();
  return 0;
}

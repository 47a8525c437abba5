// vcm_ts_b200/csrc/motion.cu -- motion glue: SpyNet pyramid pooling, flow x2 up / half down,
// and the bilinear backward warp.  All HBM-bound gather/stream kernels over NHWC views.
//
//   vcm_avgpool2    F.avg_pool2d(x, 2, 2)                        video_net.py:131-133
//   vcm_flow_up2    bilinearupsacling(flow) * gain               video_net.py:58-63, :140
//   vcm_flow_down2  bilineardownsacling(mv) * gain (gain = 1/2)  video_net.py:66-71, video_model.py:237-238
//   vcm_flow_warp   torch_warp: grid_sample(bilinear, border, align_corners=True) at
//                   linspace(-1,1) + flow/((size-1)/2)           video_net.py:32-55
//
// The warp restates ATen's fp32 arithmetic step by step (normalise -> unnormalise -> clip ->
// floor -> 4 taps) instead of the algebraically equal x + u, so border clipping and sub-pixel
// weights round the way the reference's do.
#include "common.cuh"

namespace {

inline int ew_grid(int64_t total) {
  int64_t g = (total + 255) / 256;
  return (int)(g > 148 * 16 ? 148 * 16 : (g < 1 ? 1 : g));
}

__global__ void avgpool2_kernel(const float* __restrict__ src, float* __restrict__ dst, int N, int H,
                                int W, int C, int src_cs, int dst_cs) {
  const int Ho = H / 2, Wo = W / 2;
  const int64_t total = (int64_t)N * Ho * Wo * C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    int64_t p = i / C;
    const int ox = (int)(p % Wo);
    p /= Wo;
    const int oy = (int)(p % Ho);
    const int n = (int)(p / Ho);
    const float* s = src + (((int64_t)n * H + 2 * oy) * W + 2 * ox) * src_cs + c;
    float v = __ldg(s);
    v = __fadd_rn(v, __ldg(s + src_cs));
    v = __fadd_rn(v, __ldg(s + (int64_t)W * src_cs));
    v = __fadd_rn(v, __ldg(s + (int64_t)(W + 1) * src_cs));
    dst[(((int64_t)n * Ho + oy) * Wo + ox) * dst_cs + c] = __fdiv_rn(v, 4.f);
  }
}

// align_corners=False x2 upsample of a 2-channel field: src index = d/2 - 0.25 clamped at 0
__global__ void flow_up2_kernel(const float* __restrict__ src, float* __restrict__ dst, int N, int H,
                                int W, int src_cs, int dst_cs, float gain) {
  const int Ho = 2 * H, Wo = 2 * W;
  const int64_t total = (int64_t)N * Ho * Wo;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int ox = (int)(i % Wo);
    int64_t p = i / Wo;
    const int oy = (int)(p % Ho);
    const int n = (int)(p / Ho);
    float sy = fmaxf((oy + 0.5f) * 0.5f - 0.5f, 0.f);
    float sx = fmaxf((ox + 0.5f) * 0.5f - 0.5f, 0.f);
    const int y0 = (int)sy, x0 = (int)sx;
    const int y1 = y0 + (y0 < H - 1 ? 1 : 0), x1 = x0 + (x0 < W - 1 ? 1 : 0);
    const float ly = sy - y0, lx = sx - x0;
    const float hy = 1.f - ly, hx = 1.f - lx;
    const float* b = src + (int64_t)n * H * W * src_cs;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const float p00 = __ldg(b + ((int64_t)y0 * W + x0) * src_cs + c);
      const float p01 = __ldg(b + ((int64_t)y0 * W + x1) * src_cs + c);
      const float p10 = __ldg(b + ((int64_t)y1 * W + x0) * src_cs + c);
      const float p11 = __ldg(b + ((int64_t)y1 * W + x1) * src_cs + c);
      const float top = __fadd_rn(__fmul_rn(hx, p00), __fmul_rn(lx, p01));
      const float bot = __fadd_rn(__fmul_rn(hx, p10), __fmul_rn(lx, p11));
      const float v = __fadd_rn(__fmul_rn(hy, top), __fmul_rn(ly, bot));
      dst[i * dst_cs + c] = __fmul_rn(v, gain);
    }
  }
}

// align_corners=False half-size bilinear of an even-sized field == 2x2 mean, then * gain
__global__ void flow_down2_kernel(const float* __restrict__ src, float* __restrict__ dst, int N,
                                  int H, int W, int src_cs, int dst_cs, float gain) {
  const int Ho = H / 2, Wo = W / 2;
  const int64_t total = (int64_t)N * Ho * Wo;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int ox = (int)(i % Wo);
    int64_t p = i / Wo;
    const int oy = (int)(p % Ho);
    const int n = (int)(p / Ho);
    const float* s = src + (((int64_t)n * H + 2 * oy) * W + 2 * ox) * src_cs;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const float p00 = __ldg(s + c), p01 = __ldg(s + src_cs + c);
      const float p10 = __ldg(s + (int64_t)W * src_cs + c), p11 = __ldg(s + (int64_t)(W + 1) * src_cs + c);
      const float top = __fadd_rn(__fmul_rn(0.5f, p00), __fmul_rn(0.5f, p01));
      const float bot = __fadd_rn(__fmul_rn(0.5f, p10), __fmul_rn(0.5f, p11));
      const float v = __fadd_rn(__fmul_rn(0.5f, top), __fmul_rn(0.5f, bot));
      dst[i * dst_cs + c] = __fmul_rn(v, gain);
    }
  }
}

// torch.linspace(-1, 1, steps)[i] in fp32 (two-sided evaluation, RangeFactories.cpp)
__device__ __forceinline__ float linspace_pm1(int i, int steps) {
  const float step = __fdiv_rn(2.f, (float)(steps - 1));
  return (i < steps / 2) ? __fadd_rn(-1.f, __fmul_rn(step, (float)i))
                         : __fsub_rn(1.f, __fmul_rn(step, (float)(steps - 1 - i)));
}

struct WarpTaps {
  int x0, x1, y0, y1;
  float nw, ne, sw, se;
  bool x1_ok, y1_ok;
};

__device__ __forceinline__ WarpTaps warp_taps(float u, float v, int ox, int oy, int W, int H) {
  const float sxf = __fdiv_rn((float)W - 1.f, 2.f), syf = __fdiv_rn((float)H - 1.f, 2.f);
  const float gx = __fadd_rn(linspace_pm1(ox, W), __fdiv_rn(u, sxf));
  const float gy = __fadd_rn(linspace_pm1(oy, H), __fdiv_rn(v, syf));
  float x = __fmul_rn(__fadd_rn(gx, 1.f), sxf);
  float y = __fmul_rn(__fadd_rn(gy, 1.f), syf);
  x = fminf((float)(W - 1), fmaxf(x, 0.f));
  y = fminf((float)(H - 1), fmaxf(y, 0.f));
  const float xw = floorf(x), yn = floorf(y);
  const float w = __fsub_rn(x, xw), e = __fsub_rn(1.f, w);
  const float nn = __fsub_rn(y, yn), s = __fsub_rn(1.f, nn);
  WarpTaps t;
  t.x0 = (int)xw; t.y0 = (int)yn;
  t.x1 = t.x0 + 1; t.y1 = t.y0 + 1;
  t.x1_ok = t.x1 <= W - 1; t.y1_ok = t.y1 <= H - 1;
  if (!t.x1_ok) t.x1 = W - 1;
  if (!t.y1_ok) t.y1 = H - 1;
  t.nw = __fmul_rn(s, e); t.ne = __fmul_rn(s, w);
  t.sw = __fmul_rn(nn, e); t.se = __fmul_rn(nn, w);
  return t;
}

__device__ __forceinline__ float warp_mix(const WarpTaps& t, float a, float b, float c, float d) {
  // out-of-range east / south taps contribute exactly 0 (masked gather in ATen)
  const float t0 = __fmul_rn(a, t.nw);
  const float t1 = t.x1_ok ? __fmul_rn(b, t.ne) : 0.f;
  const float t2 = t.y1_ok ? __fmul_rn(c, t.sw) : 0.f;
  const float t3 = (t.x1_ok && t.y1_ok) ? __fmul_rn(d, t.se) : 0.f;
  return __fadd_rn(__fadd_rn(__fadd_rn(t0, t1), t2), t3);
}

// one thread = one pixel x 4 channels (VEC) or x 1 channel
template <typename T, bool VEC>
__global__ void flow_warp_kernel(const T* __restrict__ feat, const float* __restrict__ flow,
                                 T* __restrict__ out, int N, int H, int W, int C, int feat_cs,
                                 int flow_cs, int out_cs) {
  const int G = VEC ? C / 4 : C;
  const int64_t total = (int64_t)N * H * W * G;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int g = (int)(i % G);
    const int64_t p = i / G;
    const int ox = (int)(p % W);
    const int64_t q = p / W;
    const int oy = (int)(q % H);
    const int n = (int)(q / H);
    const float2 uv = make_float2(__ldg(flow + p * flow_cs), __ldg(flow + p * flow_cs + 1));
    const WarpTaps t = warp_taps(uv.x, uv.y, ox, oy, W, H);
    const T* b = feat + (int64_t)n * H * W * feat_cs;
    const int c = VEC ? g * 4 : g;
    const T* p00 = b + ((int64_t)t.y0 * W + t.x0) * feat_cs + c;
    const T* p01 = b + ((int64_t)t.y0 * W + t.x1) * feat_cs + c;
    const T* p10 = b + ((int64_t)t.y1 * W + t.x0) * feat_cs + c;
    const T* p11 = b + ((int64_t)t.y1 * W + t.x1) * feat_cs + c;
    if (VEC) {
      const float4 a = ld4<T>(p00), bb = ld4<T>(p01), cc = ld4<T>(p10), dd = ld4<T>(p11);
      float4 r;
      r.x = warp_mix(t, a.x, bb.x, cc.x, dd.x);
      r.y = warp_mix(t, a.y, bb.y, cc.y, dd.y);
      r.z = warp_mix(t, a.z, bb.z, cc.z, dd.z);
      r.w = warp_mix(t, a.w, bb.w, cc.w, dd.w);
      st4<T>(out + p * out_cs + c, r);
    } else {
      st_from_float(out + p * out_cs + c,
                    warp_mix(t, ld_as_float(p00), ld_as_float(p01), ld_as_float(p10), ld_as_float(p11)));
    }
  }
}

// bf16, 8 channels (16 bytes) per thread: same per-channel arithmetic, half the instructions and
// twice the bytes in flight per thread of the 4-channel variant
// A block owns an 8 x 2*(256 / G / 8) pixel patch (G = C / 8 threads per pixel), not a run of a row: the four
// taps of neighbouring output pixels overlap in both directions, so most of the 4x gather amplification is
// served by L1 instead of L2.  grid = (tiles_x, tiles_y, N).
__global__ void __launch_bounds__(256) flow_warp8_bf16_kernel(const __nv_bfloat16* __restrict__ feat,
                                                              const float* __restrict__ flow,
                                                              __nv_bfloat16* __restrict__ out, int N, int H, int W,
                                                              int C, int feat_cs, int flow_cs, int out_cs) {
  const int G = C / 8;
  const int th = 256 / G / 8;
  constexpr int PP = 2;   // pixels per thread (rows oy and oy + th): both flow loads, then all eight tap loads
  const int g = threadIdx.x % G, pix = threadIdx.x / G;
  const int ox_raw = blockIdx.x * 8 + (pix & 7);
  const int n = blockIdx.z;
  const bool x_in = ox_raw < W;                  // no early exit: the tap shuffles below need whole warps
  const int ox = x_in ? ox_raw : W - 1;
  const __nv_bfloat16* b = feat + (int64_t)n * H * W * feat_cs + g * 8;
  int oy[PP];
  int64_t p[PP];
#pragma unroll
  for (int j = 0; j < PP; ++j) {
    oy[j] = blockIdx.y * (PP * th) + j * th + (pix >> 3);
    p[j] = ((int64_t)n * H + min(oy[j], H - 1)) * W + ox;
  }
  WarpTaps t[PP];
  uint4 v[PP][4];
  // The G lanes of a pixel are adjacent.  The coordinate arithmetic (four correctly rounded divisions among ~70
  // instructions) is done ONCE per (pixel, row) and shuffled to the lanes that need it.  With G a power of two
  // >= 2 a warp holds 32 / G pixels x PP rows <= 32 tap sets: lane L computes the set of (pixel L % PW, row
  // L / PW) -- all sets of the warp in one pass instead of one divergent pass per row by every G-th lane
  // (ncu r02: this kernel is issue-bound, ~350 instructions per warp, 140 of them those two passes).
  const bool coop = G > 1 && (G & (G - 1)) == 0 && G <= 32;
  if (coop) {
    const int PW = 32 / G;                               // pixels per warp
    const int lane = threadIdx.x & 31;
    const int wpix0 = (threadIdx.x >> 5) * PW;           // first pixel slot of this warp in the block
    WarpTaps mine;
    mine.x0 = mine.y0 = 0; mine.nw = mine.ne = mine.sw = mine.se = 0.f;
    if (lane < PW * PP) {
      const int pl = wpix0 + lane % PW, jl = lane / PW;
      const int oxl = min(blockIdx.x * 8 + (pl & 7), W - 1);
      const int oyl = min((int)(blockIdx.y * (PP * th) + jl * th + (pl >> 3)), H - 1);
      const float2 uvl = __ldg(reinterpret_cast<const float2*>(flow + (((int64_t)n * H + oyl) * W + oxl) * flow_cs));
      mine = warp_taps(uvl.x, uvl.y, oxl, oyl, W, H);
    }
#pragma unroll
    for (int j = 0; j < PP; ++j) {
      const int src = (lane / G) + j * PW;
      t[j].x0 = __shfl_sync(0xffffffffu, mine.x0, src);
      t[j].y0 = __shfl_sync(0xffffffffu, mine.y0, src);
      t[j].nw = __shfl_sync(0xffffffffu, mine.nw, src);
      t[j].ne = __shfl_sync(0xffffffffu, mine.ne, src);
      t[j].sw = __shfl_sync(0xffffffffu, mine.sw, src);
      t[j].se = __shfl_sync(0xffffffffu, mine.se, src);
      t[j].x1_ok = t[j].x0 + 1 <= W - 1;
      t[j].y1_ok = t[j].y0 + 1 <= H - 1;
      t[j].x1 = t[j].x1_ok ? t[j].x0 + 1 : W - 1;
      t[j].y1 = t[j].y1_ok ? t[j].y0 + 1 : H - 1;
    }
  } else {
#pragma unroll
    for (int j = 0; j < PP; ++j) {
      const float2 uv = __ldg(reinterpret_cast<const float2*>(flow + p[j] * flow_cs));
      t[j] = warp_taps(uv.x, uv.y, ox, min(oy[j], H - 1), W, H);
    }
  }
#pragma unroll
  for (int j = 0; j < PP; ++j) {
    v[j][0] = __ldg(reinterpret_cast<const uint4*>(b + ((int64_t)t[j].y0 * W + t[j].x0) * feat_cs));
    v[j][1] = __ldg(reinterpret_cast<const uint4*>(b + ((int64_t)t[j].y0 * W + t[j].x1) * feat_cs));
    v[j][2] = __ldg(reinterpret_cast<const uint4*>(b + ((int64_t)t[j].y1 * W + t[j].x0) * feat_cs));
    v[j][3] = __ldg(reinterpret_cast<const uint4*>(b + ((int64_t)t[j].y1 * W + t[j].x1) * feat_cs));
  }
#pragma unroll
  for (int j = 0; j < PP; ++j) {
    if (oy[j] >= H || !x_in) continue;
    // ncu (r02): this kernel was instruction-issue bound (issue slots 81 % busy, DRAM 31 %).  The result is
    // rounded to bf16 anyway, so the four taps are combined with one multiply + three FMAs per channel (masked
    // taps get weight 0) instead of the seven separately rounded operations of the fp32 kernel.
    const float wnw = t[j].nw, wne = t[j].x1_ok ? t[j].ne : 0.f, wsw = t[j].y1_ok ? t[j].sw : 0.f,
                wse = (t[j].x1_ok && t[j].y1_ok) ? t[j].se : 0.f;
    const __nv_bfloat162* h00 = reinterpret_cast<const __nv_bfloat162*>(&v[j][0]);
    const __nv_bfloat162* h01 = reinterpret_cast<const __nv_bfloat162*>(&v[j][1]);
    const __nv_bfloat162* h10 = reinterpret_cast<const __nv_bfloat162*>(&v[j][2]);
    const __nv_bfloat162* h11 = reinterpret_cast<const __nv_bfloat162*>(&v[j][3]);
    uint4 r;
    __nv_bfloat162* hr = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 a = __bfloat1622float2(h00[k]), bb = __bfloat1622float2(h01[k]);
      const float2 cc = __bfloat1622float2(h10[k]), dd = __bfloat1622float2(h11[k]);
      hr[k] = __floats2bfloat162_rn(fmaf(dd.x, wse, fmaf(cc.x, wsw, fmaf(bb.x, wne, a.x * wnw))),
                                    fmaf(dd.y, wse, fmaf(cc.y, wsw, fmaf(bb.y, wne, a.y * wnw))));
    }
    *reinterpret_cast<uint4*>(out + p[j] * out_cs + g * 8) = r;
  }
}

// ------------------------------------------------------------------------------------------------
// Fused SpyNet glue (video_net.py:118-146).  Same arithmetic, statement by statement, as the separate
// kernels above, so the results are bit-identical to avgpool2 / flow_up2 / flow_warp chains.
//
// spynet_pyramid: both NCHW fp32 frames -> the three coarser levels of both image pyramids plus the NHWC
// level-0 copies, in ONE pass: a thread owns one 8x8 block of one colour plane of one image, i.e. one pixel
// of the coarsest level (avg_pool2d applied three times, each level from the rounded level below).
//   im1 level k -> lvl[k] channel c (8-channel level-input buffers), im2 level k -> im2[k] (3 channels)
struct PyrOut {
  float* lvl[4];
  float* im2[4];
};

__global__ void spynet_pyramid_kernel(const float* __restrict__ x, const float* __restrict__ ref, PyrOut O, int N,
                                      int H, int W) {
  const int H3 = H / 8, W3 = W / 8;
  const int64_t total = (int64_t)N * 2 * 3 * H3 * W3;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int bx = (int)(i % W3);
    int64_t r = i / W3;
    const int by = (int)(r % H3);
    r /= H3;
    const int c = (int)(r % 3);
    r /= 3;
    const int img = (int)(r % 2);
    const int n = (int)(r / 2);
    const float* src = (img == 0 ? x : ref) + ((int64_t)n * 3 + c) * H * W + (int64_t)(8 * by) * W + 8 * bx;
    float l0[8][8];
#pragma unroll
    for (int yy = 0; yy < 8; ++yy) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(src + (int64_t)yy * W));
      const float4 b = __ldg(reinterpret_cast<const float4*>(src + (int64_t)yy * W + 4));
      l0[yy][0] = a.x; l0[yy][1] = a.y; l0[yy][2] = a.z; l0[yy][3] = a.w;
      l0[yy][4] = b.x; l0[yy][5] = b.y; l0[yy][6] = b.z; l0[yy][7] = b.w;
    }
    float* d0 = img == 0 ? O.lvl[0] : O.im2[0];
    const int cs = img == 0 ? 8 : 3;
#pragma unroll
    for (int yy = 0; yy < 8; ++yy)
#pragma unroll
      for (int xx = 0; xx < 8; ++xx)
        d0[(((int64_t)n * H + 8 * by + yy) * W + 8 * bx + xx) * cs + c] = l0[yy][xx];
    float l1[4][4], l2[2][2];
#pragma unroll
    for (int yy = 0; yy < 4; ++yy)
#pragma unroll
      for (int xx = 0; xx < 4; ++xx) {
        float v = l0[2 * yy][2 * xx];
        v = __fadd_rn(v, l0[2 * yy][2 * xx + 1]);
        v = __fadd_rn(v, l0[2 * yy + 1][2 * xx]);
        v = __fadd_rn(v, l0[2 * yy + 1][2 * xx + 1]);
        l1[yy][xx] = __fdiv_rn(v, 4.f);
      }
    float* d1 = img == 0 ? O.lvl[1] : O.im2[1];
#pragma unroll
    for (int yy = 0; yy < 4; ++yy)
#pragma unroll
      for (int xx = 0; xx < 4; ++xx)
        d1[(((int64_t)n * (H / 2) + 4 * by + yy) * (W / 2) + 4 * bx + xx) * cs + c] = l1[yy][xx];
#pragma unroll
    for (int yy = 0; yy < 2; ++yy)
#pragma unroll
      for (int xx = 0; xx < 2; ++xx) {
        float v = l1[2 * yy][2 * xx];
        v = __fadd_rn(v, l1[2 * yy][2 * xx + 1]);
        v = __fadd_rn(v, l1[2 * yy + 1][2 * xx]);
        v = __fadd_rn(v, l1[2 * yy + 1][2 * xx + 1]);
        l2[yy][xx] = __fdiv_rn(v, 4.f);
      }
    float* d2 = img == 0 ? O.lvl[2] : O.im2[2];
#pragma unroll
    for (int yy = 0; yy < 2; ++yy)
#pragma unroll
      for (int xx = 0; xx < 2; ++xx)
        d2[(((int64_t)n * (H / 4) + 2 * by + yy) * (W / 4) + 2 * bx + xx) * cs + c] = l2[yy][xx];
    float v = l2[0][0];
    v = __fadd_rn(v, l2[0][1]);
    v = __fadd_rn(v, l2[1][0]);
    v = __fadd_rn(v, l2[1][1]);
    float* d3 = img == 0 ? O.lvl[3] : O.im2[3];
    d3[(((int64_t)n * H3 + by) * W3 + bx) * cs + c] = __fdiv_rn(v, 4.f);
  }
}

// spynet_level_input: one level's network input in one pass (video_net.py:139-146):
//   flow_up = bilinear_x2(flow_prev) * 2 -> channels 6,7 ; warp(im2_level, flow_up) -> channels 3..5
// of the 8-channel fp32 level buffer whose channels 0..2 already hold im1_level, and (buf16 != null) the
// bf16 copy of all eight channels padded to 16 that the tcgen05 7x7 conv reads.  One thread per pixel.
__global__ void spynet_level_input_kernel(const float* __restrict__ flow_prev, const float* __restrict__ im2,
                                          float* __restrict__ lvl, __nv_bfloat16* __restrict__ buf16, int N, int h,
                                          int w, int prev_cs) {
  const int hp = h / 2, wp = w / 2;
  const int64_t total = (int64_t)N * h * w;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int ox = (int)(i % w);
    int64_t p = i / w;
    const int oy = (int)(p % h);
    const int n = (int)(p / h);
    // ---- flow_up2 (gain 2)
    float sy = fmaxf((oy + 0.5f) * 0.5f - 0.5f, 0.f);
    float sx = fmaxf((ox + 0.5f) * 0.5f - 0.5f, 0.f);
    const int y0 = (int)sy, x0 = (int)sx;
    const int y1 = y0 + (y0 < hp - 1 ? 1 : 0), x1 = x0 + (x0 < wp - 1 ? 1 : 0);
    const float ly = sy - y0, lx = sx - x0;
    const float hy = 1.f - ly, hx = 1.f - lx;
    const float* b = flow_prev + (int64_t)n * hp * wp * prev_cs;
    float uv[2];
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const float p00 = __ldg(b + ((int64_t)y0 * wp + x0) * prev_cs + c);
      const float p01 = __ldg(b + ((int64_t)y0 * wp + x1) * prev_cs + c);
      const float p10 = __ldg(b + ((int64_t)y1 * wp + x0) * prev_cs + c);
      const float p11 = __ldg(b + ((int64_t)y1 * wp + x1) * prev_cs + c);
      const float top = __fadd_rn(__fmul_rn(hx, p00), __fmul_rn(lx, p01));
      const float bot = __fadd_rn(__fmul_rn(hx, p10), __fmul_rn(lx, p11));
      uv[c] = __fmul_rn(__fadd_rn(__fmul_rn(hy, top), __fmul_rn(ly, bot)), 2.f);
    }
    // ---- warp of the 3-channel reference level
    const WarpTaps t = warp_taps(uv[0], uv[1], ox, oy, w, h);
    const float* f = im2 + (int64_t)n * h * w * 3;
    float o[8];
    float* L = lvl + i * 8;
    o[0] = L[0]; o[1] = L[1]; o[2] = L[2];
#pragma unroll
    for (int c = 0; c < 3; ++c)
      o[3 + c] = warp_mix(t, __ldg(f + ((int64_t)t.y0 * w + t.x0) * 3 + c), __ldg(f + ((int64_t)t.y0 * w + t.x1) * 3 + c),
                          __ldg(f + ((int64_t)t.y1 * w + t.x0) * 3 + c), __ldg(f + ((int64_t)t.y1 * w + t.x1) * 3 + c));
    o[6] = uv[0];
    o[7] = uv[1];
    *reinterpret_cast<float4*>(L + 4) = make_float4(o[4], o[5], o[6], o[7]);
    L[3] = o[3];
    if (buf16) {
      __nv_bfloat162 h2[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) h2[k] = __floats2bfloat162_rn(o[2 * k], o[2 * k + 1]);
      uint4* d = reinterpret_cast<uint4*>(buf16 + i * 16);
      d[0] = *reinterpret_cast<uint4*>(h2);
      d[1] = make_uint4(0u, 0u, 0u, 0u);
    }
  }
}

}  // namespace

extern "C" int vcm_spynet_pyramid(const float* x, const float* ref, float* const* lvl, float* const* im2, int32_t N,
                                  int32_t H, int32_t W, void* stream) {
  VCM_CHECK_ARG(x && ref && lvl && im2 && N > 0 && H % 8 == 0 && W % 8 == 0, "bad argument (H, W multiples of 8)");
  VCM_CHECK_ARG(((uintptr_t)x & 15) == 0 && ((uintptr_t)ref & 15) == 0, "frames must be 16-byte aligned");
  PyrOut O;
  for (int k = 0; k < 4; ++k) {
    VCM_CHECK_ARG(lvl[k] && im2[k], "null level buffer");
    O.lvl[k] = lvl[k];
    O.im2[k] = im2[k];
  }
  const int64_t total = (int64_t)N * 6 * (H / 8) * (W / 8);
  spynet_pyramid_kernel<<<ew_grid(total), 128, 0, (cudaStream_t)stream>>>(x, ref, O, N, H, W);
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_spynet_level_input(const float* flow_prev, const float* im2, float* lvl, void* buf16, int32_t N,
                                      int32_t h, int32_t w, int32_t prev_cstride, void* stream) {
  VCM_CHECK_ARG(flow_prev && im2 && lvl && N > 0 && h > 1 && w > 1 && h % 2 == 0 && w % 2 == 0, "bad argument");
  VCM_CHECK_ARG(((uintptr_t)lvl & 15) == 0 && ((uintptr_t)buf16 & 15) == 0, "level buffers must be 16-byte aligned");
  const int64_t total = (int64_t)N * h * w;
  spynet_level_input_kernel<<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(
      flow_prev, im2, lvl, (__nv_bfloat16*)buf16, N, h, w, prev_cstride);
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_avgpool2(const float* src, float* dst, int32_t N, int32_t H, int32_t W, int32_t C,
                            int32_t src_cstride, int32_t dst_cstride, void* stream) {
  VCM_CHECK_ARG(src && dst && N > 0 && H > 1 && W > 1 && C > 0, "bad argument");
  const int64_t total = (int64_t)N * (H / 2) * (W / 2) * C;
  avgpool2_kernel<<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(src, dst, N, H, W, C, src_cstride,
                                                                    dst_cstride);
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_flow_up2(const float* src, float* dst, int32_t N, int32_t H, int32_t W,
                            int32_t src_cstride, int32_t dst_cstride, float gain, void* stream) {
  VCM_CHECK_ARG(src && dst && N > 0 && H > 0 && W > 0, "bad argument");
  const int64_t total = (int64_t)N * H * W * 4;
  flow_up2_kernel<<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(src, dst, N, H, W, src_cstride,
                                                                    dst_cstride, gain);
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_flow_down2(const float* src, float* dst, int32_t N, int32_t H, int32_t W,
                              int32_t src_cstride, int32_t dst_cstride, float gain, void* stream) {
  VCM_CHECK_ARG(src && dst && N > 0 && H > 1 && W > 1 && H % 2 == 0 && W % 2 == 0, "bad argument");
  const int64_t total = (int64_t)N * (H / 2) * (W / 2);
  flow_down2_kernel<<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(src, dst, N, H, W, src_cstride,
                                                                      dst_cstride, gain);
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_flow_warp(const void* feat, const float* flow, void* out, int32_t N, int32_t H,
                             int32_t W, int32_t C, int32_t feat_cstride, int32_t flow_cstride,
                             int32_t out_cstride, int32_t dtype, void* stream) {
  VCM_CHECK_ARG(feat && flow && out && N > 0 && H > 1 && W > 1 && C > 0, "bad argument");
  const int esz = dtype == VCM_F32 ? 4 : 2;
  const bool vec = (C % 4 == 0) && (feat_cstride % 4 == 0) && (out_cstride % 4 == 0) &&
                   ((uintptr_t)feat % (4 * esz) == 0) && ((uintptr_t)out % (4 * esz) == 0);
  const int64_t total = (int64_t)N * H * W * (vec ? C / 4 : C);
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == VCM_BF16 && C % 8 == 0 && C <= 256 && 256 % (C / 8) == 0 && N <= 65535 && feat_cstride % 8 == 0 && out_cstride % 8 == 0 && flow_cstride % 2 == 0 &&
      (uintptr_t)feat % 16 == 0 && (uintptr_t)out % 16 == 0 && (uintptr_t)flow % 8 == 0) {
    const int th = 256 / (C / 8) / 8;
    flow_warp8_bf16_kernel<<<dim3((W + 7) / 8, (H + 2 * th - 1) / (2 * th), N), 256, 0, s>>>(
        (const __nv_bfloat16*)feat, flow, (__nv_bfloat16*)out, N, H, W, C, feat_cstride, flow_cstride, out_cstride);
    VCM_CHECK_LAUNCH();
    return 0;
  }
  VCM_DISPATCH_DTYPE(dtype, {
    if (vec)
      flow_warp_kernel<T, true><<<ew_grid(total), 256, 0, s>>>((const T*)feat, flow, (T*)out, N, H, W, C,
                                                             feat_cstride, flow_cstride, out_cstride);
    else
      flow_warp_kernel<T, false><<<ew_grid(total), 256, 0, s>>>((const T*)feat, flow, (T*)out, N, H, W, C,
                                                              feat_cstride, flow_cstride, out_cstride);
  });
  VCM_CHECK_LAUNCH();
  return 0;
}

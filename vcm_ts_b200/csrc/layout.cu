// vcm_ts_b200/csrc/layout.cu -- HBM-bound glue kernels: NCHW<->NHWC, channel-slice copy,
// per-channel scale, 2x2 max-pool, SELayer (deterministic mean + FC + sigmoid).
// Reference: video_net.py:149-162 (SELayer), :185,207,210 (max_pool2d),
// video_model.py:280,297,306,326 (per-channel q scaling), video_coder.py:99-117 (NCHW frames).
#include "common.cuh"

namespace {

// NCHW fp32 -> NHWC: a 32(pixels) x 32(channels) tile is transposed through shared memory so
// both the read (along W) and the write (along C) are coalesced.
template <typename T>
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ src, T* __restrict__ dst, int C,
                                    int64_t HW, int dst_cs) {
  __shared__ float tile[32][33];
  const int n = blockIdx.z;
  const int64_t p0 = (int64_t)blockIdx.x * 32;
  const int c0 = blockIdx.y * 32;
  const int tx = threadIdx.x, ty = threadIdx.y;  // 32 x 8
  for (int j = ty; j < 32; j += 8) {
    const int c = c0 + j;
    const int64_t p = p0 + tx;
    tile[j][tx] = (c < C && p < HW) ? __ldg(src + ((int64_t)n * C + c) * HW + p) : 0.f;
  }
  __syncthreads();
  for (int j = ty; j < 32; j += 8) {
    const int64_t p = p0 + j;
    const int c = c0 + tx;
    if (c < C && p < HW) st_from_float(dst + ((int64_t)n * HW + p) * dst_cs + c, tile[tx][j]);
  }
}

template <typename T>
__global__ void nhwc_to_nchw_kernel(const T* __restrict__ src, float* __restrict__ dst, int C,
                                    int64_t HW, int src_cs, int do_clamp, float lo, float hi) {
  __shared__ float tile[32][33];
  const int n = blockIdx.z;
  const int64_t p0 = (int64_t)blockIdx.x * 32;
  const int c0 = blockIdx.y * 32;
  const int tx = threadIdx.x, ty = threadIdx.y;
  for (int j = ty; j < 32; j += 8) {
    const int64_t p = p0 + j;
    const int c = c0 + tx;
    tile[j][tx] = (c < C && p < HW) ? ld_as_float(src + ((int64_t)n * HW + p) * src_cs + c) : 0.f;
  }
  __syncthreads();
  for (int j = ty; j < 32; j += 8) {
    const int c = c0 + j;
    const int64_t p = p0 + tx;
    if (c < C && p < HW) {
      float v = tile[tx][j];
      if (do_clamp) v = fminf(fmaxf(v, lo), hi);
      dst[((int64_t)n * C + c) * HW + p] = v;
    }
  }
}

template <typename T, typename TO>
__global__ void copy_channels_kernel(const T* __restrict__ src, TO* __restrict__ dst, int64_t npix,
                                     int C, int src_cs, int dst_cs) {
  const int64_t total = npix * C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = i / C;
    const int c = (int)(i - p * C);
    st_from_float(dst + p * dst_cs + c, ld_as_float(src + p * src_cs + c));
  }
}

template <typename T>
__global__ void scale_channels_kernel(const T* __restrict__ src, T* __restrict__ dst,
                                      const float* __restrict__ vec, int64_t hw, int C, int src_cs,
                                      int dst_cs, int op, int64_t total) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = i / C;
    const int c = (int)(i - p * C);
    const int n = (int)(p / hw);
    const float v = ld_as_float(src + p * src_cs + c);
    const float s = __ldg(vec + (int64_t)n * C + c);
    st_from_float(dst + p * dst_cs + c, op == 0 ? __fdiv_rn(v, s) : __fmul_rn(v, s));
  }
}

template <typename T>
__global__ void maxpool2_kernel(const T* __restrict__ src, T* __restrict__ dst, int N, int H, int W,
                                int C, int src_cs, int dst_cs) {
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
    const T* s = src + (((int64_t)n * H + 2 * oy) * W + 2 * ox) * src_cs + c;
    const float a = ld_as_float(s), b = ld_as_float(s + src_cs);
    const float e = ld_as_float(s + (int64_t)W * src_cs), f = ld_as_float(s + (int64_t)(W + 1) * src_cs);
    st_from_float(dst + (((int64_t)n * Ho + oy) * Wo + ox) * dst_cs + c, fmaxf(fmaxf(a, b), fmaxf(e, f)));
  }
}

// bf16, 8 channels (16 bytes) per thread
__global__ void maxpool2_bf16x8_kernel(const __nv_bfloat16* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                       int N, int H, int W, int C, int src_cs, int dst_cs) {
  const int Ho = H / 2, Wo = W / 2, G = C / 8;
  const int64_t total = (int64_t)N * Ho * Wo * G;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int g = (int)(i % G);
    int64_t p = i / G;
    const int ox = (int)(p % Wo);
    p /= Wo;
    const int oy = (int)(p % Ho);
    const int n = (int)(p / Ho);
    const __nv_bfloat16* s = src + (((int64_t)n * H + 2 * oy) * W + 2 * ox) * src_cs + g * 8;
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(s));
    const uint4 b = __ldg(reinterpret_cast<const uint4*>(s + src_cs));
    const uint4 e = __ldg(reinterpret_cast<const uint4*>(s + (int64_t)W * src_cs));
    const uint4 f = __ldg(reinterpret_cast<const uint4*>(s + (int64_t)(W + 1) * src_cs));
    const __nv_bfloat162* ha = reinterpret_cast<const __nv_bfloat162*>(&a);
    const __nv_bfloat162* hb = reinterpret_cast<const __nv_bfloat162*>(&b);
    const __nv_bfloat162* he = reinterpret_cast<const __nv_bfloat162*>(&e);
    const __nv_bfloat162* hf = reinterpret_cast<const __nv_bfloat162*>(&f);
    uint4 r;
    __nv_bfloat162* hr = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
    for (int k = 0; k < 4; ++k) hr[k] = __hmax2(__hmax2(ha[k], hb[k]), __hmax2(he[k], hf[k]));
    *reinterpret_cast<uint4*>(dst + (((int64_t)n * Ho + oy) * Wo + ox) * dst_cs + g * 8) = r;
  }
}

// ---- SELayer -------------------------------------------------------------------------------
// stage 1: block b of sample n sums a fixed pixel range for every channel (fixed order ->
// deterministic); stage 2: one block per sample adds the partials in index order, divides by
// H*W, and applies the two bias-free Linear layers.
constexpr int SE_PIX_PER_BLOCK = 4096;
constexpr int SE_PIX_PER_BLOCK_VEC = 1024;   // vectorised (bf16) path: 4x the blocks, fills the GPU at 1/4 resolution

template <typename T>
__global__ void se_partial_kernel(const T* __restrict__ x, float* __restrict__ partial, int64_t hw,
                                  int C, int cs, int nblk) {
  extern __shared__ float sm[];  // [rows][C]
  const int n = blockIdx.y, b = blockIdx.x;
  const int rows = blockDim.x / C;  // pixel lanes per step (blockDim multiple of C, C <= 256)
  const int c = threadIdx.x % C, r = threadIdx.x / C;
  const int64_t p0 = (int64_t)b * SE_PIX_PER_BLOCK;
  const int64_t p1 = (p0 + SE_PIX_PER_BLOCK < hw) ? p0 + SE_PIX_PER_BLOCK : hw;
  float s = 0.f;
  if (r < rows)
    for (int64_t p = p0 + r; p < p1; p += rows) s += ld_as_float(x + ((int64_t)n * hw + p) * cs + c);
  if (r < rows) sm[r * C + c] = s;
  __syncthreads();
  if (r == 0) {
    float tot = 0.f;
    for (int k = 0; k < rows; ++k) tot += sm[k * C + c];
    partial[((int64_t)n * nblk + b) * C + c] = tot;
  }
}

// Vectorised variant (C % 8 == 0, 16-byte aligned rows): a thread owns 8 channels of a pixel lane and
// issues 16-byte loads, four in flight; same fixed summation structure on every run.
template <typename T>
__global__ void __launch_bounds__(256) se_partial_vec_kernel(const T* __restrict__ x, float* __restrict__ partial,
                                                             int64_t hw, int C, int cs, int nblk) {
  extern __shared__ float sm[];  // [rows][C]
  const int n = blockIdx.y, b = blockIdx.x;
  const int c8 = C >> 3, rows = 256 / c8;
  const int u = threadIdx.x % c8, r = threadIdx.x / c8;
  const int64_t p0 = (int64_t)b * SE_PIX_PER_BLOCK_VEC;
  const int64_t p1 = (p0 + SE_PIX_PER_BLOCK_VEC < hw) ? p0 + SE_PIX_PER_BLOCK_VEC : hw;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (r < rows) {
    const T* base = x + (int64_t)n * hw * cs + 8 * u;
    int64_t p = p0 + r;
    for (; p + 3 * (int64_t)rows < p1; p += 4 * (int64_t)rows) {
      float4 v[4][2];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const T* q = base + (p + (int64_t)j * rows) * cs;
        v[j][0] = ld4<T>(q);
        v[j][1] = ld4<T>(q + 4);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        acc[0] += v[j][0].x; acc[1] += v[j][0].y; acc[2] += v[j][0].z; acc[3] += v[j][0].w;
        acc[4] += v[j][1].x; acc[5] += v[j][1].y; acc[6] += v[j][1].z; acc[7] += v[j][1].w;
      }
    }
    for (; p < p1; p += rows) {
      const T* q = base + p * cs;
      const float4 a = ld4<T>(q), c = ld4<T>(q + 4);
      acc[0] += a.x; acc[1] += a.y; acc[2] += a.z; acc[3] += a.w;
      acc[4] += c.x; acc[5] += c.y; acc[6] += c.z; acc[7] += c.w;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) sm[r * C + 8 * u + i] = acc[i];
  }
  __syncthreads();
  if (threadIdx.x < C) {
    float tot = 0.f;
    for (int k = 0; k < rows; ++k) tot += sm[k * C + threadIdx.x];
    partial[((int64_t)n * nblk + b) * C + threadIdx.x] = tot;
  }
}

__global__ void se_fc_kernel(const float* __restrict__ partial, const float* __restrict__ w1,
                             const float* __restrict__ w2, float* __restrict__ scale, int64_t hw,
                             int C, int Cmid, int nblk) {
  extern __shared__ float sm[];  // mean[C] + hid[Cmid]
  float* mean = sm;
  float* hid = sm + C;
  const int n = blockIdx.x;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s = 0.f;
    for (int b = 0; b < nblk; ++b) s += partial[((int64_t)n * nblk + b) * C + c];
    mean[c] = s / (float)hw;
  }
  __syncthreads();
  for (int j = threadIdx.x; j < Cmid; j += blockDim.x) {
    float s = 0.f;
    for (int c = 0; c < C; ++c) s = fmaf(mean[c], w1[j * C + c], s);
    hid[j] = fmaxf(s, 0.f);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s = 0.f;
    for (int j = 0; j < Cmid; ++j) s = fmaf(hid[j], w2[c * Cmid + j], s);
    scale[(int64_t)n * C + c] = 1.f / (1.f + expf(-s));
  }
}

// Same, with the sum over the nblk partials striped over 1024/(C/4) thread groups, each thread
// adding float4 quarter-rows (fixed order: stripe sums, then groups in index order).  The serial
// loop above is a chain of ~2000 dependent L2 round trips at 1080p (20-100 us); this one is 32 per
// thread, eight in flight.
constexpr int SE_FC_THREADS = 1024;
__global__ void __launch_bounds__(SE_FC_THREADS) se_fc_striped_kernel(const float* __restrict__ partial,
                                                                      const float* __restrict__ w1,
                                                                      const float* __restrict__ w2,
                                                                      float* __restrict__ scale, int64_t hw, int C,
                                                                      int Cmid, int nblk) {
  extern __shared__ float sm[];  // part[groups][C] + mean[C] + hid[Cmid]
  const int c4 = C >> 2, groups = SE_FC_THREADS / c4;
  float* part = sm;
  float* mean = sm + groups * C;
  float* hid = mean + C;
  const int n = blockIdx.x;
  const int u = threadIdx.x % c4, grp = threadIdx.x / c4;
  {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    const float4* p = reinterpret_cast<const float4*>(partial + (int64_t)n * nblk * C) + u;
#pragma unroll 8
    for (int b = grp; b < nblk; b += groups) {
      const float4 v = __ldg(p + (int64_t)b * c4);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    reinterpret_cast<float4*>(part + grp * C)[u] = s;
  }
  __syncthreads();
  if (threadIdx.x < C) {
    float s = 0.f;
    for (int k = 0; k < groups; ++k) s += part[k * C + threadIdx.x];
    mean[threadIdx.x] = s / (float)hw;
  }
  __syncthreads();
  for (int j = threadIdx.x; j < Cmid; j += blockDim.x) {
    float s = 0.f;
    for (int k = 0; k < C; ++k) s = fmaf(mean[k], w1[j * C + k], s);
    hid[j] = fmaxf(s, 0.f);
  }
  __syncthreads();
  for (int k = threadIdx.x; k < C; k += blockDim.x) {
    float s = 0.f;
    for (int j = 0; j < Cmid; ++j) s = fmaf(hid[j], w2[k * Cmid + j], s);
    scale[(int64_t)n * C + k] = 1.f / (1.f + expf(-s));
  }
}

// ---- frame I/O ---------------------------------------------------------
// uint8 HWC frame (what PNG decoding yields, png_reader.py:37-39) -> float NCHW `u8 / 255.f`, zero-padded
// right / bottom to (Hp, Wp) (test_video.py:118-125, video_coder.py:112); and the way back,
// save_torch_image (stream_helper.py:151-154): crop, clip(rint(v * 255), 0, 255).
__global__ void u8hwc_to_frame_kernel(const uint8_t* __restrict__ src, float* __restrict__ dst, int N, int H,
                                      int W, int Hp, int Wp) {
  const int64_t total = (int64_t)N * 3 * Hp * Wp;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % Wp);
    int64_t r = i / Wp;
    const int y = (int)(r % Hp);
    r /= Hp;
    const int c = (int)(r % 3);
    const int n = (int)(r / 3);
    float v = 0.f;
    if (y < H && x < W) v = __fdiv_rn((float)src[(((int64_t)n * H + y) * W + x) * 3 + c], 255.f);
    dst[i] = v;
  }
}

__global__ void frame_to_u8hwc_kernel(const float* __restrict__ src, uint8_t* __restrict__ dst, int N, int H,
                                      int W, int Hs, int Ws) {
  const int64_t total = (int64_t)N * H * W * 3;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % 3);
    int64_t r = i / 3;
    const int x = (int)(r % W);
    r /= W;
    const int y = (int)(r % H);
    const int n = (int)(r / H);
    const float v = rintf(__fmul_rn(src[(((int64_t)n * 3 + c) * Hs + y) * Ws + x], 255.f));
    dst[i] = (uint8_t)fminf(fmaxf(v, 0.f), 255.f);
  }
}

inline int ew_grid(int64_t total) {
  int64_t g = (total + 255) / 256;
  return (int)(g > 148 * 16 ? 148 * 16 : (g < 1 ? 1 : g));
}

}  // namespace

extern "C" int vcm_nchw_to_nhwc(const float* src, void* dst, int32_t N, int32_t C, int32_t H,
                                int32_t W, int32_t dst_cstride, int32_t dtype, void* stream) {
  VCM_CHECK_ARG(src && dst && N > 0 && C > 0 && H > 0 && W > 0 && dst_cstride >= C, "bad argument");
  const int64_t HW = (int64_t)H * W;
  dim3 grid(cdiv(HW, 32), cdiv(C, 32), N), block(32, 8);
  VCM_DISPATCH_DTYPE(dtype, (nchw_to_nhwc_kernel<T><<<grid, block, 0, (cudaStream_t)stream>>>(
                                src, (T*)dst, C, HW, dst_cstride)));
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_u8hwc_to_frame(const uint8_t* src, float* dst, int32_t N, int32_t H, int32_t W, int32_t Hp,
                                  int32_t Wp, void* stream) {
  VCM_CHECK_ARG(src && dst && N > 0 && H > 0 && W > 0 && Hp >= H && Wp >= W, "bad argument");
  u8hwc_to_frame_kernel<<<ew_grid((int64_t)N * 3 * Hp * Wp), 256, 0, (cudaStream_t)stream>>>(src, dst, N, H, W, Hp, Wp);
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_frame_to_u8hwc(const float* src, uint8_t* dst, int32_t N, int32_t H, int32_t W, int32_t Hs,
                                  int32_t Ws, void* stream) {
  VCM_CHECK_ARG(src && dst && N > 0 && H > 0 && W > 0 && Hs >= H && Ws >= W, "bad argument");
  frame_to_u8hwc_kernel<<<ew_grid((int64_t)N * H * W * 3), 256, 0, (cudaStream_t)stream>>>(src, dst, N, H, W, Hs, Ws);
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_nhwc_to_nchw(const void* src, float* dst, int32_t N, int32_t C, int32_t H,
                                int32_t W, int32_t src_cstride, int32_t dtype, int32_t do_clamp,
                                float lo, float hi, void* stream) {
  VCM_CHECK_ARG(src && dst && N > 0 && C > 0 && H > 0 && W > 0 && src_cstride >= C, "bad argument");
  const int64_t HW = (int64_t)H * W;
  dim3 grid(cdiv(HW, 32), cdiv(C, 32), N), block(32, 8);
  VCM_DISPATCH_DTYPE(dtype, (nhwc_to_nchw_kernel<T><<<grid, block, 0, (cudaStream_t)stream>>>(
                                (const T*)src, dst, C, HW, src_cstride, do_clamp, lo, hi)));
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_copy_channels(const void* src, void* dst, int64_t npix, int32_t C,
                                 int32_t src_cstride, int32_t dst_cstride, int32_t src_dtype,
                                 int32_t dst_dtype, void* stream) {
  VCM_CHECK_ARG(src && dst && npix > 0 && C > 0, "bad argument");
  const int g = ew_grid(npix * C);
  cudaStream_t s = (cudaStream_t)stream;
  using bf = __nv_bfloat16;
  if (src_dtype == VCM_F32 && dst_dtype == VCM_F32)
    copy_channels_kernel<float, float><<<g, 256, 0, s>>>((const float*)src, (float*)dst, npix, C, src_cstride, dst_cstride);
  else if (src_dtype == VCM_F32 && dst_dtype == VCM_BF16)
    copy_channels_kernel<float, bf><<<g, 256, 0, s>>>((const float*)src, (bf*)dst, npix, C, src_cstride, dst_cstride);
  else if (src_dtype == VCM_BF16 && dst_dtype == VCM_F32)
    copy_channels_kernel<bf, float><<<g, 256, 0, s>>>((const bf*)src, (float*)dst, npix, C, src_cstride, dst_cstride);
  else if (src_dtype == VCM_BF16 && dst_dtype == VCM_BF16)
    copy_channels_kernel<bf, bf><<<g, 256, 0, s>>>((const bf*)src, (bf*)dst, npix, C, src_cstride, dst_cstride);
  else {
    vcm_set_error("vcm_copy_channels: bad dtype");
    return 1;
  }
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_scale_channels(const void* src, void* dst, const float* vec, int32_t N, int64_t hw,
                                  int32_t C, int32_t src_cstride, int32_t dst_cstride, int32_t op,
                                  int32_t dtype, void* stream) {
  VCM_CHECK_ARG(src && dst && vec && N > 0 && hw > 0 && C > 0 && (op == 0 || op == 1), "bad argument");
  const int64_t total = (int64_t)N * hw * C;
  VCM_DISPATCH_DTYPE(dtype, (scale_channels_kernel<T><<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(
                                (const T*)src, (T*)dst, vec, hw, C, src_cstride, dst_cstride, op, total)));
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_maxpool2(const void* src, void* dst, int32_t N, int32_t H, int32_t W, int32_t C,
                            int32_t src_cstride, int32_t dst_cstride, int32_t dtype, void* stream) {
  VCM_CHECK_ARG(src && dst && N > 0 && H > 1 && W > 1 && C > 0, "bad argument");
  const int64_t total = (int64_t)N * (H / 2) * (W / 2) * C;
  if (dtype == VCM_BF16 && C % 8 == 0 && src_cstride % 8 == 0 && dst_cstride % 8 == 0 &&
      (uintptr_t)src % 16 == 0 && (uintptr_t)dst % 16 == 0) {
    maxpool2_bf16x8_kernel<<<ew_grid(total / 8), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)src, (__nv_bfloat16*)dst, N, H, W, C, src_cstride, dst_cstride);
    VCM_CHECK_LAUNCH();
    return 0;
  }
  VCM_DISPATCH_DTYPE(dtype, (maxpool2_kernel<T><<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(
                                (const T*)src, (T*)dst, N, H, W, C, src_cstride, dst_cstride)));
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int32_t vcm_se_partial_count(int64_t hw) { return cdiv(hw, SE_PIX_PER_BLOCK_VEC); }   // upper bound of both paths

extern "C" int vcm_se_scale(const void* x, float* partial, const float* w1, const float* w2,
                            float* scale, int32_t N, int64_t hw, int32_t C, int32_t Cmid,
                            int32_t cstride, int32_t dtype, void* stream) {
  VCM_CHECK_ARG(x && partial && w1 && w2 && scale && N > 0 && hw > 0, "bad argument");
  VCM_CHECK_ARG(C > 0 && C <= 256 && Cmid > 0 && Cmid <= 64, "SELayer channels out of range");
  const int nblk = cdiv(hw, SE_PIX_PER_BLOCK);
  const int rows = 512 / C >= 1 ? 512 / C : 1;
  const int threads = rows * C;
  cudaStream_t s = (cudaStream_t)stream;
  // exact (fp32) mode keeps the original summation order: its symbols are pinned against the oracle
  const bool vec = dtype == VCM_BF16 && C % 8 == 0 && cstride % 8 == 0 && ((uintptr_t)x & 15) == 0 && 256 % (C / 8) == 0;
  if (vec) {
    const int vrows = 256 / (C / 8);
    const int vblk = cdiv(hw, SE_PIX_PER_BLOCK_VEC);
    VCM_DISPATCH_DTYPE(dtype, (se_partial_vec_kernel<T><<<dim3(vblk, N), 256, vrows * C * sizeof(float), s>>>(
                                  (const T*)x, partial, hw, C, cstride, vblk)));
    VCM_CHECK_LAUNCH();
    if (C % 4 == 0 && SE_FC_THREADS % (C / 4) == 0 && ((uintptr_t)partial & 15) == 0) {
      se_fc_striped_kernel<<<N, SE_FC_THREADS, (4 * SE_FC_THREADS + C + Cmid) * sizeof(float), s>>>(
          partial, w1, w2, scale, hw, C, Cmid, vblk);
    } else {
      se_fc_kernel<<<N, 256, (C + Cmid) * sizeof(float), s>>>(partial, w1, w2, scale, hw, C, Cmid, vblk);
    }
    VCM_CHECK_LAUNCH();
    return 0;
  }
  VCM_DISPATCH_DTYPE(dtype, (se_partial_kernel<T><<<dim3(nblk, N), threads, threads * sizeof(float), s>>>(
                                (const T*)x, partial, hw, C, cstride, nblk)));
  VCM_CHECK_LAUNCH();
  se_fc_kernel<<<N, 256, (C + Cmid) * sizeof(float), s>>>(partial, w1, w2, scale, hw, C, Cmid, nblk);
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_se_scale_from_sums(const float* sums, int32_t rows, const float* w1, const float* w2,
                                      float* scale, int64_t hw, int32_t C, int32_t Cmid, void* stream) {
  VCM_CHECK_ARG(sums && w1 && w2 && scale && rows > 0 && hw > 0 && C > 0 && Cmid > 0, "bad argument");
  cudaStream_t s = (cudaStream_t)stream;
  if (C % 4 == 0 && SE_FC_THREADS % (C / 4) == 0 && ((uintptr_t)sums & 15) == 0) {
    se_fc_striped_kernel<<<1, SE_FC_THREADS, (4 * SE_FC_THREADS + C + Cmid) * sizeof(float), s>>>(
        sums, w1, w2, scale, hw, C, Cmid, rows);
  } else {
    se_fc_kernel<<<1, 256, (C + Cmid) * sizeof(float), s>>>(sums, w1, w2, scale, hw, C, Cmid, rows);
  }
  VCM_CHECK_LAUNCH();
  return 0;
}

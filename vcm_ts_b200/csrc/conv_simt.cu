// vcm_ts_b200/csrc/conv_simt.cu -- exact-mode convolution engine: fp32 FMA implicit GEMM.
//
// Replaces every nn.Conv2d call of the DCVC-HEM path (video_model.py:17-224, video_net.py:74-276,
// layers.py:18-127, image_model.py:16-48) when bit-level agreement of the quantised symbols with
// the fp32 reference matters more than speed, and serves the shapes the tcgen05 engine does not
// take (stride 2, Cin in {2,3}, Cout in {2,3,8}).
//
// Tiling: CTA = 8x8 output pixels x BN output channels, 256 threads, 4 x (BN/16) register tile,
// K loop over (tap, 16 input channels) with register-staged double buffering in shared memory.
// NHWC activations make the A-operand loads 16-byte vectors along channels; the packed weight
// layout [tap][Cin][Cout] makes the B-operand loads 16-byte vectors along Cout.
#include "conv_common.cuh"

namespace {

constexpr int BM = 64;       // 8 x 8 output pixels
constexpr int BK = 16;
constexpr int A_PAD = 4;

template <typename T, typename TO, int BN>
__global__ void __launch_bounds__(256) conv_simt_kernel(const ConvParams P) {
  constexpr int TN = BN / 16;
  __shared__ __align__(16) float As[2][BK][BM + A_PAD];
  __shared__ __align__(16) float Bs[2][BK][BN];

  const int t = threadIdx.x;
  const int tile = blockIdx.x;
  const int ty0 = (tile / P.tiles_x) * 8;
  const int tx0 = (tile % P.tiles_x) * 8;
  const int n0 = blockIdx.y * BN;
  const int n = blockIdx.z;

  // A-load role
  const int ap = t >> 2, aq = t & 3;
  const int a_oy = ty0 + (ap >> 3), a_ox = tx0 + (ap & 7);
  const bool a_pix_ok = (a_oy < P.Ho) && (a_ox < P.Wo);
  const T* xin = reinterpret_cast<const T*>(P.x) + (size_t)n * P.H * P.W * P.in_cs;
  // B-load role
  const int br = t >> 4;
  const int bc = (BN == 64) ? (t & 15) * 4 : (t & 15);
  // compute role
  const int cy = t >> 4, cx = t & 15;

  float acc[4][TN];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  const int nchunks = P.K * P.K * P.cin_chunks;
  float4 ra;
  float rb[TN == 4 ? 4 : 1];

  auto load_global = [&](int it) {
    const int tap = it / P.cin_chunks;
    const int k0 = (it - tap * P.cin_chunks) * BK;
    const int ky = tap / P.K, kx = tap - ky * P.K;
    // ---- A
    ra = make_float4(0.f, 0.f, 0.f, 0.f);
    const int iy = a_oy * P.stride + ky - P.pad;
    const int ix = a_ox * P.stride + kx - P.pad;
    const int c = k0 + aq * 4;
    if (a_pix_ok && iy >= 0 && iy < P.H && ix >= 0 && ix < P.W && c < P.Cin) {
      const T* p = xin + ((size_t)iy * P.W + ix) * P.in_cs + c;
      if (P.vecA) {
        ra = ld4<T>(p);
      } else {
        ra.x = ld_as_float(p);
        if (c + 1 < P.Cin) ra.y = ld_as_float(p + 1);
        if (c + 2 < P.Cin) ra.z = ld_as_float(p + 2);
        if (c + 3 < P.Cin) ra.w = ld_as_float(p + 3);
      }
      if (P.flags & VCM_CONV_PRE_ACT) {
        ra.x = leaky(ra.x, P.pre_slope); ra.y = leaky(ra.y, P.pre_slope);
        ra.z = leaky(ra.z, P.pre_slope); ra.w = leaky(ra.w, P.pre_slope);
      }
    }
    // ---- B
    const int cin = k0 + br;
#pragma unroll
    for (int j = 0; j < (TN == 4 ? 4 : 1); ++j) rb[j] = 0.f;
    if (cin < P.Cin) {
      const float* wp = P.w + ((size_t)tap * P.Cin + cin) * P.Cout + n0 + bc;
      if (BN == 64) {
        if (P.vecB && n0 + bc + 3 < P.Cout) {
          const float4 v = __ldg(reinterpret_cast<const float4*>(wp));
          rb[0] = v.x; rb[1] = v.y; rb[2] = v.z; rb[3] = v.w;
        } else {
#pragma unroll
          for (int j = 0; j < (TN == 4 ? 4 : 1); ++j)
            if (n0 + bc + j < P.Cout) rb[j] = __ldg(wp + j);
        }
      } else {
        if (n0 + bc < P.Cout) rb[0] = __ldg(wp);
      }
    }
  };
  auto store_smem = [&](int buf) {
    As[buf][aq * 4 + 0][ap] = ra.x;
    As[buf][aq * 4 + 1][ap] = ra.y;
    As[buf][aq * 4 + 2][ap] = ra.z;
    As[buf][aq * 4 + 3][ap] = ra.w;
    if (BN == 64) {
      *reinterpret_cast<float4*>(&Bs[buf][br][bc]) = make_float4(rb[0], rb[TN == 4 ? 1 : 0],
                                                                 rb[TN == 4 ? 2 : 0], rb[TN == 4 ? 3 : 0]);
    } else {
      Bs[buf][br][bc] = rb[0];
    }
  };

  load_global(0);
  store_smem(0);
  __syncthreads();
  int buf = 0;
  for (int it = 0; it < nchunks; ++it) {
    const bool more = (it + 1 < nchunks);
    if (more) load_global(it + 1);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[buf][k][cy * 4]);
      float b[TN];
      if (TN == 4) {
        const float4 bv = *reinterpret_cast<const float4*>(&Bs[buf][k][cx * 4]);
        b[0] = bv.x; b[TN == 4 ? 1 : 0] = bv.y; b[TN == 4 ? 2 : 0] = bv.z; b[TN == 4 ? 3 : 0] = bv.w;
      } else {
        b[0] = Bs[buf][k][cx];
      }
#pragma unroll
      for (int j = 0; j < TN; ++j) {
        acc[0][j] = fmaf(a.x, b[j], acc[0][j]);
        acc[1][j] = fmaf(a.y, b[j], acc[1][j]);
        acc[2][j] = fmaf(a.z, b[j], acc[2][j]);
        acc[3][j] = fmaf(a.w, b[j], acc[3][j]);
      }
    }
    if (more) store_smem(buf ^ 1);
    __syncthreads();
    buf ^= 1;
  }

#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p = cy * 4 + i;
    const int oy = ty0 + (p >> 3), ox = tx0 + (p & 7);
    if (oy >= P.Ho || ox >= P.Wo) continue;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int co = n0 + cx * TN + j;
      if (co < P.Cout) conv_epilogue_store<TO>(P, n, oy, ox, co, acc[i][j], __ldg(P.bias + co));
    }
  }
}

// OIHW -> [tap][Cin_p][Cout] with optional input-channel gather
__global__ void pack_weight_kernel(const float* __restrict__ w, float* __restrict__ out, int Cout,
                                   int Cin, int K, int Cin_p, const int* __restrict__ in_map) {
  const int64_t total = (int64_t)K * K * Cin_p * Cout;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int co = (int)(i % Cout);
    const int ci_p = (int)((i / Cout) % Cin_p);
    const int tap = (int)(i / ((int64_t)Cout * Cin_p));
    const int ci = in_map ? in_map[ci_p] : ci_p;
    float v = 0.f;
    if (ci >= 0 && ci < Cin) v = w[((int64_t)co * Cin + ci) * K * K + tap];
    out[i] = v;
  }
}

}  // namespace

extern "C" int vcm_pack_conv_weight(const float* w_oihw, float* w_packed, int32_t Cout, int32_t Cin,
                                    int32_t K, int32_t Cin_p, const int32_t* in_map, void* stream) {
  VCM_CHECK_ARG(w_oihw && w_packed && Cout > 0 && Cin > 0 && K > 0 && Cin_p > 0, "bad argument");
  const int64_t total = (int64_t)K * K * Cin_p * Cout;
  pack_weight_kernel<<<cdiv(total, 256) > 4096 ? 4096 : cdiv(total, 256), 256, 0,
                       (cudaStream_t)stream>>>(w_oihw, w_packed, Cout, Cin, K, Cin_p, in_map);
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_conv2d_fwd(const vcm_conv_desc* d, const void* x, const float* w_packed,
                              const float* bias, const void* res, const float* res_scale,
                              const void* res2, void* y, void* y2, void* stream) {
  VCM_CHECK_ARG(d && x && w_packed && bias && y, "null argument");
  ConvParams P;
  memset(&P, 0, sizeof(P));
  VCM_CHECK_ARG(conv_fill_params(d, P) == 0, "unsupported conv descriptor");
  VCM_CHECK_ARG(!(P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE)) || res, "residual flag without tensor");
  VCM_CHECK_ARG(!(P.flags & VCM_CONV_RES_SCALE) || res_scale, "RES_SCALE without scale");
  VCM_CHECK_ARG(!(P.flags & VCM_CONV_RES2) || res2, "RES2 without tensor");
  VCM_CHECK_ARG(!(P.flags & VCM_CONV_OUT2) || y2, "OUT2 without tensor");
  VCM_CHECK_ARG(!(P.flags & VCM_CONV_SE_SUM), "VCM_CONV_SE_SUM is a tensor-core engine feature");
  P.y2 = y2;
  P.x = x; P.w = w_packed; P.bias = bias; P.res = res; P.res_scale = res_scale; P.res2 = res2; P.y = y;
  P.cin_chunks = cdiv(P.Cin, BK);
  VCM_CHECK_ARG((d->in_dtype == VCM_F32 || d->in_dtype == VCM_BF16) &&
                    (d->out_dtype == VCM_F32 || d->out_dtype == VCM_BF16), "bad dtype");
  const int esz = d->in_dtype == VCM_F32 ? 4 : 2;
  P.vecA = (P.Cin % 4 == 0) && (P.in_cs % 4 == 0) && (((uintptr_t)x) % (4 * esz) == 0);
  P.vecB = (P.Cout % 4 == 0) && (((uintptr_t)w_packed) % 16 == 0);
  P.tiles_x = cdiv(P.Wo, 8);
  P.tiles_y = cdiv(P.Ho, 8);
  const bool narrow = P.Cout <= 16;
  const int BN = narrow ? 16 : 64;
  dim3 grid(P.tiles_x * P.tiles_y, cdiv(P.Cout, BN), P.N);
  VCM_CHECK_ARG(grid.y <= 65535 && grid.z <= 65535, "grid too large");
  cudaStream_t s = (cudaStream_t)stream;
#define LAUNCH_SIMT(TI, TO_)                                          \
  do {                                                                \
    if (narrow)                                                       \
      conv_simt_kernel<TI, TO_, 16><<<grid, 256, 0, s>>>(P);          \
    else                                                              \
      conv_simt_kernel<TI, TO_, 64><<<grid, 256, 0, s>>>(P);          \
  } while (0)
  if (d->in_dtype == VCM_F32 && d->out_dtype == VCM_F32) LAUNCH_SIMT(float, float);
  else if (d->in_dtype == VCM_F32) LAUNCH_SIMT(float, __nv_bfloat16);
  else if (d->out_dtype == VCM_F32) LAUNCH_SIMT(__nv_bfloat16, float);
  else LAUNCH_SIMT(__nv_bfloat16, __nv_bfloat16);
#undef LAUNCH_SIMT
  VCM_CHECK_LAUNCH();
  return 0;
}

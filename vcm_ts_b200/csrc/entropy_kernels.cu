// vcm_ts_b200/csrc/entropy_kernels.cu -- fused entropy-parameter / quantise / scale-index /
// likelihood epilogues of the dual-spatial-prior (checkerboard) coder.
//
//   vcm_dual_prior_pass0/1   common_model.py:82-217 (get_mask, process_with_mask,
//                            forward_dual_prior, compress_dual_prior, decompress_dual_prior)
//                            + entropy_models.py:264-268 (GaussianEncoder.build_indexes)
//   vcm_round_latent         video_model.py:284,310 / image_model.py:63 (z_hat = round(z))
//   vcm_symbols_to_latent    entropy_models.py:189-195 / :276-281 (decoded ints -> float tensor)
//   vcm_latent_bits          common_model.py:51-69 (probs_to_bits, Laplace / Normal CDF)
//   vcm_factorized_bits      common_model.py:71-73, entropy_models.py:54-73,105-117
//   vcm_sq_err_sum           video_model.py:544-545
//
// The reference runs ~60 elementwise ATen launches per dual-prior call; here each checkerboard
// pass is ONE kernel that reads (y, q_step, scales, means) once and writes int16 symbols, uint8
// scale indexes (in the reference's NCHW coding order), the reconstruction and the spatial-prior
// input.  Integer outputs are bit-exact for identical float inputs: division is IEEE (__fdiv_rn),
// rounding is round-half-to-even (rintf == torch.round), masks multiply exactly as the reference.
#include "common.cuh"

namespace {

inline int ew_grid(int64_t total) {
  int64_t g = (total + 255) / 256;
  return (int)(g > 148 * 16 ? 148 * 16 : (g < 1 ? 1 : g));
}

__device__ __forceinline__ uint8_t scale_index(float sc, float lmin, float lstep) {
  const float s = fmaxf(sc, 1e-5f);
  // The index is floor((ln s - lmin) / step).  logf (<= 1 ulp of |ln s| <= 11.6, i.e. <= 2.8e-5 in the
  // quotient) decides it unless the quotient lands within 2e-4 of an integer; only then is the correctly rounded fp32 log
  // taken through double -- the best match for the reference's CPU logf -- which is ~60x slower on B200
  // (ncu r02: 53 us per pass with the double log on every element, DRAM 1.7 % of peak).
  float v = __fdiv_rn(__fsub_rn(logf(s), lmin), lstep);
  const float fr = v - floorf(v);
  if (fr < 2e-4f || fr > 1.f - 2e-4f) v = __fdiv_rn(__fsub_rn((float)log((double)s), lmin), lstep);
  v = fminf(fmaxf(v, 0.f), 255.f);
  return (uint8_t)(int)v;
}

// Symbols travel as int16 (the reference uses int32, entropy_models.py:183-187).  The quantised value
// itself is clamped to the int16 range BEFORE the reconstruction is formed, so the encoder's y_hat and
// the decoder's (which only sees the int16 symbol) agree for any input; |q| > 32767 never occurs for a
// codec operating on [0,1] frames and would cost > 4 kB of bypass digits per symbol in the reference.
__device__ __forceinline__ float clamp_sym(float q) { return fminf(fmaxf(q, -32768.f), 32767.f); }
__device__ __forceinline__ int16_t to_sym(float q) { return (int16_t)(int)q; }

// Both passes work on tiles of DP_TP consecutive pixels of one image: activations are NHWC (channels fastest) but
// the coder's symbol / index planes are NCHW (pixels fastest).  A first version indexed everything in plane
// order, so every 2-byte activation access of a warp hit 32 different cache lines (ncu r02: 53 us per launch at
// 1.7 % of the DRAM peak, long-scoreboard stall 49 cycles per issue).  Now the NHWC side is walked channel-fastest
// (coalesced), symbols and indexes cross over through shared memory and the planes are written pixel-fastest.
// Per element the float operations are the same statements as before (bit-exact stage).
constexpr int DP_TP = 32;          // pixels per tile
constexpr int DP_THREADS = 256;

template <typename T>
__global__ void __launch_bounds__(DP_THREADS)
dual_prior_pass0_kernel(const vcm_dp_desc D, const T* __restrict__ y, const T* __restrict__ fusion,
                        T* __restrict__ params, float* __restrict__ yq, int16_t* __restrict__ sym,
                        uint8_t* __restrict__ idx, float* __restrict__ y_q_full, float* __restrict__ scales_full) {
  extern __shared__ uint8_t dp_smem[];
  const int C = D.C, Ch = D.C / 2;
  const int64_t hw = (int64_t)D.h * D.w;
  int16_t* s_sym = reinterpret_cast<int16_t*>(dp_smem);                    // [Ch][DP_TP + 1]
  uint8_t* s_idx = dp_smem + (size_t)Ch * (DP_TP + 1) * sizeof(int16_t);   // [Ch][DP_TP + 1]
  const int tiles = (int)((hw + DP_TP - 1) / DP_TP);
  for (int t = blockIdx.x; t < tiles * D.N; t += gridDim.x) {
    const int n = t / tiles;
    const int64_t p0 = (int64_t)(t - n * tiles) * DP_TP;
    const int np = (int)(hw - p0 < DP_TP ? hw - p0 : DP_TP);
    const int64_t plane0 = (int64_t)n * Ch * hw + p0;          // + c * hw + p: the NCHW plane position
    if (D.mode == 2) {
      for (int e = threadIdx.x; e < Ch * DP_TP; e += DP_THREADS) {
        const int c = e / DP_TP, pp = e % DP_TP;
        if (pp < np) s_sym[c * (DP_TP + 1) + pp] = sym[plane0 + (int64_t)c * hw + pp];
      }
    }
    __syncthreads();
    for (int e = threadIdx.x; e < np * C; e += DP_THREADS) {
      const int pl = e / C, cc = e - pl * C;
      const int k = cc >= Ch ? 1 : 0, c = cc - k * Ch;
      const int64_t plin = p0 + pl;
      const int hy = (int)(plin / D.w), wx = (int)(plin - (int64_t)hy * D.w);
      const int64_t pix = (int64_t)n * hw + plin;
      const bool m0 = ((hy + wx) & 1) == 0;
      const float qs = fmaxf(ld_as_float(fusion + pix * D.fusion_cstride + cc), 0.5f);
      const float sc = ld_as_float(fusion + pix * D.fusion_cstride + C + cc);
      const float mu = ld_as_float(fusion + pix * D.fusion_cstride + 2 * C + cc);
      T* pp = params + pix * D.params_cstride;
      st_from_float(pp + C + cc, mu);
      st_from_float(pp + 2 * C + cc, sc);
      st_from_float(pp + 3 * C + cc, qs);
      const bool active = (k == 0) ? m0 : !m0;
      float yhat = 0.f, q = 0.f;
      if (D.mode == 0) {
        const float v = __fdiv_rn(ld_as_float(y + pix * D.y_cstride + cc), qs);
        yq[pix * C + cc] = v;
        if (active) {
          q = clamp_sym(rintf(__fsub_rn(v, mu)));
          yhat = __fadd_rn(q, mu);
          s_sym[c * (DP_TP + 1) + pl] = to_sym(q);
          s_idx[c * (DP_TP + 1) + pl] = scale_index(sc, D.log_scale_min, D.log_scale_step);
        }
      } else if (D.mode == 1) {
        if (active) s_idx[c * (DP_TP + 1) + pl] = scale_index(sc, D.log_scale_min, D.log_scale_step);
      } else {
        if (active) {
          q = (float)s_sym[c * (DP_TP + 1) + pl];
          yhat = __fadd_rn(q, mu);
        }
      }
      if (D.mode != 1) {
        st_from_float(pp + cc, yhat);
        if (y_q_full) {
          y_q_full[pix * C + cc] = q;
          scales_full[pix * C + cc] = active ? sc : 0.f;
        }
      }
    }
    __syncthreads();
    if (D.mode != 2) {
      // every (c, pixel) position is coded by exactly one of the two channel halves in this pass
      for (int e = threadIdx.x; e < Ch * DP_TP; e += DP_THREADS) {
        const int c = e / DP_TP, pp = e % DP_TP;
        if (pp < np) {
          if (D.mode == 0) sym[plane0 + (int64_t)c * hw + pp] = s_sym[c * (DP_TP + 1) + pp];
          idx[plane0 + (int64_t)c * hw + pp] = s_idx[c * (DP_TP + 1) + pp];
        }
      }
    }
    __syncthreads();
  }
}

template <typename T>
__global__ void __launch_bounds__(DP_THREADS)
dual_prior_pass1_kernel(const vcm_dp_desc D, const float* __restrict__ yq, const T* __restrict__ prior,
                        const T* __restrict__ params, const float* __restrict__ q_outer, T* __restrict__ y_hat,
                        int16_t* __restrict__ sym, uint8_t* __restrict__ idx, float* __restrict__ y_q_full,
                        float* __restrict__ scales_full) {
  extern __shared__ uint8_t dp_smem[];
  const int C = D.C, Ch = D.C / 2;
  const int64_t hw = (int64_t)D.h * D.w;
  int16_t* s_sym = reinterpret_cast<int16_t*>(dp_smem);
  uint8_t* s_idx = dp_smem + (size_t)Ch * (DP_TP + 1) * sizeof(int16_t);
  const int tiles = (int)((hw + DP_TP - 1) / DP_TP);
  for (int t = blockIdx.x; t < tiles * D.N; t += gridDim.x) {
    const int n = t / tiles;
    const int64_t p0 = (int64_t)(t - n * tiles) * DP_TP;
    const int np = (int)(hw - p0 < DP_TP ? hw - p0 : DP_TP);
    const int64_t plane0 = (int64_t)n * Ch * hw + p0;
    if (D.mode == 2) {
      for (int e = threadIdx.x; e < Ch * DP_TP; e += DP_THREADS) {
        const int c = e / DP_TP, pp = e % DP_TP;
        if (pp < np) s_sym[c * (DP_TP + 1) + pp] = sym[plane0 + (int64_t)c * hw + pp];
      }
    }
    __syncthreads();
    for (int e = threadIdx.x; e < np * C; e += DP_THREADS) {
      const int pl = e / C, cc = e - pl * C;
      const int k = cc >= Ch ? 1 : 0, c = cc - k * Ch;
      const int64_t plin = p0 + pl;
      const int hy = (int)(plin / D.w), wx = (int)(plin - (int64_t)hy * D.w);
      const int64_t pix = (int64_t)n * hw + plin;
      const bool m0 = ((hy + wx) & 1) == 0;
      const bool active = (k == 0) ? !m0 : m0;  // pass 1 codes the complementary positions
      const float sc = ld_as_float(prior + pix * D.prior_cstride + k * C + c);
      const float mu = ld_as_float(prior + pix * D.prior_cstride + k * C + Ch + c);
      const T* pp = params + pix * D.params_cstride;
      float part = 0.f, q = 0.f;
      if (D.mode == 0) {
        if (active) {
          q = clamp_sym(rintf(__fsub_rn(yq[pix * C + cc], mu)));
          part = __fadd_rn(q, mu);
          s_sym[c * (DP_TP + 1) + pl] = to_sym(q);
          s_idx[c * (DP_TP + 1) + pl] = scale_index(sc, D.log_scale_min, D.log_scale_step);
        }
      } else if (D.mode == 1) {
        if (active) s_idx[c * (DP_TP + 1) + pl] = scale_index(sc, D.log_scale_min, D.log_scale_step);
      } else {
        if (active) {
          q = (float)s_sym[c * (DP_TP + 1) + pl];
          part = __fadd_rn(q, mu);
        }
      }
      if (D.mode != 1) {
        const float first = ld_as_float(pp + cc);
        const float qs = ld_as_float(pp + 3 * C + cc);
        float v = __fadd_rn(first, part);
        v = __fmul_rn(v, qs);
        v = __fmul_rn(v, __ldg(q_outer + (int64_t)n * C + cc));
        st_from_float(y_hat + pix * D.yhat_cstride + cc, v);
        if (y_q_full && active) {
          y_q_full[pix * C + cc] = q;
          scales_full[pix * C + cc] = sc;
        }
      }
    }
    __syncthreads();
    if (D.mode != 2) {
      for (int e = threadIdx.x; e < Ch * DP_TP; e += DP_THREADS) {
        const int c = e / DP_TP, pp = e % DP_TP;
        if (pp < np) {
          if (D.mode == 0) sym[plane0 + (int64_t)c * hw + pp] = s_sym[c * (DP_TP + 1) + pp];
          idx[plane0 + (int64_t)c * hw + pp] = s_idx[c * (DP_TP + 1) + pp];
        }
      }
    }
    __syncthreads();
  }
}

template <typename T>
__global__ void round_latent_kernel(const T* __restrict__ z, T* __restrict__ z_hat,
                                    int16_t* __restrict__ sym, int N, int h, int w, int C, int z_cs,
                                    int zhat_cs) {
  const int64_t total = (int64_t)N * C * h * w;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int wx = (int)(i % w);
    int64_t r = i / w;
    const int hy = (int)(r % h);
    r /= h;
    const int c = (int)(r % C);
    const int n = (int)(r / C);
    const int64_t pix = ((int64_t)n * h + hy) * w + wx;
    const float q = clamp_sym(rintf(ld_as_float(z + pix * z_cs + c)));
    st_from_float(z_hat + pix * zhat_cs + c, q);
    if (sym) sym[i] = to_sym(q);
  }
}

template <typename T>
__global__ void symbols_to_latent_kernel(const int16_t* __restrict__ sym, T* __restrict__ z_hat,
                                         int N, int h, int w, int C, int zhat_cs) {
  const int64_t total = (int64_t)N * C * h * w;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int wx = (int)(i % w);
    int64_t r = i / w;
    const int hy = (int)(r % h);
    r /= h;
    const int c = (int)(r % C);
    const int n = (int)(r / C);
    const int64_t pix = ((int64_t)n * h + hy) * w + wx;
    st_from_float(z_hat + pix * zhat_cs + c, (float)sym[i]);
  }
}

// ---- reductions (deterministic: fixed chunking, fixed shuffle tree, ordered final sum) ------
constexpr int RED_CHUNK = 8192;

__device__ __forceinline__ float block_sum_256(float v) {
  __shared__ float ws[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
  if (threadIdx.x == 0)
    for (int k = 0; k < 8; ++k) t += ws[k];
  return t;  // valid in thread 0
}

__device__ __forceinline__ float bits_from_prob(float p) {
  const float b = -1.0f * logf(p + 1e-5f) / 0.69314718055994530942f;
  return fmaxf(b, 0.f);
}

__device__ __forceinline__ float laplace_cdf(float x, float scale) {
  const float sg = (x > 0.f) ? 1.f : ((x < 0.f) ? -1.f : 0.f);
  return 0.5f - 0.5f * sg * expm1f(-fabsf(x) / scale);
}
__device__ __forceinline__ float normal_cdf(float x, float scale) {
  return 0.5f * (1.f + erff(x * (1.f / scale) / 1.41421356237309504880f));
}

__global__ void __launch_bounds__(256)
latent_bits_kernel(const float* __restrict__ yq, const float* __restrict__ sc, float* __restrict__ partial,
                   int64_t per_sample, int nblk, int dist) {
  const int n = blockIdx.y, b = blockIdx.x;
  const int64_t i0 = (int64_t)b * RED_CHUNK;
  const int64_t i1 = i0 + RED_CHUNK < per_sample ? i0 + RED_CHUNK : per_sample;
  float acc = 0.f;
  for (int64_t i = i0 + threadIdx.x; i < i1; i += 256) {
    const float v = yq[(int64_t)n * per_sample + i];
    float s = sc[(int64_t)n * per_sample + i];
    float p;
    if (dist == 0) {
      s = fminf(fmaxf(s, 1e-5f), 1e10f);
      p = laplace_cdf(v + 0.5f, s) - laplace_cdf(v - 0.5f, s);
    } else {
      s = fminf(fmaxf(s, 0.11f), 1e10f);
      p = normal_cdf(v + 0.5f, s) - normal_cdf(v - 0.5f, s);
    }
    acc += bits_from_prob(p);
  }
  const float t = block_sum_256(acc);
  if (threadIdx.x == 0) partial[(int64_t)n * nblk + b] = t;
}

__device__ __forceinline__ float softplus_f(float x) { return x > 20.f ? x : log1pf(expf(x)); }

__device__ __forceinline__ float factorized_cdf(float x, const float* __restrict__ fp, int C, int c) {
  // fp rows: f1.h f1.b f1.a f2.h f2.b f2.a f3.h f3.b f3.a f4.h f4.b
#pragma unroll
  for (int l = 0; l < 3; ++l) {
    x = x * softplus_f(__ldg(fp + (3 * l) * C + c)) + __ldg(fp + (3 * l + 1) * C + c);
    x = x + tanhf(x) * tanhf(__ldg(fp + (3 * l + 2) * C + c));
  }
  x = x * softplus_f(__ldg(fp + 9 * C + c)) + __ldg(fp + 10 * C + c);
  return 1.f / (1.f + expf(-x));
}

template <typename T>
__global__ void __launch_bounds__(256)
factorized_bits_kernel(const T* __restrict__ z, const float* __restrict__ fp, float* __restrict__ partial,
                       int64_t hw, int C, int cs, int nblk) {
  const int n = blockIdx.y, b = blockIdx.x;
  const int64_t per_sample = hw * C;
  const int64_t i0 = (int64_t)b * RED_CHUNK;
  const int64_t i1 = i0 + RED_CHUNK < per_sample ? i0 + RED_CHUNK : per_sample;
  float acc = 0.f;
  for (int64_t i = i0 + threadIdx.x; i < i1; i += 256) {
    const int64_t p = i / C;
    const int c = (int)(i - p * C);
    const float v = ld_as_float(z + ((int64_t)n * hw + p) * cs + c);
    const float pr = factorized_cdf(v + 0.5f, fp, C, c) - factorized_cdf(v - 0.5f, fp, C, c);
    acc += bits_from_prob(pr);
  }
  const float t = block_sum_256(acc);
  if (threadIdx.x == 0) partial[(int64_t)n * nblk + b] = t;
}

__global__ void __launch_bounds__(256)
sq_err_kernel(const float* __restrict__ a, const float* __restrict__ bb, float* __restrict__ partial,
              int64_t per_sample, int nblk) {
  const int n = blockIdx.y, b = blockIdx.x;
  const int64_t i0 = (int64_t)b * RED_CHUNK;
  const int64_t i1 = i0 + RED_CHUNK < per_sample ? i0 + RED_CHUNK : per_sample;
  float acc = 0.f;
  for (int64_t i = i0 + threadIdx.x; i < i1; i += 256) {
    const float d = a[(int64_t)n * per_sample + i] - bb[(int64_t)n * per_sample + i];
    acc = fmaf(d, d, acc);
  }
  const float t = block_sum_256(acc);
  if (threadIdx.x == 0) partial[(int64_t)n * nblk + b] = t;
}

__global__ void final_sum_kernel(const float* __restrict__ partial, float* __restrict__ out, int nblk) {
  // one warp per sample; lane-strided double accumulation in a fixed order
  const int n = blockIdx.x;
  double acc = 0.0;
  for (int b = threadIdx.x; b < nblk; b += 32) acc += (double)partial[(int64_t)n * nblk + b];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (threadIdx.x == 0) out[n] = (float)acc;
}

int check_dp(const vcm_dp_desc* d) {
  if (!d) return 1;
  if (d->N <= 0 || d->h <= 0 || d->w <= 0 || d->C <= 0 || (d->C & 1)) return 1;
  if (d->mode < 0 || d->mode > 2) return 1;
  return 0;
}

}  // namespace

extern "C" int vcm_dual_prior_pass0(const vcm_dp_desc* d, const void* y, const void* fusion,
                                    void* params, float* yq, int16_t* sym, uint8_t* idx,
                                    float* y_q_full, float* scales_full, void* stream) {
  VCM_CHECK_ARG(check_dp(d) == 0, "bad descriptor");
  VCM_CHECK_ARG(fusion && params, "null tensor");
  VCM_CHECK_ARG(d->mode != 0 || (y && yq && sym && idx), "encode mode needs y, yq, sym, idx");
  VCM_CHECK_ARG(d->mode != 1 || idx, "index mode needs idx");
  VCM_CHECK_ARG(d->mode != 2 || sym, "reconstruct mode needs sym");
  VCM_CHECK_ARG((y_q_full == nullptr) == (scales_full == nullptr), "y_q_full/scales_full go together");
  const int64_t tiles = ((int64_t)d->h * d->w + DP_TP - 1) / DP_TP * d->N;
  const int grid = (int)(tiles < 148 * 8 ? tiles : 148 * 8);
  const size_t smem = (size_t)(d->C / 2) * (DP_TP + 1) * 3;
  VCM_CHECK_ARG(smem <= 48 * 1024, "C too large for the symbol tile");
  VCM_DISPATCH_DTYPE(d->dtype, (dual_prior_pass0_kernel<T><<<grid, DP_THREADS, smem, (cudaStream_t)stream>>>(
                                   *d, (const T*)y, (const T*)fusion, (T*)params, yq, sym, idx, y_q_full,
                                   scales_full)));
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_dual_prior_pass1(const vcm_dp_desc* d, const float* yq, const void* fusion,
                                    const void* prior, const void* params, const float* q_outer,
                                    void* y_hat, int16_t* sym, uint8_t* idx, float* y_q_full,
                                    float* scales_full, void* stream) {
  (void)fusion;
  VCM_CHECK_ARG(check_dp(d) == 0, "bad descriptor");
  VCM_CHECK_ARG(prior && params, "null tensor");
  VCM_CHECK_ARG(d->mode != 0 || (yq && sym && idx && y_hat && q_outer), "encode mode needs yq, sym, idx, y_hat");
  VCM_CHECK_ARG(d->mode != 1 || idx, "index mode needs idx");
  VCM_CHECK_ARG(d->mode != 2 || (sym && y_hat && q_outer), "reconstruct mode needs sym, y_hat");
  VCM_CHECK_ARG((y_q_full == nullptr) == (scales_full == nullptr), "y_q_full/scales_full go together");
  const int64_t tiles = ((int64_t)d->h * d->w + DP_TP - 1) / DP_TP * d->N;
  const int grid = (int)(tiles < 148 * 8 ? tiles : 148 * 8);
  const size_t smem = (size_t)(d->C / 2) * (DP_TP + 1) * 3;
  VCM_CHECK_ARG(smem <= 48 * 1024, "C too large for the symbol tile");
  VCM_DISPATCH_DTYPE(d->dtype, (dual_prior_pass1_kernel<T><<<grid, DP_THREADS, smem, (cudaStream_t)stream>>>(
                                   *d, yq, (const T*)prior, (const T*)params, q_outer, (T*)y_hat, sym, idx,
                                   y_q_full, scales_full)));
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_round_latent(const void* z, void* z_hat, int16_t* sym, int32_t N, int32_t h,
                                int32_t w, int32_t C, int32_t z_cstride, int32_t zhat_cstride,
                                int32_t dtype, void* stream) {
  VCM_CHECK_ARG(z && z_hat && N > 0 && h > 0 && w > 0 && C > 0, "bad argument");
  const int64_t total = (int64_t)N * C * h * w;
  VCM_DISPATCH_DTYPE(dtype, (round_latent_kernel<T><<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(
                                (const T*)z, (T*)z_hat, sym, N, h, w, C, z_cstride, zhat_cstride)));
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_symbols_to_latent(const int16_t* sym, void* z_hat, int32_t N, int32_t h, int32_t w,
                                     int32_t C, int32_t zhat_cstride, int32_t dtype, void* stream) {
  VCM_CHECK_ARG(sym && z_hat && N > 0 && h > 0 && w > 0 && C > 0, "bad argument");
  const int64_t total = (int64_t)N * C * h * w;
  VCM_DISPATCH_DTYPE(dtype, (symbols_to_latent_kernel<T><<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(
                                sym, (T*)z_hat, N, h, w, C, zhat_cstride)));
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int32_t vcm_bits_partial_count(int64_t per_sample) { return cdiv(per_sample, RED_CHUNK); }

extern "C" int vcm_latent_bits(const float* y_q, const float* scales, float* partial, float* out_bits,
                               int32_t N, int64_t per_sample, int32_t dist, void* stream) {
  VCM_CHECK_ARG(y_q && scales && partial && out_bits && N > 0 && per_sample > 0, "bad argument");
  VCM_CHECK_ARG(dist == 0 || dist == 1, "dist must be 0 (laplace) or 1 (normal)");
  const int nblk = cdiv(per_sample, RED_CHUNK);
  cudaStream_t s = (cudaStream_t)stream;
  latent_bits_kernel<<<dim3(nblk, N), 256, 0, s>>>(y_q, scales, partial, per_sample, nblk, dist);
  VCM_CHECK_LAUNCH();
  final_sum_kernel<<<N, 32, 0, s>>>(partial, out_bits, nblk);
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_factorized_bits(const void* z_hat, const float* fparams, float* partial,
                                   float* out_bits, int32_t N, int64_t hw, int32_t C, int32_t cstride,
                                   int32_t dtype, void* stream) {
  VCM_CHECK_ARG(z_hat && fparams && partial && out_bits && N > 0 && hw > 0 && C > 0, "bad argument");
  const int nblk = cdiv(hw * C, RED_CHUNK);
  cudaStream_t s = (cudaStream_t)stream;
  VCM_DISPATCH_DTYPE(dtype, (factorized_bits_kernel<T><<<dim3(nblk, N), 256, 0, s>>>(
                                (const T*)z_hat, fparams, partial, hw, C, cstride, nblk)));
  VCM_CHECK_LAUNCH();
  final_sum_kernel<<<N, 32, 0, s>>>(partial, out_bits, nblk);
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_sq_err_sum(const float* a, const float* b, float* partial, float* out, int32_t N,
                              int64_t per_sample, void* stream) {
  VCM_CHECK_ARG(a && b && partial && out && N > 0 && per_sample > 0, "bad argument");
  const int nblk = cdiv(per_sample, RED_CHUNK);
  cudaStream_t s = (cudaStream_t)stream;
  sq_err_kernel<<<dim3(nblk, N), 256, 0, s>>>(a, b, partial, per_sample, nblk);
  VCM_CHECK_LAUNCH();
  final_sum_kernel<<<N, 32, 0, s>>>(partial, out, nblk);
  VCM_CHECK_LAUNCH();
  return 0;
}

// ---- allocation-free variant of vcm_sq_err_sum ------------------------------------------
// No counterpart in the reference: added for callers that
// may not call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
// The partial sums of the two-stage reduction live in a file-scope device array instead of a
// caller-supplied buffer.  Every GPU that loads the library holds the 32 KiB until exit; all calls
// on one device share it, so calls that overlap on the GPU (different streams) race.  Nothing has to
// start at zero: every slot the second stage reads is written by the first stage of the same call.
constexpr int NOALLOC_MAX_PARTIALS = 8192;   // N * ceil(per_sample / RED_CHUNK) <= 8192, i.e. 64 Mi elements
static __device__ __align__(256) float g_noalloc_partial[NOALLOC_MAX_PARTIALS];

extern "C" int vcm_sq_err_sum_noalloc(const float* a, const float* b, float* out, int32_t N,
                                      int64_t per_sample, void* stream) {
  VCM_CHECK_ARG(a && b && out && N > 0 && per_sample > 0, "bad argument");
  const int64_t nblk64 = (per_sample + RED_CHUNK - 1) / RED_CHUNK;
  if (nblk64 * N > NOALLOC_MAX_PARTIALS) {
    vcm_set_error("%s: N * ceil(per_sample / %d) = %lld partial sums exceed the built-in scratch of %d "
                  "(use vcm_sq_err_sum with caller scratch)", __func__, RED_CHUNK, (long long)(nblk64 * N),
                  NOALLOC_MAX_PARTIALS);
    return 1;
  }
  // device address of the scratch array, looked up once per device with that device current (the
  // array's name is only a host-side handle here); idempotent, so racing first calls are harmless
  constexpr int kMaxDev = 64;
  static float* scratch_ptr[kMaxDev] = {nullptr};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDev) {
    vcm_set_error("%s: cudaGetDevice: %s (device %d)", __func__, cudaGetErrorString(e), dev);
    return 1;
  }
  float* partial = scratch_ptr[dev];
  if (!partial) {
    e = cudaGetSymbolAddress((void**)&partial, g_noalloc_partial);
    if (e != cudaSuccess || !partial) {
      vcm_set_error("%s: cudaGetSymbolAddress: %s", __func__, cudaGetErrorString(e));
      return 1;
    }
    scratch_ptr[dev] = partial;
  }
  const int nblk = (int)nblk64;
  cudaStream_t s = (cudaStream_t)stream;
  sq_err_kernel<<<dim3(nblk, N), 256, 0, s>>>(a, b, partial, per_sample, nblk);
  VCM_CHECK_LAUNCH();
  final_sum_kernel<<<N, 32, 0, s>>>(partial, out, nblk);
  VCM_CHECK_LAUNCH();
  return 0;
}

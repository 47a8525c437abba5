// vcm_ts_b200/csrc/conv_tc.cu -- tensor-core convolution engine for sm_100a:
// implicit GEMM on tcgen05.mma (bf16 x bf16 -> fp32 in TMEM), operands fed by TMA.
//
// Replaces the nn.Conv2d stacks of the DCVC-HEM path (video_model.py:17-224, video_net.py:74-276,
// layers.py:18-127, image_model.py:16-48) in bf16 mode.
//
// Mapping.  GEMM M = 128 output pixels (an 8 x 16 spatial tile), N = BN output channels (Cout padded
// to 16 and, above 256, split into equal N tiles), K = taps x Cin.  For every (tap, CHUNK input
// channels) the producer thread issues
//   * one 4-D TMA load of the NHWC activation box {CHUNK, 16*s, 8*s, 1} (element stride s) at the
//     tap-shifted coordinate: out-of-image taps are zero-filled by TMA, which IS the conv padding,
//     and a channel-slice view of a concat buffer is just a different base pointer / extent;
//   * one 2-D TMA load of the pre-packed weight slab [BN][CHUNK];
// both land 128/64/32-byte swizzled (K-major) in a multi-stage shared-memory ring guarded by
// full/empty mbarriers.  One elected thread issues CHUNK/16 tcgen05.mma per stage into one of two
// TMEM accumulator buffers and commits to the stage's empty barrier.  Two epilogue warpgroups
// (warps 4-7 / 8-11, one per accumulator buffer, setmaxnreg-enlarged register budget) drain TMEM with tcgen05.ld and run the fused epilogue (bias,
// LeakyReLU, residuals, SE scale, PixelShuffle / channel-slice store, activated twin); residual
// operands are fetched before the accumulator is awaited so their latency hides behind the MMAs.
// CTAs are persistent: grid = min(work items, #SM), work items round-robin, N tiles of one spatial
// tile adjacent so the activation boxes are re-read from L2.
#include <mutex>
#include <stdlib.h>

#include "conv_common.cuh"
#include "tc_epilogue.cuh"
#include "tc_host.cuh"
using namespace tc;

namespace {

constexpr int TH = 8, TW = 16;     // output tile: 8 rows x 16 cols = 128 GEMM rows
constexpr int NUM_THREADS = 384;   // warp 0 TMA, warp 1 MMA + TMEM alloc, warps 2-3 idle, warps 4-7 / 8-11 epilogue
constexpr int MAX_STAGES = 12;

struct TcGeom {
  // split-bf16 parity mode (VCM_CONV_SPLIT3): a stage holds the three bf16 planes of the activation box
  // (a_plane bytes apart) and the three weight parts (b_slab bytes apart); the six products x_i*w_j,
  // i + j <= 2, go to nsacc TMEM accumulators per item ([0] corrections, [1..] main term round-robin over
  // the stages), which the epilogue adds in fp32 (see conv_halo.cu mma_role_split)
  int split, pstride, a_plane, b_slab, nsacc;
  int nchunks, ntaps, kiters, stages;
  int a_bytes, b_bytes, b_off, stage_bytes;
  int tiles_x, tiles_y, n_tiles, num_items;   // items = spatial tiles x N tiles
  int bn, cout_p;                             // N per MMA, padded Cout (rows per weight slab set)
  int tmem_cols;
  int scalar_epi;                             // tc_epilogue_mode()
  int epi_base, bar_base, bias_base, epi_bw;
  uint32_t idesc;
};

// ---------------------------------------------------------------------------------- kernel
template <int CHUNK>
__global__ void __launch_bounds__(NUM_THREADS, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const ConvParams P, const TcGeom G) {
  constexpr uint32_t LAYOUT = CHUNK == 64 ? 2u : (CHUNK == 32 ? 4u : 6u);
  constexpr uint32_t SBO = 8u * CHUNK * 2u;  // 8 rows of CHUNK bf16
  constexpr int KSTEPS = CHUNK / 16;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + G.bar_base);
  uint64_t* full = bars;
  uint64_t* empty = bars + MAX_STAGES;
  uint64_t* tfull = bars + 2 * MAX_STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int sp_tiles = G.tiles_x * G.tiles_y;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int s = 0; s < G.stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], 4);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, (uint32_t)G.tmem_cols);
  {   // bias vector -> shared memory (the vector epilogue reads it per channel block)
    float* bs_ = reinterpret_cast<float*>(smem + G.bias_base);
    const int nb = G.cout_p < EPI_BIAS_MAX ? G.cout_p : EPI_BIAS_MAX;
    for (int i = threadIdx.x; i < nb; i += blockDim.x) bs_[i] = i < P.Cout ? __ldg(P.bias + i) : 0.f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t bias_smem = smem_u32(smem + G.bias_base);
  pdl_trigger();

  if (warp < 4) {
  reg_dec<56>();
  if (warp == 0) {
    // ================================ TMA producer (one thread) ================================
    if (lane == 0) {
      pdl_wait();   // activations are the previous kernel's output (this producer also loads the weights)
      int stage = 0;
      uint32_t phase = 0;
      for (int item = blockIdx.x; item < G.num_items; item += gridDim.x) {
        const int nt = item % G.n_tiles, sp = item / G.n_tiles;
        const int n = sp / sp_tiles, rem = sp - n * sp_tiles;
        const int y0 = (rem / G.tiles_x) * TH, x0 = (rem % G.tiles_x) * TW;
        for (int tap = 0; tap < G.ntaps; ++tap) {
          const int ky = tap / P.K, kx = tap - ky * P.K;
          const int cy = y0 * P.stride + ky - P.pad, cx = x0 * P.stride + kx - P.pad;
          for (int ch = 0; ch < G.nchunks; ++ch) {
            mbar_wait(&empty[stage], phase ^ 1u);
            uint8_t* sa = smem + (size_t)stage * G.stage_bytes;
            if (G.split) {
              mbar_expect_tx(&full[stage], 3u * (uint32_t)(G.a_bytes + G.b_bytes));
              for (int p = 0; p < 3; ++p) {
                tma_load_4d(smem_u32(sa + p * G.a_plane), &tmA, &full[stage], p * G.pstride + ch * CHUNK, cx, cy, n);
                tma_load_2d(smem_u32(sa + G.b_off + p * G.b_slab), &tmB, &full[stage], 0,
                            ((p * G.ntaps + tap) * G.nchunks + ch) * G.cout_p + nt * G.bn);
              }
            } else {
              mbar_expect_tx(&full[stage], (uint32_t)(G.a_bytes + G.b_bytes));
              tma_load_4d(smem_u32(sa), &tmA, &full[stage], ch * CHUNK, cx, cy, n);
              tma_load_2d(smem_u32(sa + G.b_off), &tmB, &full[stage], 0,
                          (tap * G.nchunks + ch) * G.cout_p + nt * G.bn);
            }
            if (++stage == G.stages) { stage = 0; phase ^= 1u; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issue (warp-uniform loop, one elected lane) ==========
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int item = blockIdx.x; item < G.num_items; item += gridDim.x) {
      mbar_wait(&tempty[acc], acc_phase ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * (G.split ? G.nsacc : 1) * G.bn);
      uint32_t started = 0;
      int mctr = 0;
      for (int it = 0; it < G.kiters; ++it) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(smem + (size_t)stage * G.stage_bytes);
          const uint64_t adesc = make_smem_desc(sa, SBO, LAYOUT);
          const uint64_t bdesc = make_smem_desc(sa + G.b_off, SBO, LAYOUT);
          if (G.split) {
            const uint32_t aplane16 = (uint32_t)G.a_plane >> 4, bslab16 = (uint32_t)G.b_slab >> 4;
            for (int pj = 2; pj >= 0; --pj)
              for (int pi = 0; pi <= 2 - pj; ++pi) {
                uint32_t sel = 0;
                if (pi == 0 && pj == 0) {
                  sel = 1u + (uint32_t)mctr;
                  if (++mctr == G.nsacc - 1) mctr = 0;
                }
                const uint32_t dd = d_tmem + sel * (uint32_t)G.bn;
                const uint64_t ad = adesc + (uint64_t)((uint32_t)pi * aplane16);
                const uint64_t bd = bdesc + (uint64_t)((uint32_t)pj * bslab16);
#pragma unroll
                for (int k = 0; k < KSTEPS; ++k)
                  umma_bf16(dd, ad + (uint64_t)(2 * k), bd + (uint64_t)(2 * k), G.idesc,
                            k > 0 ? 1u : ((started >> sel) & 1u));
                started |= 1u << sel;
              }
          } else {
#pragma unroll
            for (int k = 0; k < KSTEPS; ++k) {
              // advance 16 bf16 = 32 bytes along K inside the swizzle row: +2 in the >>4 address field
              umma_bf16(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), G.idesc,
                        (it > 0 || k > 0) ? 1u : 0u);
            }
          }
          umma_commit(&empty[stage]);
        }
        __syncwarp();
        if (++stage == G.stages) { stage = 0; phase ^= 1u; }
      }
      if (elect_one()) umma_commit(&tfull[acc]);
      __syncwarp();
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1u;
    }
  }
  } else {
    reg_inc<224>();
    // ============ epilogue: warpgroup g (warps 4-7 / 8-11) owns accumulator buffer g ===========
    const int g = (warp - 4) >> 2;
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    const uint32_t stage = smem_u32(smem + G.epi_base) + (uint32_t)((warp - 4) * EPI_STAGE_FLOATS * 4);
    pdl_wait();
    uint32_t acc_phase = 0;
    int local = 0;
    for (int item = blockIdx.x; item < G.num_items; item += gridDim.x, ++local) {
      if ((local & 1) != g) continue;
      const int nt = item % G.n_tiles, sp = item / G.n_tiles;
      const int n = sp / sp_tiles, rem = sp - n * sp_tiles;
      const int oy0 = (rem / G.tiles_x) * TH, ox0 = (rem % G.tiles_x) * TW;
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(g * (G.split ? G.nsacc : 1) * G.bn);
      const int cbase = nt * G.bn;
      if (G.split) {
        const int m = q * 32 + lane;
        epilogue_thread_scalar<float>(P, G.bn, n, oy0 + m / TW, ox0 + m % TW, cbase, taddr, &tfull[g], acc_phase,
                                      &tempty[g], G.nsacc);
      } else if (G.scalar_epi == 0) {
        if (G.epi_bw == 32)
          epilogue_warp_vec<32, TW>(P, G.bn, n, oy0, ox0, cbase, taddr, stage, q, lane, &tfull[g], acc_phase, &tempty[g], bias_smem);
        else
          epilogue_warp_vec<16, TW>(P, G.bn, n, oy0, ox0, cbase, taddr, stage, q, lane, &tfull[g], acc_phase, &tempty[g], bias_smem);
      } else {
        const int m = q * 32 + lane;
        const int oy = oy0 + m / TW, ox = ox0 + m % TW;
        if (G.scalar_epi == 2)
          epilogue_thread_scalar<float>(P, G.bn, n, oy, ox, cbase, taddr, &tfull[g], acc_phase, &tempty[g]);
        else
          epilogue_thread_scalar<bf>(P, G.bn, n, oy, ox, cbase, taddr, &tfull[g], acc_phase, &tempty[g]);
      }
      acc_phase ^= 1u;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)G.tmem_cols);
  }
}

// OIHW fp32 -> bf16 [tap][chunk][Cout_p][CHUNK] (K-major rows the weight TMA box is cut from);
// rows co >= Cout and channels mapped to -1 are zero.
__global__ void pack_umma_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out, int Cout,
                                 int Cout_p, int Cin, int K, int Cin_p, int chunk,
                                 const int* __restrict__ in_map) {
  const int nchunks = Cin_p / chunk;
  const int64_t total = (int64_t)K * K * Cin_p * Cout_p;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int kk = (int)(i % chunk);
    int64_t r = i / chunk;
    const int co = (int)(r % Cout_p);
    r /= Cout_p;
    const int ch = (int)(r % nchunks);
    const int tap = (int)(r / nchunks);
    const int ci_p = ch * chunk + kk;
    const int ci = in_map ? in_map[ci_p] : ci_p;
    float v = 0.f;
    if (co < Cout && ci >= 0 && ci < Cin) v = w[((int64_t)co * Cin + ci) * K * K + tap];
    out[i] = __float2bfloat16_rn(v);
  }
}

// the same for the three bf16 parts of the fp32 weights: out[part][tap][chunk][Cout_p][CHUNK],
// w0 = bf16(w), w1 = bf16(w - w0), w2 = bf16(w - w0 - w1) (each difference is exact in fp32)
__global__ void pack_umma_split3_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out, int Cout,
                                        int Cout_p, int Cin, int K, int Cin_p, int chunk,
                                        const int* __restrict__ in_map) {
  const int nchunks = Cin_p / chunk;
  const int64_t total = (int64_t)K * K * Cin_p * Cout_p;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int kk = (int)(i % chunk);
    int64_t r = i / chunk;
    const int co = (int)(r % Cout_p);
    r /= Cout_p;
    const int ch = (int)(r % nchunks);
    const int tap = (int)(r / nchunks);
    const int ci_p = ch * chunk + kk;
    const int ci = in_map ? in_map[ci_p] : ci_p;
    float v = 0.f;
    if (co < Cout && ci >= 0 && ci < Cin) v = w[((int64_t)co * Cin + ci) * K * K + tap];
    const __nv_bfloat16 w0 = __float2bfloat16_rn(v);
    const float r1 = __fsub_rn(v, __bfloat162float(w0));
    const __nv_bfloat16 w1 = __float2bfloat16_rn(r1);
    const __nv_bfloat16 w2 = __float2bfloat16_rn(__fsub_rn(r1, __bfloat162float(w1)));
    out[i] = w0;
    out[i + total] = w1;
    out[i + 2 * total] = w2;
  }
}

// fp32 view -> three bf16 planes (vcm_split_bf16x3); one thread = 8 channels of one pixel
__global__ void split_bf16x3_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t npix,
                                    int C, int Cp, int src_cs, int dst_cs, int dst_ps, int pre_act, float slope,
                                    int vec) {
  const int G8 = Cp / 8;
  const int64_t total = npix * G8;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = i / G8;
    const int c0 = (int)(i - p * G8) * 8;
    float v[8];
    const float* s = src + p * src_cs + c0;
    if (vec && c0 + 8 <= C) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(s)), b = __ldg(reinterpret_cast<const float4*>(s + 4));
      v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = c0 + j < C ? __ldg(s + j) : 0.f;
    }
    float x0[8], x1[8], x2[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float x = v[j];
      if (pre_act) x = leaky(x, slope);
      x0[j] = __bfloat162float(__float2bfloat16_rn(x));
      const float r1 = __fsub_rn(x, x0[j]);
      x1[j] = __bfloat162float(__float2bfloat16_rn(r1));
      x2[j] = __fsub_rn(r1, x1[j]);
    }
    __nv_bfloat16* o = dst + p * dst_cs + c0;
    *reinterpret_cast<uint4*>(o) = pack8(x0);
    *reinterpret_cast<uint4*>(o + dst_ps) = pack8(x1);
    *reinterpret_cast<uint4*>(o + 2 * dst_ps) = pack8(x2);
  }
}

// ---------------------------------------------------------------------------------- host
constexpr int SMEM_BAR_BYTES = 512;

}  // namespace

extern "C" int vcm_conv2d_tc_supported(const vcm_conv_desc* d) {
  if (!d) return 0;
  if (d->in_dtype != VCM_BF16) return 0;
  if (d->out_dtype != VCM_BF16 && d->out_dtype != VCM_F32) return 0;
  if (!(d->K == 1 || d->K == 3 || d->K == 7)) return 0;
  if (!(d->stride == 1 || d->stride == 2)) return 0;
  if (d->Cin % 16 || d->Cout < 1 || d->Cout > 2048) return 0;
  if (d->flags & VCM_CONV_PRE_ACT) return 0;
  // the epilogue computes leaky_relu as max(v, v*slope): exact for slopes in [0, 1]
  if ((d->flags & VCM_CONV_POST_ACT) && !(d->post_slope >= 0.f && d->post_slope <= 1.f)) return 0;
  if ((d->flags & VCM_CONV_OUT2) && !(d->out2_slope >= 0.f && d->out2_slope <= 1.f)) return 0;
  if (d->in_cstride % 8) return 0;
  if ((d->flags & VCM_CONV_SPLIT3) && (d->in_pstride % 8 || d->in_pstride < d->Cin ||
                                       d->in_cstride < 2 * d->in_pstride + d->Cin || d->out_dtype != VCM_F32))
    return 0;
  if ((d->flags & VCM_CONV_PIXEL_SHUFFLE) && d->Cout % 4) return 0;
  if (d->stride == 2 && ((d->H & 1) || (d->W & 1))) return 0;
  return 1;
}

extern "C" int64_t vcm_conv2d_tc_weight_bytes(int32_t Cout, int32_t Cin, int32_t K) {
  if (Cout <= 0 || Cin <= 0 || Cin % 16 || Cout > 2048) return 0;
  return (int64_t)K * K * Cin * round_up(Cout, 16) * 2;
}

extern "C" int vcm_pack_conv_weight_umma(const float* w_oihw, void* w_umma, int32_t Cout, int32_t Cin,
                                         int32_t K, int32_t Cin_p, const int32_t* in_map, void* stream) {
  VCM_CHECK_ARG(w_oihw && w_umma && Cout > 0 && Cin > 0 && K > 0 && Cin_p > 0 && Cin_p % 16 == 0,
                "bad argument");
  const int chunk = pick_chunk(Cin_p);
  const int cout_p = round_up(Cout, 16);
  const int64_t total = (int64_t)K * K * Cin_p * cout_p;
  const int grid = cdiv(total, 256) > 4096 ? 4096 : cdiv(total, 256);
  pack_umma_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(w_oihw, (__nv_bfloat16*)w_umma, Cout, cout_p, Cin,
                                                           K, Cin_p, chunk, in_map);
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_pack_conv_weight_umma_split3(const float* w_oihw, void* w_umma, int32_t Cout, int32_t Cin,
                                                int32_t K, int32_t Cin_p, const int32_t* in_map, void* stream) {
  VCM_CHECK_ARG(w_oihw && w_umma && Cout > 0 && Cin > 0 && K > 0 && Cin_p > 0 && Cin_p % 16 == 0,
                "bad argument");
  const int chunk = pick_chunk_split(Cin_p, K);
  const int cout_p = round_up(Cout, 16);
  const int64_t total = (int64_t)K * K * Cin_p * cout_p;
  const int grid = cdiv(total, 256) > 4096 ? 4096 : cdiv(total, 256);
  pack_umma_split3_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(w_oihw, (__nv_bfloat16*)w_umma, Cout, cout_p,
                                                                  Cin, K, Cin_p, chunk, in_map);
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_split_bf16x3(const float* src, void* dst, int64_t npix, int32_t C, int32_t Cp,
                                int32_t src_cstride, int32_t dst_cstride, int32_t dst_pstride,
                                int32_t pre_act, float pre_slope, void* stream) {
  VCM_CHECK_ARG(src && dst && npix > 0 && C > 0 && Cp >= C && Cp % 8 == 0, "bad argument");
  VCM_CHECK_ARG(dst_cstride % 8 == 0 && dst_pstride % 8 == 0 && ((uintptr_t)dst & 15) == 0,
                "planes must be 16-byte aligned");
  const int vec = (src_cstride % 4 == 0 && ((uintptr_t)src & 15) == 0) ? 1 : 0;
  const int64_t total = npix * (Cp / 8);
  const int grid = cdiv(total, 256) > 148 * 16 ? 148 * 16 : cdiv(total, 256);
  split_bf16x3_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst, npix, C, Cp, src_cstride,
                                                               dst_cstride, dst_pstride, pre_act, pre_slope, vec);
  VCM_CHECK_LAUNCH();
  return 0;
}

static bool force_se_pertap() {
  static const bool f = getenv("VCM_TC_PERTAP") && getenv("VCM_TC_PERTAP")[0] == '1';
  return f;
}

extern "C" int32_t vcm_conv2d_se_sum_rows(const vcm_conv_desc* d) {
  if (force_se_pertap()) return 0;
  if (!d || !vcm_conv2d_tc_supported(d) || d->in_dtype != VCM_BF16 || d->out_dtype != VCM_BF16) return 0;
  ConvParams P;
  memset(&P, 0, sizeof(P));
  if (conv_fill_params(d, P) != 0 || P.stride != 1 || (P.flags & VCM_CONV_SPLIT3)) return 0;
  // 16-byte aligned stand-ins: the epilogue choice looks at pointer alignment; a launch whose real pointers
  // make another choice fails instead of skipping the sums
  void* al = reinterpret_cast<void*>((uintptr_t)4096);
  P.x = al; P.bias = reinterpret_cast<const float*>(al); P.res = al;
  P.res_scale = reinterpret_cast<const float*>(al); P.res2 = al; P.y = al; P.y2 = al;
  P.flags |= VCM_CONV_SE_SUM;
  int rows = 0;
  const int rc = vcm_conv_halo_launch(d, P, al, al, nullptr, &rows);
  return rc == 0 ? rows : 0;
}

extern "C" int vcm_conv2d_tc_fwd(const vcm_conv_desc* d, const void* x, const void* w_umma,
                                 const float* bias, const void* res, const float* res_scale,
                                 const void* res2, void* y, void* y2, void* stream) {
  VCM_CHECK_ARG(d && x && w_umma && bias && y, "null argument");
  if (!vcm_conv2d_tc_supported(d)) {
    vcm_set_error("vcm_conv2d_tc_fwd: shape not supported by the tensor-core engine");
    return 2;
  }
  ConvParams P;
  memset(&P, 0, sizeof(P));
  VCM_CHECK_ARG(conv_fill_params(d, P) == 0, "unsupported conv descriptor");
  VCM_CHECK_ARG(!(P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE)) || res, "residual flag without tensor");
  VCM_CHECK_ARG(!(P.flags & VCM_CONV_RES_SCALE) || res_scale, "RES_SCALE without scale");
  VCM_CHECK_ARG(!(P.flags & VCM_CONV_RES2) || res2, "RES2 without tensor");
  VCM_CHECK_ARG(!(P.flags & VCM_CONV_OUT2) || y2, "OUT2 without tensor");
  if (((uintptr_t)x & 15) != 0) {   // TMA needs a 16-byte aligned base: the caller falls back to vcm_conv2d_fwd
    vcm_set_error("vcm_conv2d_tc_fwd: x is not 16-byte aligned (not supported by the tensor-core engine)");
    return 2;
  }
  VCM_CHECK_ARG(((uintptr_t)w_umma & 15) == 0, "weights must be 16-byte aligned");
  P.x = x; P.bias = bias; P.res = res; P.res_scale = res_scale; P.res2 = res2; P.y = y; P.y2 = y2;
  VCM_CHECK_ARG(!(P.flags & VCM_CONV_SE_SUM) || (y2 && !(P.flags & VCM_CONV_OUT2) && P.stride == 1 && !force_se_pertap()),
                "VCM_CONV_SE_SUM needs y2 = the sums array, no OUT2, a stride-1 conv on the halo / row engines");

  // stride-1 shapes go to the halo-tile engine (conv_halo.cu) unless VCM_TC_PERTAP=1 forces this one
  static const bool force_pertap = getenv("VCM_TC_PERTAP") && getenv("VCM_TC_PERTAP")[0] == '1';
  if (P.stride == 1 && !force_pertap) {
    const int rc = vcm_conv_halo_launch(d, P, x, w_umma, (cudaStream_t)stream);
    if (rc != 3) return rc;
  }
  VCM_CHECK_ARG(!(P.flags & VCM_CONV_SE_SUM), "VCM_CONV_SE_SUM: this shape runs on the per-tap engine, which has no fused sums");

  EncodeTiledFn enc = get_encode_fn();
  VCM_CHECK_ARG(enc != nullptr, "cuTensorMapEncodeTiled unavailable");

  TcGeom G;
  memset(&G, 0, sizeof(G));
  G.split = (P.flags & VCM_CONV_SPLIT3) ? 1 : 0;
  G.pstride = d->in_pstride;
  const int chunk = G.split ? pick_chunk_split(P.Cin, P.K) : pick_chunk(P.Cin);
  pick_n_tiling(P.Cout, G.cout_p, G.bn, G.n_tiles, G.split ? 128 : 256);
  G.nchunks = P.Cin / chunk;
  G.ntaps = P.K * P.K;
  G.kiters = G.ntaps * G.nchunks;
  G.a_bytes = 128 * chunk * 2;
  G.b_bytes = G.bn * chunk * 2;
  G.a_plane = round_up(G.a_bytes, 1024);
  G.b_slab = round_up(G.b_bytes, 1024);
  const int nrep = G.split ? 3 : 1;
  G.b_off = nrep * G.a_plane;
  G.stage_bytes = G.b_off + nrep * G.b_slab;
  const int bias_bytes = round_up((G.cout_p < EPI_BIAS_MAX ? G.cout_p : EPI_BIAS_MAX) * 4, 128);
  int stages = (SMEM_LIMIT - 1024 - SMEM_BAR_BYTES - bias_bytes - EPI_STAGE_BYTES) / G.stage_bytes;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  if (stages > G.kiters * 2) stages = G.kiters * 2 > 2 ? G.kiters * 2 : 2;
  VCM_CHECK_ARG(stages >= 2, "stage does not fit shared memory");
  G.stages = stages;
  G.tiles_x = cdiv(P.Wo, TW);
  G.tiles_y = cdiv(P.Ho, TH);
  G.num_items = G.tiles_x * G.tiles_y * P.N * G.n_tiles;
  G.nsacc = 1;
  if (G.split) {
    G.nsacc = 512 / (2 * G.bn);
    if (G.nsacc > 8) G.nsacc = 8;
    if (G.nsacc - 1 > G.kiters) G.nsacc = 1 + G.kiters;
  }
  int cols = 32;
  while (cols < 2 * G.nsacc * G.bn) cols <<= 1;
  G.tmem_cols = cols;
  G.idesc = make_idesc_bf16(128, G.bn);
  G.scalar_epi = tc_epilogue_mode(d, P);
  G.epi_bw = tc_epilogue_bw(d, G.bn);
  G.epi_base = G.stages * G.stage_bytes;
  G.bar_base = G.epi_base + EPI_STAGE_BYTES;
  G.bias_base = G.bar_base + SMEM_BAR_BYTES;

  CUtensorMap tmA, tmB;
  {
    cuuint64_t dims[4] = {(cuuint64_t)(P.Cin + (G.split ? 2 * G.pstride : 0)), (cuuint64_t)P.W, (cuuint64_t)P.H,
                          (cuuint64_t)P.N};
    cuuint64_t strides[3] = {(cuuint64_t)P.in_cs * 2, (cuuint64_t)P.W * P.in_cs * 2,
                             (cuuint64_t)P.H * P.W * P.in_cs * 2};
    cuuint32_t box[4] = {(cuuint32_t)chunk, (cuuint32_t)(TW * P.stride), (cuuint32_t)(TH * P.stride), 1};
    cuuint32_t estr[4] = {1, (cuuint32_t)P.stride, (cuuint32_t)P.stride, 1};
    CUresult r = enc(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_of(chunk), CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      vcm_set_error("vcm_conv2d_tc_fwd: cuTensorMapEncodeTiled(A) failed (%d)", (int)r);
      return 1;
    }
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)chunk, (cuuint64_t)G.kiters * nrep * G.cout_p};
    cuuint64_t strides[1] = {(cuuint64_t)chunk * 2};
    cuuint32_t box[2] = {(cuuint32_t)chunk, (cuuint32_t)G.bn};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(w_umma), dims, strides, box,
                     estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_of(chunk), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      vcm_set_error("vcm_conv2d_tc_fwd: cuTensorMapEncodeTiled(B) failed (%d)", (int)r);
      return 1;
    }
  }
  int smem_bytes = G.bias_base + bias_bytes + 1024;
  if (smem_bytes < 120 * 1024) smem_bytes = 120 * 1024;  // one CTA per SM (TMEM is allocated per CTA)
  const int grid = G.num_items < num_sms() ? G.num_items : num_sms();
  cudaStream_t s = (cudaStream_t)stream;
#define LAUNCH_TC(CH)                                                                                   \
  do {                                                                                                  \
    static bool attr_set[MAX_DEVICES] = {false};                                                        \
    cudaError_t e = ensure_max_smem(conv_tc_kernel<CH>, attr_set, SMEM_LIMIT);                          \
    if (e != cudaSuccess) {                                                                             \
      vcm_set_error("vcm_conv2d_tc_fwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));              \
      return 1;                                                                                         \
    }                                                                                                   \
    cudaLaunchConfig_t cfg = {};                                                                        \
    cfg.gridDim = dim3(grid);                                                                           \
    cfg.blockDim = dim3(NUM_THREADS);                                                                   \
    cfg.dynamicSmemBytes = smem_bytes;                                                                  \
    cfg.stream = s;                                                                                     \
    cudaLaunchAttribute attr[1];                                                                        \
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                                    \
    attr[0].val.programmaticStreamSerializationAllowed = 1;                                             \
    cfg.attrs = attr;                                                                                   \
    cfg.numAttrs = (P.flags & VCM_CONV_PDL) ? 1 : 0;                                                    \
    cudaLaunchKernelEx(&cfg, conv_tc_kernel<CH>, tmA, tmB, P, G);                                       \
  } while (0)
  if (chunk == 64) LAUNCH_TC(64);
  else if (chunk == 32) LAUNCH_TC(32);
  else LAUNCH_TC(16);
#undef LAUNCH_TC
  VCM_CHECK_LAUNCH();
  return 0;
}

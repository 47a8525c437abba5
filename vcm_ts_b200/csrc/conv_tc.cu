// vcm_ts_b200/csrc/conv_tc.cu -- tensor-core convolution engine for sm_100a:
// implicit GEMM on tcgen05.mma (bf16 x bf16 -> fp32 in TMEM), operands fed by TMA.
//
// Replaces the nn.Conv2d stacks of the DCVC-HEM path (video_model.py:17-224, video_net.py:74-276,
// layers.py:18-127, image_model.py:16-48) in bf16 mode.
//
// Mapping.  GEMM M = 128 output pixels (an 8 x 16 spatial tile), N = Cout (<= 256), K = taps x Cin.
// For every (tap, CHUNK input channels) the producer thread issues
//   * one 4-D TMA load of the NHWC activation box {CHUNK, 16*s, 8*s, 1} (element stride s) at the
//     tap-shifted coordinate: out-of-image taps are zero-filled by TMA, which IS the conv padding,
//     and a channel-slice view of a concat buffer is just a different base pointer / extent;
//   * one 2-D TMA load of the pre-packed weight slab [Cout][CHUNK];
// both land 128/64/32-byte swizzled (K-major) in a multi-stage shared-memory ring guarded by
// full/empty mbarriers.  One elected thread issues CHUNK/16 tcgen05.mma per stage into one of two
// TMEM accumulator buffers and commits to the stage's empty barrier; four epilogue warps drain
// the other accumulator with tcgen05.ld and run the fused epilogue (bias, LeakyReLU, residuals,
// SE scale, PixelShuffle / channel-slice store, activated twin) with 16-byte vector stores.
// CTAs are persistent: grid = min(tiles, #SM), tiles round-robin.
#include <cuda.h>

#include <mutex>
#include <unordered_map>

#include "conv_common.cuh"

namespace {

constexpr int TH = 8, TW = 16;   // output tile: 8 rows x 16 cols = 128 GEMM rows
constexpr int NUM_THREADS = 192; // warp 0 TMA, warp 1 MMA + TMEM alloc, warps 2..5 epilogue

struct TcGeom {
  int nchunks, ntaps, kiters, stages;
  int a_bytes, b_bytes, stage_bytes;
  int tiles_x, tiles_y, num_tiles;
  int tmem_cols;
  uint32_t idesc;
};

// ---------------------------------------------------------------------------------- PTX
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(addr), "r"(parity)
        : "memory");
  } while (!ok);
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* tm, uint64_t* bar, int c0,
                                            int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* tm, uint64_t* bar, int c0,
                                            int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* tm) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tm)) : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major shared-memory matrix descriptor (sm_100 "version 1").  layout: 2 = 128B swizzle,
// 4 = 64B, 6 = 32B.  sbo = byte distance between consecutive 8-row groups.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t sbo_bytes, uint32_t layout,
                                                   uint32_t base_offset = 0) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= 1ull << 46;
  d |= (uint64_t)(base_offset & 7u) << 49;
  d |= (uint64_t)layout << 61;
  return d;
}

__host__ __device__ constexpr uint32_t make_idesc_bf16(int M, int N) {
  // c_format F32 (1) @4, a_format BF16 (1) @7, b_format BF16 (1) @10, K-major A/B, N>>3 @17, M>>4 @24
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ---------------------------------------------------------------------------------- epilogue
__device__ __forceinline__ void unpack8(const uint4 r, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 r;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return r;
}
__device__ __forceinline__ void unpack4(const uint2 r, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
  const float2 a = __bfloat1622float2(h[0]), b = __bfloat1622float2(h[1]);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y;
}
__device__ __forceinline__ uint2 pack4(const float* f) {
  uint2 r;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
  h[0] = __floats2bfloat162_rn(f[0], f[1]);
  h[1] = __floats2bfloat162_rn(f[2], f[3]);
  return r;
}

// 16 consecutive output channels [c0, c0+16) of one pixel; same op order as conv_epilogue_store
__device__ __forceinline__ void epilogue16(const ConvParams& P, int n, int oy, int ox, int c0, float* v) {
  using bf = __nv_bfloat16;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float4 b = __ldg(reinterpret_cast<const float4*>(P.bias + c0) + q);
    v[4 * q] += b.x; v[4 * q + 1] += b.y; v[4 * q + 2] += b.z; v[4 * q + 3] += b.w;
  }
  if (P.flags & VCM_CONV_POST_ACT) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = leaky(v[i], P.post_slope);
  }
  if (!(P.flags & VCM_CONV_PIXEL_SHUFFLE)) {
    const size_t pix = ((size_t)n * P.Ho + oy) * P.Wo + ox;
    float r[16];
    if (P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE)) {
      const uint4* rp = reinterpret_cast<const uint4*>(reinterpret_cast<const bf*>(P.res) + pix * P.res_cs + c0);
      unpack8(__ldg(rp), r);
      unpack8(__ldg(rp + 1), r + 8);
      if (P.flags & VCM_CONV_RES) {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = r[i] + v[i];
      } else {
        const float* sp = P.res_scale + (size_t)n * P.Cout + c0;
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = v[i] + __fmul_rn(r[i], __ldg(sp + i));
      }
    }
    if (P.flags & VCM_CONV_RES2) {
      const uint4* rp = reinterpret_cast<const uint4*>(reinterpret_cast<const bf*>(P.res2) + pix * P.res2_cs + c0);
      unpack8(__ldg(rp), r);
      unpack8(__ldg(rp + 1), r + 8);
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = r[i] + v[i];
    }
    uint4* op = reinterpret_cast<uint4*>(reinterpret_cast<bf*>(P.y) + pix * P.out_cs + c0);
    op[0] = pack8(v);
    op[1] = pack8(v + 8);
    if (P.flags & VCM_CONV_OUT2) {
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = leaky(v[i], P.out2_slope);
      uint4* op2 = reinterpret_cast<uint4*>(reinterpret_cast<bf*>(P.y2) + pix * P.out2_cs + c0);
      op2[0] = pack8(v);
      op2[1] = pack8(v + 8);
    }
  } else {
    // PixelShuffle(2): channel co = 4*oc + 2*dy + dx -> pixel (2oy+dy, 2ox+dx), channel oc
    const int oc0 = c0 >> 2, Cres = P.Cout >> 2;
#pragma unroll
    for (int pos = 0; pos < 4; ++pos) {
      const int dy = pos >> 1, dx = pos & 1;
      const size_t pix = ((size_t)n * (2 * P.Ho) + 2 * oy + dy) * (2 * P.Wo) + 2 * ox + dx;
      float w[4] = {v[pos], v[4 + pos], v[8 + pos], v[12 + pos]};
      float r[4];
      if (P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE)) {
        unpack4(__ldg(reinterpret_cast<const uint2*>(reinterpret_cast<const bf*>(P.res) + pix * P.res_cs + oc0)), r);
        if (P.flags & VCM_CONV_RES) {
#pragma unroll
          for (int i = 0; i < 4; ++i) w[i] = r[i] + w[i];
        } else {
          const float* sp = P.res_scale + (size_t)n * Cres + oc0;
#pragma unroll
          for (int i = 0; i < 4; ++i) w[i] = w[i] + __fmul_rn(r[i], __ldg(sp + i));
        }
      }
      if (P.flags & VCM_CONV_RES2) {
        unpack4(__ldg(reinterpret_cast<const uint2*>(reinterpret_cast<const bf*>(P.res2) + pix * P.res2_cs + oc0)), r);
#pragma unroll
        for (int i = 0; i < 4; ++i) w[i] = r[i] + w[i];
      }
      *reinterpret_cast<uint2*>(reinterpret_cast<bf*>(P.y) + pix * P.out_cs + oc0) = pack4(w);
      if (P.flags & VCM_CONV_OUT2) {
#pragma unroll
        for (int i = 0; i < 4; ++i) w[i] = leaky(w[i], P.out2_slope);
        *reinterpret_cast<uint2*>(reinterpret_cast<bf*>(P.y2) + pix * P.out2_cs + oc0) = pack4(w);
      }
    }
  }
}

// ---------------------------------------------------------------------------------- kernel
template <int CHUNK>
__global__ void __launch_bounds__(NUM_THREADS, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const ConvParams P, const TcGeom G) {
  constexpr uint32_t LAYOUT = CHUNK == 64 ? 2u : (CHUNK == 32 ? 4u : 6u);
  constexpr uint32_t SBO = 8u * CHUNK * 2u;  // 8 rows of CHUNK bf16
  constexpr int KSTEPS = CHUNK / 16;
  constexpr int MAX_STAGES = 12;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)G.stages * G.stage_bytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + MAX_STAGES;
  uint64_t* tfull = bars + 2 * MAX_STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int s = 0; s < G.stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], 128);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, (uint32_t)G.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================================ TMA producer (one thread) ================================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < G.num_tiles; tile += gridDim.x) {
        const int n = tile / (G.tiles_x * G.tiles_y);
        const int rem = tile - n * (G.tiles_x * G.tiles_y);
        const int y0 = (rem / G.tiles_x) * TH, x0 = (rem % G.tiles_x) * TW;
        for (int tap = 0; tap < G.ntaps; ++tap) {
          const int ky = tap / P.K, kx = tap - ky * P.K;
          const int cy = y0 * P.stride + ky - P.pad, cx = x0 * P.stride + kx - P.pad;
          for (int ch = 0; ch < G.nchunks; ++ch) {
            mbar_wait(&empty[stage], phase ^ 1u);
            uint8_t* sa = smem + (size_t)stage * G.stage_bytes;
            mbar_expect_tx(&full[stage], (uint32_t)(G.a_bytes + G.b_bytes));
            tma_load_4d(smem_u32(sa), &tmA, &full[stage], ch * CHUNK, cx, cy, n);
            tma_load_2d(smem_u32(sa + G.a_bytes), &tmB, &full[stage], 0, (tap * G.nchunks + ch) * P.Cout);
            if (++stage == G.stages) { stage = 0; phase ^= 1u; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer (one thread) ==================================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < G.num_tiles; tile += gridDim.x) {
        mbar_wait(&tempty[acc], acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * P.Cout);
        for (int it = 0; it < G.kiters; ++it) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + (size_t)stage * G.stage_bytes);
          const uint64_t adesc = make_smem_desc(sa, SBO, LAYOUT);
          const uint64_t bdesc = make_smem_desc(sa + G.a_bytes, SBO, LAYOUT);
#pragma unroll
          for (int k = 0; k < KSTEPS; ++k) {
            // advance 16 bf16 = 32 bytes along K inside the swizzle row: +2 in the >>4 address field
            umma_bf16(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), G.idesc,
                      (it > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty[stage]);
          if (++stage == G.stages) { stage = 0; phase ^= 1u; }
        }
        umma_commit(&tfull[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1u;
      }
    }
  } else {
    // ================================ epilogue (4 warps) =======================================
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    const int m = q * 32 + lane;
    const int py = m / TW, px = m % TW;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < G.num_tiles; tile += gridDim.x) {
      const int n = tile / (G.tiles_x * G.tiles_y);
      const int rem = tile - n * (G.tiles_x * G.tiles_y);
      const int oy = (rem / G.tiles_x) * TH + py, ox = (rem % G.tiles_x) * TW + px;
      const bool valid = oy < P.Ho && ox < P.Wo;
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * P.Cout);
      for (int c0 = 0; c0 < P.Cout; c0 += 16) {
        uint32_t r[16];
        tmem_ld16(t0 + (uint32_t)c0, r);
        if (valid) {
          float v[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
          epilogue16(P, n, oy, ox, c0, v);
        }
      }
      tc_fence_before();
      mbar_arrive(&tempty[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1u;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)G.tmem_cols);
  }
}

// OIHW fp32 -> bf16 [tap][chunk][Cout][CHUNK] (K-major rows the weight TMA box is cut from)
__global__ void pack_umma_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out, int Cout,
                                 int Cin, int K, int Cin_p, int chunk, const int* __restrict__ in_map) {
  const int nchunks = Cin_p / chunk;
  const int64_t total = (int64_t)K * K * Cin_p * Cout;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int kk = (int)(i % chunk);
    int64_t r = i / chunk;
    const int co = (int)(r % Cout);
    r /= Cout;
    const int ch = (int)(r % nchunks);
    const int tap = (int)(r / nchunks);
    const int ci_p = ch * chunk + kk;
    const int ci = in_map ? in_map[ci_p] : ci_p;
    float v = 0.f;
    if (ci >= 0 && ci < Cin) v = w[((int64_t)co * Cin + ci) * K * K + tap];
    out[i] = __float2bfloat16_rn(v);
  }
}

// ---------------------------------------------------------------------------------- probe
// One 128 x 64 x 64 MMA with an identity B: D[m][n] = A[row(m)][n], where the tensor core fetches
// row(m) from a 128B-swizzled buffer starting `start_row` rows in, 8-row groups `sbo` bytes apart.
__global__ void __launch_bounds__(128, 1) umma_probe_kernel(int start_row, int sbo_bytes, int use_base_offset,
                                                            float* __restrict__ out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  constexpr int A_ROWS = 320;
  uint8_t* sA = smem;                      // A_ROWS x 128 B
  uint8_t* sB = smem + A_ROWS * 128;       // 64 x 128 B
  uint64_t* bar = reinterpret_cast<uint64_t*>(sB + 64 * 128);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
  for (int i = threadIdx.x; i < A_ROWS * 64; i += blockDim.x) {
    const int p = i / 64, c = i % 64;
    const uint32_t off = p * 128 + (((c >> 3) ^ (p & 7)) << 4) + (c & 7) * 2;
    *reinterpret_cast<__nv_bfloat16*>(sA + off) = __float2bfloat16_rn((float)((p * 64 + c) % 251));
  }
  for (int i = threadIdx.x; i < 64 * 64; i += blockDim.x) {
    const int nrow = i / 64, k = i % 64;
    const uint32_t off = nrow * 128 + (((k >> 3) ^ (nrow & 7)) << 4) + (k & 7) * 2;
    *reinterpret_cast<__nv_bfloat16*>(sB + off) = __float2bfloat16_rn(nrow == k ? 1.f : 0.f);
  }
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    fence_barrier_init();
  }
  fence_proxy_async();
  if ((threadIdx.x >> 5) == 0) tmem_alloc(slot, 64);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *slot;
  if (threadIdx.x == 0) {
    const uint32_t a0 = smem_u32(sA) + (uint32_t)start_row * 128u;
    const uint32_t bo = use_base_offset ? ((a0 >> 7) & 7u) : 0u;
    const uint64_t adesc = make_smem_desc(a0, (uint32_t)sbo_bytes, 2u, bo);
    const uint64_t bdesc = make_smem_desc(smem_u32(sB), 1024u, 2u);
    const uint32_t idesc = make_idesc_bf16(128, 64);
#pragma unroll
    for (int k = 0; k < 4; ++k) umma_bf16(tmem_base, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, k > 0);
    umma_commit(bar);
  }
  mbar_wait(bar, 0);
  tc_fence_after();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int c0 = 0; c0 < 64; c0 += 16) {
    uint32_t r[16];
    tmem_ld16(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, r);
#pragma unroll
    for (int i = 0; i < 16; ++i) out[(warp * 32 + lane) * 64 + c0 + i] = __uint_as_float(r[i]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 64);
  }
}

// ---------------------------------------------------------------------------------- host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int pick_chunk(int cin) { return cin % 64 == 0 ? 64 : (cin % 32 == 0 ? 32 : 16); }
CUtensorMapSwizzle swizzle_of(int chunk) {
  return chunk == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : (chunk == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
}
int round_up(int v, int m) { return (v + m - 1) / m * m; }

int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

constexpr int SMEM_BAR_BYTES = 512;
constexpr int SMEM_LIMIT = 227 * 1024;

}  // namespace

extern "C" int vcm_conv2d_tc_supported(const vcm_conv_desc* d) {
  if (!d) return 0;
  if (d->in_dtype != VCM_BF16 || d->out_dtype != VCM_BF16) return 0;
  if (!(d->K == 1 || d->K == 3 || d->K == 7)) return 0;
  if (!(d->stride == 1 || d->stride == 2)) return 0;
  if (d->Cin % 16 || d->Cout % 16 || d->Cout < 16 || d->Cout > 256) return 0;
  if (d->flags & VCM_CONV_PRE_ACT) return 0;
  if (d->in_cstride % 8 || d->out_cstride % 8) return 0;
  if ((d->flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE)) && d->res_cstride % 8) return 0;
  if ((d->flags & VCM_CONV_RES2) && d->res2_cstride % 8) return 0;
  if ((d->flags & VCM_CONV_OUT2) && d->out2_cstride % 8) return 0;
  if (d->stride == 2 && ((d->H & 1) || (d->W & 1))) return 0;
  return 1;
}

extern "C" int64_t vcm_conv2d_tc_weight_bytes(int32_t Cout, int32_t Cin, int32_t K) {
  if (Cout <= 0 || Cin <= 0 || Cin % 16 || Cout % 16 || Cout > 256) return 0;
  return (int64_t)K * K * Cin * Cout * 2;
}

extern "C" int vcm_pack_conv_weight_umma(const float* w_oihw, void* w_umma, int32_t Cout, int32_t Cin,
                                         int32_t K, int32_t Cin_p, const int32_t* in_map, void* stream) {
  VCM_CHECK_ARG(w_oihw && w_umma && Cout > 0 && Cin > 0 && K > 0 && Cin_p > 0 && Cin_p % 16 == 0,
                "bad argument");
  const int chunk = pick_chunk(Cin_p);
  const int64_t total = (int64_t)K * K * Cin_p * Cout;
  const int grid = cdiv(total, 256) > 4096 ? 4096 : cdiv(total, 256);
  pack_umma_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(w_oihw, (__nv_bfloat16*)w_umma, Cout, Cin, K,
                                                           Cin_p, chunk, in_map);
  VCM_CHECK_LAUNCH();
  return 0;
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
  const bool ps = P.flags & VCM_CONV_PIXEL_SHUFFLE;
  const uintptr_t amask = ps ? 7 : 15;
  VCM_CHECK_ARG(((uintptr_t)x & 15) == 0 && ((uintptr_t)w_umma & 15) == 0 && ((uintptr_t)y & amask) == 0 &&
                    ((uintptr_t)res & amask) == 0 && ((uintptr_t)res2 & amask) == 0 &&
                    ((uintptr_t)y2 & amask) == 0 && ((uintptr_t)bias & 15) == 0,
                "pointer alignment (16 B) not met");
  P.x = x; P.bias = bias; P.res = res; P.res_scale = res_scale; P.res2 = res2; P.y = y; P.y2 = y2;

  EncodeTiledFn enc = get_encode_fn();
  VCM_CHECK_ARG(enc != nullptr, "cuTensorMapEncodeTiled unavailable");

  TcGeom G;
  const int chunk = pick_chunk(P.Cin);
  G.nchunks = P.Cin / chunk;
  G.ntaps = P.K * P.K;
  G.kiters = G.ntaps * G.nchunks;
  G.a_bytes = 128 * chunk * 2;
  G.b_bytes = round_up(P.Cout * chunk * 2, 1024);
  G.stage_bytes = G.a_bytes + G.b_bytes;
  int stages = (SMEM_LIMIT - 1024 - SMEM_BAR_BYTES) / G.stage_bytes;
  if (stages > 12) stages = 12;
  if (stages > G.kiters * 2) stages = G.kiters * 2 > 2 ? G.kiters * 2 : 2;
  VCM_CHECK_ARG(stages >= 2, "stage does not fit shared memory");
  G.stages = stages;
  G.tiles_x = cdiv(P.Wo, TW);
  G.tiles_y = cdiv(P.Ho, TH);
  G.num_tiles = G.tiles_x * G.tiles_y * P.N;
  int cols = 32;
  while (cols < 2 * P.Cout) cols <<= 1;
  G.tmem_cols = cols;
  G.idesc = make_idesc_bf16(128, P.Cout);

  CUtensorMap tmA, tmB;
  {
    cuuint64_t dims[4] = {(cuuint64_t)P.Cin, (cuuint64_t)P.W, (cuuint64_t)P.H, (cuuint64_t)P.N};
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
    cuuint64_t dims[2] = {(cuuint64_t)chunk, (cuuint64_t)G.kiters * P.Cout};
    cuuint64_t strides[1] = {(cuuint64_t)chunk * 2};
    cuuint32_t box[2] = {(cuuint32_t)chunk, (cuuint32_t)P.Cout};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(w_umma), dims, strides, box,
                     estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_of(chunk), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      vcm_set_error("vcm_conv2d_tc_fwd: cuTensorMapEncodeTiled(B) failed (%d)", (int)r);
      return 1;
    }
  }
  // the weight slab really is P.Cout*chunk*2 bytes; a_bytes + that is what TMA reports
  const int b_tx = P.Cout * chunk * 2;
  TcGeom Gk = G;
  Gk.b_bytes = b_tx;                    // used for expect_tx ...
  Gk.stage_bytes = G.stage_bytes;       // ... while the ring stride stays 1024-aligned
  // B lives right after A inside the stage; A sizes (16/8/4 KB) keep it 1024-aligned.
  int smem_bytes = G.stages * G.stage_bytes + 1024 + SMEM_BAR_BYTES;
  if (smem_bytes < 120 * 1024) smem_bytes = 120 * 1024;  // one CTA per SM (TMEM is allocated per CTA)
  const int grid = G.num_tiles < num_sms() ? G.num_tiles : num_sms();
  cudaStream_t s = (cudaStream_t)stream;
#define LAUNCH_TC(CH)                                                                                   \
  do {                                                                                                  \
    static bool attr_set = false;                                                                       \
    if (!attr_set) {                                                                                    \
      cudaError_t e = cudaFuncSetAttribute(conv_tc_kernel<CH>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                           SMEM_LIMIT);                                                 \
      if (e != cudaSuccess) {                                                                           \
        vcm_set_error("vcm_conv2d_tc_fwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));            \
        return 1;                                                                                       \
      }                                                                                                 \
      attr_set = true;                                                                                  \
    }                                                                                                   \
    conv_tc_kernel<CH><<<grid, NUM_THREADS, smem_bytes, s>>>(tmA, tmB, P, Gk);                          \
  } while (0)
  if (chunk == 64) LAUNCH_TC(64);
  else if (chunk == 32) LAUNCH_TC(32);
  else LAUNCH_TC(16);
#undef LAUNCH_TC
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_umma_probe(int32_t start_row, int32_t sbo_bytes, int32_t use_base_offset, float* out,
                              void* stream) {
  VCM_CHECK_ARG(out && start_row >= 0 && sbo_bytes >= 128 && sbo_bytes % 16 == 0, "bad argument");
  VCM_CHECK_ARG(start_row + 15 * (sbo_bytes / 128) + 8 <= 320, "probe window exceeds the buffer");
  const int smem_bytes = 320 * 128 + 64 * 128 + 64 + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    attr_set = true;
  }
  umma_probe_kernel<<<1, 128, smem_bytes, (cudaStream_t)stream>>>(start_row, sbo_bytes, use_base_offset, out);
  VCM_CHECK_LAUNCH();
  return 0;
}

// vcm_ts_b200/csrc/tc_epilogue.cuh -- TMEM -> registers -> fused conv epilogue, shared by the two
// tcgen05 kernels (conv_tc.cu: per-tap TMA, conv_halo.cu: halo tile + resident weights).
//
// Vector path (bf16 output, Cout % 16 == 0, 16-byte aligned views).  tcgen05.ld hands every thread
// ONE pixel (TMEM lane) and BW consecutive channels; written straight to NHWC global memory that is
// 32 different 128-byte lines per warp instruction (ncu of the first engine: 32 sectors / request,
// LSU wavefronts 44 % of peak with the tensor pipe at 10 %).  So every epilogue warp owns a
// [32 pixel][BW channel] fp32 staging tile in shared memory (XOR-swizzled 16-byte chunks, bank
// conflict free both ways): it drops the accumulators there, then re-reads them so that a lane
// holds 8 consecutive channels of a pixel and the warp covers 32/UNITS whole pixels -- every global
// access (residuals, SE-scaled residual, output, activated twin) is a 16-byte piece of a contiguous
// BW*2-byte run per pixel (per-instruction address steps are warp-uniform constants: the profile of
// the first version was instruction-bound, ~1100 warp instructions per 128x64 tile, 21 % address math).  Bias / activation / residual arithmetic happens in that second mapping,
// in the same float-op order as conv_epilogue_store.
//
// PixelShuffle: the tensor-core weight packing orders the GEMM columns position-major
// (column c' = pos * Cout/4 + oc, pos = 2*dy + dx; backend.py pack_conv), so one BW block is BW
// consecutive output channels of ONE sub-pixel position and the store is the same contiguous run at
// pixel (2*oy + dy, 2*ox + dx).
//
// Scalar path (fp32 output or Cout % 16 != 0: flow / RGB heads): per-element conv_epilogue_store.
#pragma once
#include "conv_common.cuh"
#include "tc_ptx.cuh"

namespace tc {

using bf = __nv_bfloat16;

__device__ __forceinline__ void unpack8(const uint4 r, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 a = __bfloat1622float2(h[i]);
    f[2 * i] = a.x;
    f[2 * i + 1] = a.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 r;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return r;
}

// TMEM -> registers for BW columns of this thread's lane (one wait for all loads)
template <int BW>
__device__ __forceinline__ void tmem_load_block(uint32_t taddr, float* v) {
#pragma unroll
  for (int j = 0; j < BW / 16; ++j) {
    uint32_t* r = reinterpret_cast<uint32_t*>(v) + 16 * j;
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr + 16u * j));
  }
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// the same without the wait: the registers are valid only after tmem_load_wait()
template <int BW>
__device__ __forceinline__ void tmem_load_issue(uint32_t taddr, float* v) {
#pragma unroll
  for (int j = 0; j < BW / 16; ++j) {
    uint32_t* r = reinterpret_cast<uint32_t*>(v) + 16 * j;
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr + 16u * j));
  }
}
__device__ __forceinline__ void tmem_load_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ------------------------------------------------------------------------------- vector path
constexpr int EPI_STAGE_FLOATS = 32 * 32;                       // per warp: [32 px][<=32 ch] fp32
constexpr int EPI_STAGE_BYTES = 8 * EPI_STAGE_FLOATS * 4;       // 8 epilogue warps
constexpr int EPI_BIAS_MAX = 2048;                              // bias vector kept in shared memory (floats)

__device__ __forceinline__ float4 lds128(uint32_t saddr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(saddr));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t saddr, float a, float b, float c, float d) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

template <int BW>
__device__ __forceinline__ int stage_swz(int p) {
  return BW == 32 ? (p & 7) : ((p >> 1) & 3);
}

// Coalesced mapping of one warp's 32-pixel x BW-channel block: lane -> (pixel slot l = lane / UNITS,
// 8-channel unit u = lane % UNITS); warp instruction `it` covers pixels it*PIXPER .. +PIXPER-1 of the
// warp's 32 (row-major in a TWID-wide tile), so the (row, col) step of every instruction is a
// compile-time constant and all per-instruction address deltas are warp-uniform.
template <int BW, int TWID>
struct VecMap {
  static constexpr int UNITS = BW / 8;        // 16-byte units per pixel
  static constexpr int PIXPER = 32 / UNITS;   // pixels per warp instruction
  static constexpr int ITERS = UNITS;         // warp instructions per 32-pixel block
  static_assert(PIXPER % TWID == 0 || TWID % PIXPER == 0, "tile width / block width mismatch");
  __host__ __device__ static constexpr int drow(int it) { return (it * PIXPER) / TWID; }
  __host__ __device__ static constexpr int dcol(int it) { return (it * PIXPER) % TWID; }
};

// leaky_relu for slopes in [0, 1] (all the path uses): max(v, v*slope) == (v > 0 ? v : v*slope)
__device__ __forceinline__ float leaky01(float v, float slope) { return fmaxf(v, v * slope); }

// One warp, one 32-pixel x bn accumulator slab.  taddr: TMEM address of (lane quarter, first column).
// stage: shared-window address of this warp's staging tile.
// wait != nullptr: wait for the accumulator (after the first block's residuals are in flight).
// release != nullptr: arrive on it as soon as the last TMEM read of the slab has completed.
template <int BW, int TWID>
__device__ __forceinline__ void epilogue_warp_vec(const ConvParams& P, int bn, int n, int oy0, int ox0,
                                                  int cbase, uint32_t taddr, uint32_t stage, int wq, int lane,
                                                  uint64_t* wait, uint32_t parity, uint64_t* release,
                                                  uint32_t bias_smem) {
  using M = VecMap<BW, TWID>;
  constexpr int UNITS = M::UNITS, ITERS = M::ITERS;
  const int u = lane % UNITS, l = lane / UNITS;
  const int oyb = oy0 + wq * (32 / TWID) + l / TWID, oxb = ox0 + l % TWID;
  const bool ps = P.flags & VCM_CONV_PIXEL_SHUFFLE;
  const int sx = ps ? 2 : 1;
  const int Wout = sx * P.Wo, Hout = sx * P.Ho;
  const int q4 = P.Cout >> 2;
  const bool has_res = P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE);
  const bool has_res2 = P.flags & VCM_CONV_RES2;
  const float slope = (P.flags & VCM_CONV_POST_ACT) ? P.post_slope : 1.0f;   // 1.0: identity, exact
  uint32_t vmask = 0;
#pragma unroll
  for (int it = 0; it < ITERS; ++it)
    if (oyb + M::drow(it) < P.Ho && oxb + M::dcol(it) < P.Wo) vmask |= 1u << it;
  // output pixel index of this lane's first pixel (sub-pixel position added per block), and the
  // warp-uniform pixel-index step of instruction `it`
  const size_t pixb = ((size_t)n * Hout + (size_t)(sx * oyb)) * (size_t)Wout + (size_t)(sx * oxb);
  uint32_t dpix[ITERS];
#pragma unroll
  for (int it = 0; it < ITERS; ++it) dpix[it] = (uint32_t)(sx * (M::drow(it) * Wout + M::dcol(it)));

  // block-dependent part: first pixel index (PixelShuffle position) and this lane's first channel
  auto block_pos = [&](int c0, size_t& pix0, int& ch) {
    int chd = c0, dy = 0, dx = 0;
    if (ps) {
      const int pos = c0 / q4;
      chd = c0 - pos * q4;
      dy = pos >> 1;
      dx = pos & 1;
    }
    pix0 = pixb + (size_t)(dy * Wout + dx);
    ch = chd + 8 * u;
  };
  auto prefetch1 = [&](size_t pix0, int ch, uint4 (&r1)[ITERS]) {
    if (has_res) {
      const bf* rp = reinterpret_cast<const bf*>(P.res) + pix0 * (size_t)P.res_cs + ch;
#pragma unroll
      for (int it = 0; it < ITERS; ++it)
        if (vmask >> it & 1u) r1[it] = __ldg(reinterpret_cast<const uint4*>(rp + (size_t)dpix[it] * P.res_cs));
    }
  };
  auto prefetch2 = [&](size_t pix0, int ch, uint4 (&r2)[ITERS]) {
    if (has_res2) {
      const bf* rp = reinterpret_cast<const bf*>(P.res2) + pix0 * (size_t)P.res2_cs + ch;
#pragma unroll
      for (int it = 0; it < ITERS; ++it)
        if (vmask >> it & 1u) r2[it] = __ldg(reinterpret_cast<const uint4*>(rp + (size_t)dpix[it] * P.res2_cs));
    }
  };

  size_t pix0;
  int ch;
  uint4 r1[ITERS], r2[ITERS];
  block_pos(cbase, pix0, ch);
  prefetch1(pix0, ch, r1);
  prefetch2(pix0, ch, r2);
  if (wait) {
    mbar_wait(wait, parity);
    tc_fence_after();
  }
  const int sw_st = stage_swz<BW>(lane);
  const uint32_t st_row = stage + (uint32_t)(lane * BW) * 4u;
  float v[BW];
  tmem_load_issue<BW>(taddr, v);
  for (int cb = 0; cb < bn; cb += BW) {
    const bool more = cb + BW < bn;
    tmem_load_wait();
#pragma unroll
    for (int j = 0; j < BW / 4; ++j)
      sts128(st_row + (uint32_t)(4 * (j ^ sw_st)) * 4u, v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    if (more) tmem_load_issue<BW>(taddr + (uint32_t)(cb + BW), v);   // in flight while this block is finished
    __syncwarp();
    if (!more && release && lane == 0) {   // one arrival per warp (barrier count = 4 warps)
      tc_fence_before();
      mbar_arrive(release);
    }
    // the whole block in the coalesced mapping, then the staging tile is free again
    float4 a0[ITERS], a1[ITERS];
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
      const int p = it * M::PIXPER + l;
      const int sw = stage_swz<BW>(p);
      const uint32_t row = stage + (uint32_t)(p * BW) * 4u;
      a0[it] = lds128(row + (uint32_t)(4 * ((2 * u) ^ sw)) * 4u);
      a1[it] = lds128(row + (uint32_t)(4 * ((2 * u + 1) ^ sw)) * 4u);
    }
    const float4 b0 = lds128(bias_smem + (uint32_t)(cbase + cb + 8 * u) * 4u);
    const float4 b1 = lds128(bias_smem + (uint32_t)(cbase + cb + 8 * u + 4) * 4u);
    __syncwarp();
    size_t pixn = 0;
    int chn = 0;
    uint4 n1[ITERS];   // first residual double-buffered; the (rare) second one is refetched after use
    if (more && (has_res || has_res2)) {   // next block's residuals in flight while this one is finished
      block_pos(cbase + cb + BW, pixn, chn);
      prefetch1(pixn, chn, n1);
    }
    // ---- finish this block in the coalesced mapping
    {
      const int c0 = cbase + cb;
      float b[8], sc[8];
      {
        b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
      }
      if (P.flags & VCM_CONV_RES_SCALE) {
        const float* sp = P.res_scale + (size_t)n * (ps ? q4 : P.Cout) + ch;
        const float4 s0 = __ldg(reinterpret_cast<const float4*>(sp));
        const float4 s1 = __ldg(reinterpret_cast<const float4*>(sp + 4));
        sc[0] = s0.x; sc[1] = s0.y; sc[2] = s0.z; sc[3] = s0.w; sc[4] = s1.x; sc[5] = s1.y; sc[6] = s1.z; sc[7] = s1.w;
      }
      bf* yp = reinterpret_cast<bf*>(P.y) + pix0 * (size_t)P.out_cs + ch;
      bf* y2p = (P.flags & VCM_CONV_OUT2) ? reinterpret_cast<bf*>(P.y2) + pix0 * (size_t)P.out2_cs + ch : nullptr;
#pragma unroll
      for (int it = 0; it < ITERS; ++it) {
        if (!(vmask >> it & 1u)) continue;
        float w[8] = {a0[it].x, a0[it].y, a0[it].z, a0[it].w, a1[it].x, a1[it].y, a1[it].z, a1[it].w};
#pragma unroll
        for (int i = 0; i < 8; ++i) w[i] = leaky01(w[i] + b[i], slope);
        if (has_res) {
          float r[8];
          unpack8(r1[it], r);
          if (P.flags & VCM_CONV_RES) {
#pragma unroll
            for (int i = 0; i < 8; ++i) w[i] = r[i] + w[i];
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) w[i] = w[i] + __fmul_rn(r[i], sc[i]);
          }
        }
        if (has_res2) {
          float r[8];
          unpack8(r2[it], r);
#pragma unroll
          for (int i = 0; i < 8; ++i) w[i] = r[i] + w[i];
        }
        *reinterpret_cast<uint4*>(yp + (size_t)dpix[it] * P.out_cs) = pack8(w);
        if (y2p) {
#pragma unroll
          for (int i = 0; i < 8; ++i) w[i] = leaky01(w[i], P.out2_slope);
          *reinterpret_cast<uint4*>(y2p + (size_t)dpix[it] * P.out2_cs) = pack8(w);
        }
      }
    }
    if (more) {
      if (has_res || has_res2) {
        pix0 = pixn;
        ch = chn;
#pragma unroll
        for (int it = 0; it < ITERS; ++it) r1[it] = n1[it];
        prefetch2(pix0, ch, r2);
      } else {
        block_pos(cbase + cb + BW, pix0, ch);
      }
    }
  }
}

// ------------------------------------------------------------------------------- TMA-store path
// Measured (scripts/conv_bench.py, round 2): with its global loads / stores switched off the vector path above
// still costs ~3000 cycles per 128 x 64 tile -- ~1100 warp instructions of TMEM -> registers -> staging ->
// registers, address arithmetic and per-lane 16-byte global accesses -- which is MORE than the 1980 cycles the 36
// MMAs of a 3x3 64->64 tile take: every N = 64 layer was epilogue-bound.  Here the tile leaves through ONE
// TMA store: each thread (= one pixel = one TMEM lane) converts its 64 channels and writes them as eight
// 16-byte chunks into a 128B-swizzled [128 px][64 ch] bf16 image of the tile in shared memory (conflict
// free), a warpgroup barrier, one elected thread issues cp.async.bulk.tensor (the tensor map clips the tile at
// the image edge).  The residual operand arrives the same way, by a TMA load issued one item ahead, and is
// read back row-wise from shared memory.  Same float operations in the same order as the vector path.
// Handles: bf16 output, N tile of exactly 64 channels, POST_ACT, RES / RES_SCALE, OUT2; not PixelShuffle, RES2.
constexpr int EPI_TMA_TILE_BYTES = 128 * 128;                 // [128 px][64 ch] bf16
constexpr int EPI_TMA_BYTES = 4 * EPI_TMA_TILE_BYTES;         // per warpgroup: output tile + residual tile

struct EpiTma {
  const CUtensorMap *tmY, *tmY2, *tmR;
  uint32_t obuf;             // shared-window address of this warpgroup's TWO staging tiles (obuf, obuf + tile_bytes)
  uint32_t tile_bytes;       // distance between them (EPI_TMA_TILE_BYTES, or less for narrow tiles)
  uint64_t* rfull;           // [2] residual tile landed in staging tile b
  uint32_t scale_smem;       // shared-window address of a copy of res_scale [N][Cout] (0: read it from global)
  int k;                     // tiles this warpgroup has finished (staging tile k & 1, barrier phase (k >> 1) & 1)
  bool se_sum;               // VCM_CONV_SE_SUM: keep running per-channel sums of the (bf16-rounded) outputs
  float se0, se1;            // ... of this lane's channel pair over the rows its warp has produced so far
  bool leader;               // the one thread of the warpgroup that issues TMA operations
  int bar;                   // named barrier of this warpgroup (1 or 2)
};

// residual tile of this warpgroup's tile number `k` -> staging tile k & 1 (where the output is then formed)
template <int BN>
__device__ __forceinline__ void epi_tma_prefetch_res(const EpiTma& E, int k, int c, int x0, int y0, int n) {
  const int b = k & 1;
  mbar_expect_tx(&E.rfull[b], (uint32_t)(128 * BN * 2));
  tma_load_4d(E.obuf + (uint32_t)b * E.tile_bytes, E.tmR, &E.rfull[b], c, x0, y0, n);
}

// VCM_CONV_SE_SUM: one row of channel sums per epilogue warp, sums[(cta * 8 + epilogue warp)][BN]; rows of warps
// without tiles are zero.  The SE FC kernel adds the rows in index order (deterministic).
template <int BN>
__device__ __forceinline__ void epi_tma_write_se_sums(const EpiTma& E, float* sums, int warp_e, int lane) {
  constexpr int PAIRS = BN / 2;
  float a = E.se0, b = E.se1;
#pragma unroll
  for (int o = 16; o >= PAIRS; o >>= 1) {          // lanes l, l + PAIRS, ... hold the same pair over other rows
    a += __shfl_down_sync(0xffffffffu, a, o);
    b += __shfl_down_sync(0xffffffffu, b, o);
  }
  if (lane < PAIRS) {
    float* row = sums + ((size_t)blockIdx.x * 8 + (size_t)warp_e) * BN;
    row[2 * lane] = a;
    row[2 * lane + 1] = b;
  }
}

// One warp's 32 pixels of a 128-pixel x 64-channel tile (wq = TMEM lane quarter = tile rows 32*wq ..).
// The warpgroup owns two staging tiles and alternates between them.  A residual tile is loaded by TMA INTO
// the staging tile its output will be formed in and updated in place (each 16-byte chunk is read and
// rewritten by the one thread that owns it), so the second tile buys lead time instead of holding a copy:
// the residual of tile k + 1 is requested at the START of tile k, as soon as the store of tile k - 1 has
// finished reading that buffer.  (Before: one output + one residual buffer, the request went out when the
// next tile was about to start, and every tile waited a DRAM latency for it -- ncu: the wait on rfull was
// the top stall of the 1x1 + residual convs, 164 -> 134 us at 1088x1920 once it had lead.)  Without a
// residual the two tiles double-buffer the output, so nobody waits for the previous store to drain.
// (ox0, oy0, cstore): tensor-map coordinates of the tile (output pixel and channel; PixelShuffle tiles go through a
// map with element strides 2, see conv_halo.cu); cbase: first GEMM column (bias index).
// next_*: the tensor-map coordinates of this warpgroup's NEXT tile (has_next = false at the end).
// OUT2 (activated twin) is only taken together with "no residual": staging tile 0 = output, 1 = twin.
// The 64 channels are processed as two halves of 32 to keep the register footprint small.
// BN = 64 (128-byte rows, 128B swizzle), 32 (64-byte rows, 64B swizzle: chunk ^= (row >> 1) & 3) or 16 (32-byte
// rows, 32B swizzle: chunk ^= (row >> 2) & 1).
template <int BN>
__device__ __forceinline__ void epilogue_warp_tma(const ConvParams& P, EpiTma& E, int n, int oy0, int ox0, int cstore,
                                                  int cbase, uint32_t taddr, int wq, int lane, uint64_t* wait,
                                                  uint32_t parity, uint64_t* release, uint32_t bias_smem,
                                                  bool has_next, int nn, int noy0, int nox0, int ncbase) {
  const bool has_res = P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE);
  const bool twin = P.flags & VCM_CONV_OUT2;
  const float slope = (P.flags & VCM_CONV_POST_ACT) ? P.post_slope : 1.0f;
  const int m = wq * 32 + lane;                 // tile row = pixel
  const uint32_t row = (uint32_t)m * (uint32_t)(BN * 2);
  const uint32_t sw = BN == 64 ? (uint32_t)(m & 7) : (BN == 32 ? (uint32_t)((m >> 1) & 3) : (uint32_t)((m >> 2) & 1));
  constexpr int CB = BN < 32 ? BN : 32;         // channels per pass
  const int b = twin ? 0 : (E.k & 1);
  const uint32_t obuf = E.obuf + (uint32_t)b * E.tile_bytes;
  const uint32_t tbuf = E.obuf + E.tile_bytes;      // twin tile (twin mode only)
  if (twin) {
    // fixed buffers: the previous tile's stores must have finished reading them before they are overwritten
    if (E.leader) tma_store_wait_read();
    named_bar_sync(E.bar, 128);
  } else if (has_res && E.leader) {
    // staging tile 1 - b: its last store (tile k - 1) must be done reading before the next residual lands in
    // it.  Staging tile b needs no check here: the same wait one tile ago covered its last store (tile k - 2),
    // and that tile's closing barrier published it.
    tma_store_wait_read();
    if (has_next) epi_tma_prefetch_res<BN>(E, E.k + 1, ncbase, nox0, noy0, nn);
  }
  if (wait) {
    mbar_wait(wait, parity);
    tc_fence_after();
  }
  if (has_res) mbar_wait(&E.rfull[b], (uint32_t)(E.k >> 1) & 1u);
  // not unrolled: one copy of the 32-channel body keeps the hot loop of the epilogue warps small (the
  // instruction cache is shared with the MMA-issuing thread's long unrolled sequence)
#pragma unroll 1
  for (int half = 0; half < BN / CB; ++half) {
    float v[CB];
    tmem_load_block<CB>(taddr + (uint32_t)(CB * half), v);
    if (half == BN / CB - 1 && release) {      // last TMEM read of the tile
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(release);
    }
#pragma unroll
    for (int j = 0; j < CB / 4; ++j) {
      const float4 bb = lds128(bias_smem + (uint32_t)(cbase + CB * half + 4 * j) * 4u);
      v[4 * j] = leaky01(v[4 * j] + bb.x, slope);
      v[4 * j + 1] = leaky01(v[4 * j + 1] + bb.y, slope);
      v[4 * j + 2] = leaky01(v[4 * j + 2] + bb.z, slope);
      v[4 * j + 3] = leaky01(v[4 * j + 3] + bb.w, slope);
    }
#pragma unroll
    for (int j = 0; j < CB / 8; ++j) {
      const uint32_t chunk = ((uint32_t)(CB / 8 * half + j) ^ sw) << 4;
      if (has_res) {
        uint4 rr;
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(rr.x), "=r"(rr.y), "=r"(rr.z), "=r"(rr.w)
                     : "r"(obuf + row + chunk)
                     : "memory");
        float r[8];
        unpack8(rr, r);
        if (P.flags & VCM_CONV_RES) {
#pragma unroll
          for (int i = 0; i < 8; ++i) v[8 * j + i] = r[i] + v[8 * j + i];
        } else {
          // SE scales: from the shared-memory copy (16 dependent global loads per tile and thread otherwise)
          const int so = n * P.Cout + cstore + CB * half + 8 * j;
          float4 s0, s1;
          if (E.scale_smem) {
            s0 = lds128(E.scale_smem + (uint32_t)so * 4u);
            s1 = lds128(E.scale_smem + (uint32_t)so * 4u + 16u);
          } else {
            s0 = __ldg(reinterpret_cast<const float4*>(P.res_scale + so));
            s1 = __ldg(reinterpret_cast<const float4*>(P.res_scale + so + 4));
          }
          const float sc[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
#pragma unroll
          for (int i = 0; i < 8; ++i) v[8 * j + i] = v[8 * j + i] + __fmul_rn(r[i], sc[i]);
        }
      }
      const uint4 o = pack8(v + 8 * j);
      asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(obuf + row + chunk), "r"(o.x), "r"(o.y),
                   "r"(o.z), "r"(o.w)
                   : "memory");
      if (twin) {
        float t[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) t[i] = leaky01(v[8 * j + i], P.out2_slope);
        const uint4 o2 = pack8(t);
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(tbuf + row + chunk), "r"(o2.x), "r"(o2.y),
                     "r"(o2.z), "r"(o2.w)
                     : "memory");
      }
    }
  }
  if (E.se_sum) {
    // SELayer mean fused into the producing conv: the warp's 32 rows of the tile are in shared memory as the
    // bf16 values that go to HBM; lane l adds up one channel pair over them (row order, fp32).  Tiles are
    // whole (checked on the host), so every row counts.
    __syncwarp();
    constexpr int PAIRS = BN / 2, GROUPS = 32 / PAIRS, RPL = 32 / GROUPS;
    const int pair = lane % PAIRS;
    const int r0 = wq * 32 + (lane / PAIRS) * RPL;
#pragma unroll 8
    for (int i = 0; i < RPL; ++i) {
      const int r = r0 + i;
      const uint32_t swr = BN == 64 ? (uint32_t)(r & 7) : (BN == 32 ? (uint32_t)((r >> 1) & 3) : (uint32_t)((r >> 2) & 1));
      uint32_t wv;
      asm volatile("ld.shared.u32 %0, [%1];"
                   : "=r"(wv)
                   : "r"(obuf + (uint32_t)r * (uint32_t)(BN * 2) + ((((uint32_t)pair >> 2) ^ swr) << 4) +
                         ((uint32_t)pair & 3u) * 4u)
                   : "memory");
      E.se0 += __uint_as_float(wv << 16);
      E.se1 += __uint_as_float(wv & 0xffff0000u);
    }
  }
  fence_proxy_async();
  // plain output: the store of tile k - 1 (issued a whole tile ago) must be done reading staging tile 1 - b
  // before tile k + 1 writes it; the barrier below publishes that
  if (!twin && !has_res && E.leader) tma_store_wait_read();
  named_bar_sync(E.bar, 128);        // the tile is complete in shared memory
  if (E.leader) {
    tma_store_4d(E.tmY, obuf, cstore, ox0, oy0, n);
    if (twin) tma_store_4d(E.tmY2, tbuf, cstore, ox0, oy0, n);
    tma_store_commit();
  }
  ++E.k;
}

// fp32 variant for the split-bf16 parity mode: the item owns `nsum` accumulators of bn columns each; the tile is
// written as bn / 32 sub-tiles of [128 px][32 ch] fp32 (128-byte rows, 128B swizzle), one TMA store each, through
// ONE 16 KB staging buffer per warpgroup (the leader waits for the previous store's shared-memory reads).  The
// residual is read straight from global memory (a thread's 32 channels are one contiguous 128-byte run).
// Same float operations in the same order as epilogue_thread_scalar / conv_epilogue_store.
__device__ __forceinline__ void epilogue_warp_tma_f32(int bar, const ConvParams& P, const CUtensorMap* tmY, uint32_t obuf,
                                                      bool leader, int bn, int nsum, int n, int oy0, int ox0,
                                                      int cbase, uint32_t taddr, int wq, int lane, uint64_t* wait,
                                                      uint32_t parity, uint64_t* release, uint32_t bias_smem) {
  const bool has_res = P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE);
  const float slope = (P.flags & VCM_CONV_POST_ACT) ? P.post_slope : 1.0f;
  const int m = wq * 32 + lane;
  const int oy = oy0 + m / 8, ox = ox0 + m % 8;          // 8 x 16 tile (HTW = 8)
  const bool valid = oy < P.Ho && ox < P.Wo;
  const uint32_t row = (uint32_t)m * 128u;
  const uint32_t sw = (uint32_t)(m & 7);
  const size_t pix = ((size_t)n * P.Ho + (size_t)(valid ? oy : 0)) * (size_t)P.Wo + (size_t)(valid ? ox : 0);
  const float* rp = reinterpret_cast<const float*>(P.res) + pix * (size_t)P.res_cs + cbase;
  if (wait) {
    mbar_wait(wait, parity);
    tc_fence_after();
  }
  const int halves = bn / 32;
  for (int half = 0; half < halves; ++half) {
    float4 r[8];
    if (has_res && valid) {
#pragma unroll
      for (int j = 0; j < 8; ++j) r[j] = __ldg(reinterpret_cast<const float4*>(rp + 32 * half + 4 * j));
    }
    float v[32];
    tmem_load_block<32>(taddr + (uint32_t)(32 * half), v);
    for (int a = 1; a < nsum; ++a) {
      float u[32];
      tmem_load_block<32>(taddr + (uint32_t)(a * bn + 32 * half), u);
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = __fadd_rn(v[i], u[i]);
    }
    if (half == halves - 1 && release) {
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(release);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 b = lds128(bias_smem + (uint32_t)(cbase + 32 * half + 4 * j) * 4u);
      float w0 = leaky01(v[4 * j] + b.x, slope), w1 = leaky01(v[4 * j + 1] + b.y, slope);
      float w2 = leaky01(v[4 * j + 2] + b.z, slope), w3 = leaky01(v[4 * j + 3] + b.w, slope);
      if (has_res && valid) {
        if (P.flags & VCM_CONV_RES) {
          w0 = r[j].x + w0; w1 = r[j].y + w1; w2 = r[j].z + w2; w3 = r[j].w + w3;
        } else {
          const float4 sc = __ldg(reinterpret_cast<const float4*>(P.res_scale + (size_t)n * P.Cout + cbase + 32 * half + 4 * j));
          w0 = w0 + __fmul_rn(r[j].x, sc.x); w1 = w1 + __fmul_rn(r[j].y, sc.y);
          w2 = w2 + __fmul_rn(r[j].z, sc.z); w3 = w3 + __fmul_rn(r[j].w, sc.w);
        }
      }
      v[4 * j] = w0; v[4 * j + 1] = w1; v[4 * j + 2] = w2; v[4 * j + 3] = w3;
    }
    // the previous sub-tile's store must have finished reading the staging buffer
    if (leader) tma_store_wait_read();
    named_bar_sync(bar, 128);
#pragma unroll
    for (int j = 0; j < 8; ++j)
      sts128(obuf + row + (((uint32_t)j ^ sw) << 4), v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    fence_proxy_async();
    named_bar_sync(bar, 128);
    if (leader) {
      tma_store_4d(tmY, obuf, cbase + 32 * half, ox0, oy0, n);
      tma_store_commit();
    }
  }
}

// ------------------------------------------------------------------------------- scalar path
// thread = pixel (oy, ox); loops over the bn columns of the slab
// nsum > 1 (split-bf16 parity mode): the item owns nsum accumulators of bn columns each ([0] = correction
// terms, [1..] = partial sums of the main term); they are added smallest first with fp32 RN adds.
template <class T>
__device__ __forceinline__ void epilogue_thread_scalar(const ConvParams& P, int bn, int n, int oy, int ox,
                                                       int cbase, uint32_t taddr, uint64_t* wait,
                                                       uint32_t parity, uint64_t* release, int nsum = 1) {
  const bool valid = oy < P.Ho && ox < P.Wo;
  if (wait) {
    mbar_wait(wait, parity);
    tc_fence_after();
  }
  for (int cb = 0; cb < bn; cb += 16) {
    float v[16];
    tmem_load_block<16>(taddr + (uint32_t)cb, v);
    for (int a = 1; a < nsum; ++a) {
      float u[16];
      tmem_load_block<16>(taddr + (uint32_t)(a * bn + cb), u);
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = __fadd_rn(v[i], u[i]);
    }
    if (cb + 16 >= bn && release) {   // one arrival per warp
      tc_fence_before();
      __syncwarp();
      if ((threadIdx.x & 31) == 0) mbar_arrive(release);
    }
    if (!valid) continue;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int c = cbase + cb + i;
      if (c < P.Cout) {
        int co = c;   // position-major GEMM column -> reference channel 4*oc + pos
        if (P.flags & VCM_CONV_PIXEL_SHUFFLE) {
          const int q4 = P.Cout >> 2;
          co = 4 * (c % q4) + c / q4;
        }
        conv_epilogue_store<T>(P, n, oy, ox, co, v[i], __ldg(P.bias + c));
      }
    }
  }
}

}  // namespace tc

// vcm_ts_b200/csrc/conv_halo.cu -- stride-1 tensor-core convolution for sm_100a: implicit GEMM on
// tcgen05.mma with a HALO activation tile and (where they fit) shared-memory-RESIDENT weights.
//
// Why.  The per-tap engine (conv_tc.cu) re-fetches the 128-pixel activation tile and the weight slab
// for each of the K*K taps: 216 KB of L2->SM traffic per 128x64 output tile of a 3x3 64->64 conv,
// which ncu shows is what bounds it (profiles/ncu_conv64_pertap_r01.json: 4.1 GB through the
// crossbar per launch, tensor pipe 12 %).  Here
//   * ONE 4-D TMA box {CHUNK, 8+K-1, 16*MH+K-1, 1} brings the tile plus its halo (zero-filled outside
//     the image = the conv padding); every tap is then just a different START ADDRESS of the same
//     shared-memory tile: the UMMA descriptor addresses row m = 8*y + x of the A operand at
//     start + y*SBO + x*rowbytes with SBO = (8+K-1)*rowbytes, start = ((ky*(8+K-1) + kx)*rowbytes.
//     This relies on the 128/64/32-byte swizzle being a function of the absolute shared-memory
//     address (measured: vcm_umma_probe, profiles/umma_probe_r01.json), so TMA-written tiles can be
//     read at any row offset / row-group stride;
//   * weights whose taps*Cin*Cout bf16 fit beside the A ring (3x3 64->64: 72 KB) are loaded ONCE per
//     persistent CTA; otherwise weight slabs stream through their own ring, and the item grows to
//     M = 256 (two 8x16 halves, MH = 2) so each slab is used for twice the pixels.
// Traffic per 3x3 64->64 tile drops from 216 KB to 23 KB.
//
// Roles (384 threads = 3 warpgroups): warpgroup 0 = warp 0 activation TMA, warp 1 MMA issue + TMEM,
// warp 2 weight TMA, warp 3 idle (setmaxnreg.dec); warpgroups 1 / 2 (warps 4-7 / 8-11) = epilogue, one per
// TMEM accumulator buffer (setmaxnreg.inc: no spills -- with 227 KB of shared memory carved out L1 is
// tiny and every local-memory access stalls for an L2 round trip), see tc_epilogue.cuh.
#include <mutex>

#include "conv_common.cuh"
#include "tc_epilogue.cuh"
#include "tc_host.cuh"
using namespace tc;

namespace {

constexpr int HTW = 8, HTH = 16;
constexpr int H_THREADS = 384;
constexpr int H_MAX_A = 6, H_MAX_B = 12, H_MAX_ACC = 4;

struct HaloGeom {
  // split-bf16 parity mode (VCM_CONV_SPLIT3): split = 1, an activation stage holds the THREE bf16 planes
  // of one Cin chunk (a_plane_bytes apart), weight slabs are [part][tap][chunk], and an item owns
  // nsacc TMEM accumulators of bn columns: [0] = the correction terms (x_i*w_j, i+j > 0), [1..] = the
  // main term x0*w0 dealt round-robin over (chunk, tap) -- summed in fp32 by the epilogue
  int split, pstride, a_plane_bytes, nsacc;
  int nchunks, ntaps, a_stages, b_stages, resident;
  int a_bytes, a_stage_bytes;
  int b_bytes, b_slot_bytes, b_base;
  int hw, hh, mh;
  int tiles_x, tiles_y, n_tiles, num_items;
  int bn, cout_p, tmem_cols, scalar_epi;
  int nacc;   // TMEM accumulator ring depth (2 or 4): lets the MMA warp run ahead of both epilogue warpgroups
  int epi_base, bar_base, bias_base, epi_bw;
  int scale_off;   // byte offset (from bias_base) of the shared-memory copy of res_scale, 0 = none
  int tma_epi;   // bf16 tiles leave through TMA stores (tc_epilogue.cuh epilogue_warp_tma)
  int sw, nsub;  // tma_epi == 1: an N tile of bn columns leaves as nsub sub-tiles of sw (64 / 32) channels
  int ps_q4;     // PixelShuffle through TMA: Cout / 4 (sub-tiles go to the strided positions of one sub-pixel phase), else 0
  uint32_t idesc;
};

struct HaloBars {
  uint64_t *afull, *aempty, *bfull, *bempty, *tfull, *tempty;
};

// MMA-issuing warp.  The profile of a runtime-parametrised tap loop showed this warp, not the tensor
// pipe, pacing the kernel (~540 cycles per tap for 4 MMAs): so the tap loops are unrolled for the
// filter size, every operand descriptor is `base + compile-time constant` (tap (ky,kx) of the halo
// tile starts (ky*hw + kx) pixel rows further; the address field counts 16-byte units), and a whole
// (item, Cin chunk) batch of K*K*KSTEPS*MH MMAs is issued by the elected lane without leaving the
// single-thread region.  The other lanes only take part in the barrier waits.
template <int CHUNK, int K, int MH, bool RESIDENT>
__device__ __forceinline__ void mma_role(const HaloGeom& G, const HaloBars& B, uint32_t a_smem, uint32_t b_smem,
                                         uint32_t tmem_base) {
  constexpr uint32_t LAYOUT = CHUNK == 64 ? 2u : (CHUNK == 32 ? 4u : 6u);
  constexpr uint32_t ROWB = CHUNK * 2u;
  constexpr int KSTEPS = CHUNK / 16;
  constexpr uint32_t HW = HTW + K - 1;
  constexpr uint32_t AROW16 = ROWB >> 4;             // one pixel to the right
  constexpr uint32_t ALINE16 = (HW * ROWB) >> 4;     // one halo row down
  constexpr uint32_t AHALF16 = HTH * ALINE16;        // second 128-pixel half (MH == 2)
  int as = 0, bs = 0, acc = 0;   // bs / bphase live in the elected lane only (always the same lane)
  uint32_t aphase = 0, bphase = 0, acc_phase = 0;
  bool first = true;   // resident weights: wait for each tap's slabs while the first tile is issued
  const uint64_t bdesc0 = make_smem_desc(b_smem, 8u * ROWB, LAYOUT);
  const uint32_t bslot16 = (uint32_t)G.b_slot_bytes >> 4;
  const uint32_t btap16 = bslot16 * (uint32_t)G.nchunks;   // resident: slab index = tap * nchunks + ch
  const uint32_t idesc = G.idesc;
  const uint32_t bn = (uint32_t)G.bn;
  for (int item = blockIdx.x; item < G.num_items; item += gridDim.x) {
    mbar_wait(&B.tempty[acc], acc_phase ^ 1u);
    tc_fence_after();
    const uint32_t d_tmem = tmem_base + (uint32_t)acc * (uint32_t)MH * bn;
    for (int ch = 0; ch < G.nchunks; ++ch) {
      mbar_wait(&B.afull[as], aphase);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t adesc0 = make_smem_desc(a_smem + (uint32_t)as * (uint32_t)G.a_stage_bytes, HW * ROWB, LAYOUT);
        const uint64_t bch = bdesc0 + (uint64_t)((uint32_t)ch * bslot16);
        // 7x7 with 32-channel chunks: the filter-row loop stays a loop (measured 465 -> 451 us for 32->64 at
        // 1088x1920; with 64-channel chunks the unrolled form was the faster one, 513 vs 535 us for 64->32)
#pragma unroll(K <= 3 || CHUNK == 64 ? K : 1)
        for (int ky = 0; ky < K; ++ky) {
#pragma unroll
          for (int kx = 0; kx < K; ++kx) {
            const int tap = ky * K + kx;
            uint64_t bdesc;
            if (RESIDENT) {
              if (first) {
                mbar_wait(&B.bfull[tap < H_MAX_B - 1 ? tap : H_MAX_B - 1], 0);
                tc_fence_after();
              }
              bdesc = bch + (uint64_t)((uint32_t)tap * btap16);
            } else {
              mbar_wait(&B.bfull[bs], bphase);
              tc_fence_after();
              bdesc = bdesc0 + (uint64_t)((uint32_t)bs * bslot16);
            }
            const uint64_t adesc = adesc0 + (uint64_t)(ky * ALINE16 + kx * AROW16);
#pragma unroll
            for (int k = 0; k < KSTEPS; ++k)
              umma_bf16(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc,
                        (tap > 0 || k > 0) ? 1u : (ch > 0 ? 1u : 0u));
            if (MH == 2) {
#pragma unroll
              for (int k = 0; k < KSTEPS; ++k)
                umma_bf16(d_tmem + bn, adesc + (uint64_t)(AHALF16 + 2 * k), bdesc + (uint64_t)(2 * k), idesc,
                          (tap > 0 || k > 0) ? 1u : (ch > 0 ? 1u : 0u));
            }
            if (!RESIDENT) {
              umma_commit(&B.bempty[bs]);
              if (++bs == G.b_stages) { bs = 0; bphase ^= 1u; }
            }
          }
        }
        umma_commit(&B.aempty[as]);
        if (ch == G.nchunks - 1) umma_commit(&B.tfull[acc]);
      }
      first = false;
      __syncwarp();
      if (++as == G.a_stages) { as = 0; aphase ^= 1u; }
    }
    if (++acc == G.nacc) { acc = 0; acc_phase ^= 1u; }
  }
}


// Split-bf16 (parity) MMA role.  The tensor core's fp32 accumulate truncates (measured: the error of a long
// accumulation chain grows linearly, profiles/umma_acc_probe_r02.json), so
//   * the main term x0*w0 -- whose bf16 x bf16 products are exact in fp32 -- is spread over several
//     accumulators (short chains), and
//   * the five correction terms (2^-8 / 2^-16 of the main term) get their own accumulator, so that they
//     are not truncated at the main sum's ulp.
// The epilogue adds the accumulators with round-to-nearest fp32 adds.  One stage = three planes of one Cin
// chunk; for every weight slab (part j, tap) the planes i <= 2 - j are multiplied: 6 products, 3 slabs.
template <int CHUNK, int K, bool RESIDENT>
__device__ __forceinline__ void mma_role_split(const HaloGeom& G, const HaloBars& B, uint32_t a_smem,
                                               uint32_t b_smem, uint32_t tmem_base) {
  constexpr uint32_t LAYOUT = CHUNK == 64 ? 2u : (CHUNK == 32 ? 4u : 6u);
  constexpr uint32_t ROWB = CHUNK * 2u;
  constexpr int KSTEPS = CHUNK / 16;
  constexpr uint32_t HW = HTW + K - 1;
  constexpr uint32_t AROW16 = ROWB >> 4;
  constexpr uint32_t ALINE16 = (HW * ROWB) >> 4;
  int as = 0, bs = 0, acc = 0;
  uint32_t aphase = 0, bphase = 0, acc_phase = 0;
  bool first = true;
  const uint64_t bdesc0 = make_smem_desc(b_smem, 8u * ROWB, LAYOUT);
  const uint32_t bslot16 = (uint32_t)G.b_slot_bytes >> 4;
  const uint32_t btap16 = bslot16 * (uint32_t)G.nchunks;
  const uint32_t bpart16 = btap16 * (uint32_t)G.ntaps;
  const uint32_t aplane16 = (uint32_t)G.a_plane_bytes >> 4;
  const uint32_t idesc = G.idesc;
  const uint32_t bn = (uint32_t)G.bn;
  const int nmain = G.nsacc - 1;
  for (int item = blockIdx.x; item < G.num_items; item += gridDim.x) {
    mbar_wait(&B.tempty[acc], acc_phase ^ 1u);
    tc_fence_after();
    const uint32_t d_tmem = tmem_base + (uint32_t)acc * (uint32_t)G.nsacc * bn;
    uint32_t started = 0;   // accumulators already written for this item (elected lane only)
    int mctr = 0;           // main-term slab counter -> accumulator 1 + mctr % nmain
    for (int ch = 0; ch < G.nchunks; ++ch) {
      mbar_wait(&B.afull[as], aphase);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t adesc0 = make_smem_desc(a_smem + (uint32_t)as * (uint32_t)G.a_stage_bytes, HW * ROWB, LAYOUT);
        for (int pj = 2; pj >= 0; --pj) {
          const uint64_t bch = bdesc0 + (uint64_t)((uint32_t)ch * bslot16 + (uint32_t)pj * bpart16);
#pragma unroll(K <= 3 ? K : 1)
          for (int ky = 0; ky < K; ++ky) {
#pragma unroll
            for (int kx = 0; kx < K; ++kx) {
              const int tap = ky * K + kx;
              uint64_t bdesc;
              if (RESIDENT) {
                if (first) {
                  mbar_wait(&B.bfull[tap < H_MAX_B - 1 ? tap : H_MAX_B - 1], 0);
                  tc_fence_after();
                }
                bdesc = bch + (uint64_t)((uint32_t)tap * btap16);
              } else {
                mbar_wait(&B.bfull[bs], bphase);
                tc_fence_after();
                bdesc = bdesc0 + (uint64_t)((uint32_t)bs * bslot16);
              }
              const uint64_t atap = adesc0 + (uint64_t)(ky * ALINE16 + kx * AROW16);
              for (int pi = 0; pi <= 2 - pj; ++pi) {
                uint32_t sel = 0;
                if (pi == 0 && pj == 0) {
                  sel = 1u + (uint32_t)mctr;
                  if (++mctr == nmain) mctr = 0;
                }
                const uint32_t d = d_tmem + sel * bn;
                const uint64_t adesc = atap + (uint64_t)((uint32_t)pi * aplane16);
#pragma unroll
                for (int k = 0; k < KSTEPS; ++k)
                  umma_bf16(d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc,
                            k > 0 ? 1u : ((started >> sel) & 1u));
                started |= 1u << sel;
              }
              if (!RESIDENT) {
                umma_commit(&B.bempty[bs]);
                if (++bs == G.b_stages) { bs = 0; bphase ^= 1u; }
              }
            }
          }
        }
        umma_commit(&B.aempty[as]);
        if (ch == G.nchunks - 1) umma_commit(&B.tfull[acc]);
      }
      first = false;
      __syncwarp();
      if (++as == G.a_stages) { as = 0; aphase ^= 1u; }
    }
    if (++acc == G.nacc) { acc = 0; acc_phase ^= 1u; }
  }
}

template <int CHUNK>
__global__ void __launch_bounds__(H_THREADS, 1)
conv_halo_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmY, const __grid_constant__ CUtensorMap tmY2,
                 const __grid_constant__ CUtensorMap tmR, const ConvParams P, const HaloGeom G) {
  constexpr uint32_t LAYOUT = CHUNK == 64 ? 2u : (CHUNK == 32 ? 4u : 6u);
  constexpr uint32_t ROWB = CHUNK * 2u;      // bytes per pixel row of the A tile / per weight row
  constexpr int KSTEPS = CHUNK / 16;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + G.bar_base);
  uint64_t* afull = bars;
  uint64_t* aempty = afull + H_MAX_A;
  uint64_t* bfull = aempty + H_MAX_A;
  uint64_t* bempty = bfull + H_MAX_B;
  uint64_t* tfull = bempty + H_MAX_B;
  uint64_t* tempty = tfull + H_MAX_ACC;
  uint64_t* rfull = tempty + H_MAX_ACC;      // [2][2]: residual tile b of epilogue warpgroup g has landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(rfull + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int sp_tiles = G.tiles_x * G.tiles_y;
  const uint32_t sbo = (uint32_t)G.hw * ROWB;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int s = 0; s < H_MAX_A; ++s) { mbar_init(&afull[s], 1); mbar_init(&aempty[s], 1); }
    for (int s = 0; s < H_MAX_B; ++s) { mbar_init(&bfull[s], 1); mbar_init(&bempty[s], 1); }
    for (int a = 0; a < H_MAX_ACC; ++a) { mbar_init(&tfull[a], 1); mbar_init(&tempty[a], 4); }
    for (int r = 0; r < 4; ++r) mbar_init(&rfull[r], 1);
    if (G.tma_epi) {
      prefetch_tmap(&tmY);
      if (G.tma_epi == 1 && (P.flags & VCM_CONV_OUT2)) prefetch_tmap(&tmY2);
      if (G.tma_epi == 1 && (P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE))) prefetch_tmap(&tmR);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, (uint32_t)G.tmem_cols);
  {   // bias vector -> shared memory (the vector epilogue reads it per channel block)
    float* bs_ = reinterpret_cast<float*>(smem + G.bias_base);
    const int nb = G.cout_p < EPI_BIAS_MAX ? G.cout_p : EPI_BIAS_MAX;
    for (int i = threadIdx.x; i < nb; i += blockDim.x) bs_[i] = i < P.Cout ? __ldg(P.bias + i) : 0.f;
    if (G.scale_off) {   // SE scales of the residual (not under PDL: they are the predecessor's output)
      float* sc_ = reinterpret_cast<float*>(smem + G.bias_base + G.scale_off);
      for (int i = threadIdx.x; i < P.N * P.Cout; i += blockDim.x) sc_[i] = __ldg(P.res_scale + i);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t bias_smem = smem_u32(smem + G.bias_base);
  pdl_trigger();   // the next kernel of the stream may start its prologue on SMs this grid has left

  if (warp < 4) {
  reg_dec<72>();
  if (warp == 0) {
    // ============================ activation producer: one halo box per (item, chunk) ==========
    if (lane == 0) {
      pdl_wait();   // activations are the previous kernel's output
      int stage = 0;
      uint32_t phase = 0;
      for (int item = blockIdx.x; item < G.num_items; item += gridDim.x) {
        const int sp = item / G.n_tiles;
        const int n = sp / sp_tiles, rem = sp - n * sp_tiles;
        const int y0 = (rem / G.tiles_x) * (HTH * G.mh), x0 = (rem % G.tiles_x) * HTW;
        for (int ch = 0; ch < G.nchunks; ++ch) {
          mbar_wait(&aempty[stage], phase ^ 1u);
          const uint32_t dst = smem_u32(smem + (size_t)stage * G.a_stage_bytes);
          if (G.split) {
            mbar_expect_tx(&afull[stage], 3u * (uint32_t)G.a_bytes);
            for (int p = 0; p < 3; ++p)
              tma_load_4d(dst + (uint32_t)(p * G.a_plane_bytes), &tmA, &afull[stage], p * G.pstride + ch * CHUNK,
                          x0 - P.pad, y0 - P.pad, n);
          } else {
            mbar_expect_tx(&afull[stage], (uint32_t)G.a_bytes);
            tma_load_4d(dst, &tmA, &afull[stage], ch * CHUNK, x0 - P.pad, y0 - P.pad, n);
          }
          if (++stage == G.a_stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 2) {
    // ============================ weight producer ==============================================
    if (lane == 0) {
      if (!(P.flags & VCM_CONV_PDL)) pdl_wait();   // weights possibly packed by the kernel just before this one
      const int nparts = G.split ? 3 : 1;
      if (G.resident) {
        // one barrier per tap (the taps past H_MAX_B - 1 share the last one), in the order the MMA role
        // consumes them: the first tile starts after 1/ntaps of the weights has landed, not all of them
        for (int tap = 0; tap < G.ntaps; ++tap) {
          const int bi = tap < H_MAX_B - 1 ? tap : H_MAX_B - 1;
          if (tap <= H_MAX_B - 1)
            mbar_expect_tx(&bfull[bi], (uint32_t)((tap < H_MAX_B - 1 ? 1 : G.ntaps - (H_MAX_B - 1)) * G.nchunks *
                                                  nparts * G.b_bytes));
          for (int part = 0; part < nparts; ++part)
            for (int ch = 0; ch < G.nchunks; ++ch) {
              const int s = (part * G.ntaps + tap) * G.nchunks + ch;   // slab layout [part][tap][chunk]
              tma_load_2d(smem_u32(smem + G.b_base + (size_t)s * G.b_slot_bytes), &tmB, &bfull[bi], 0, s * G.cout_p);
            }
        }
      } else {
        int stage = 0;
        uint32_t phase = 0;
        for (int item = blockIdx.x; item < G.num_items; item += gridDim.x) {
          const int nt = item % G.n_tiles;
          for (int ch = 0; ch < G.nchunks; ++ch)
            for (int part = nparts - 1; part >= 0; --part)     // the order the MMA role consumes them
              for (int tap = 0; tap < G.ntaps; ++tap) {
                mbar_wait(&bempty[stage], phase ^ 1u);
                mbar_expect_tx(&bfull[stage], (uint32_t)G.b_bytes);
                tma_load_2d(smem_u32(smem + G.b_base + (size_t)stage * G.b_slot_bytes), &tmB, &bfull[stage], 0,
                            ((part * G.ntaps + tap) * G.nchunks + ch) * G.cout_p + nt * G.bn);
                if (++stage == G.b_stages) { stage = 0; phase ^= 1u; }
              }
        }
      }
    }
  } else if (warp == 1) {
    // ============================ MMA issue ====================================================
    HaloBars B{afull, aempty, bfull, bempty, tfull, tempty};
    const uint32_t a_smem = smem_u32(smem), b_smem = smem_u32(smem + G.b_base);
#define HALO_MMA(KK)                                                                              \
  do {                                                                                            \
    if (G.split && G.resident) mma_role_split<CHUNK, KK, true>(G, B, a_smem, b_smem, tmem_base);   \
    else if (G.split) mma_role_split<CHUNK, KK, false>(G, B, a_smem, b_smem, tmem_base);           \
    else if (G.resident) mma_role<CHUNK, KK, 1, true>(G, B, a_smem, b_smem, tmem_base);            \
    else if (G.mh == 1) mma_role<CHUNK, KK, 1, false>(G, B, a_smem, b_smem, tmem_base);            \
    else mma_role<CHUNK, KK, 2, false>(G, B, a_smem, b_smem, tmem_base);                           \
  } while (0)
    if (P.K == 3) HALO_MMA(3);
    else if (P.K == 1) HALO_MMA(1);
    else HALO_MMA(7);
#undef HALO_MMA
  }
  } else {
    reg_inc<216>();
    // ============ epilogue: warpgroup g (warps 4-7 / 8-11) owns accumulator buffer g ===========
    const int g = (warp - 4) >> 2;
    const int q = warp & 3;   // TMEM lane quarter this warp may access
    const uint32_t stage = smem_u32(smem + G.epi_base) + (uint32_t)((warp - 4) * EPI_STAGE_FLOATS * 4);
    pdl_wait();   // residuals / SE scales are read and outputs written only after the predecessor finished
    if (G.tma_epi == 2) {
      // split-bf16 parity mode (mh == 1): fp32 tiles through TMA stores, one staging buffer per warpgroup
      const uint32_t obuf = smem_u32(smem + G.epi_base) + (uint32_t)(g * EPI_TMA_TILE_BYTES);
      const bool leader = (q == 0 && lane == 0);
      int local = 0;
      for (int item = blockIdx.x; item < G.num_items; item += gridDim.x, ++local) {
        if ((local & 1) != g) continue;
        const int ab = local % G.nacc;
        const uint32_t acc_phase = (uint32_t)(local / G.nacc) & 1u;
        const int nt = item % G.n_tiles, sp = item / G.n_tiles;
        const int n = sp / sp_tiles, rem = sp - n * sp_tiles;
        const int ox0 = (rem % G.tiles_x) * HTW, oy0 = (rem / G.tiles_x) * HTH;
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(ab * G.nsacc * G.bn);
        epilogue_warp_tma_f32(1 + g, P, &tmY, obuf, leader, G.bn, G.nsacc, n, oy0, ox0, nt * G.bn, taddr, q, lane,
                              &tfull[ab], acc_phase, &tempty[ab], bias_smem);
      }
      if (leader) tma_store_wait_all();
    } else if (G.tma_epi) {
      // this warpgroup's k-th tile: item local = 2 * (k / mh) + g, half h = k % mh
      EpiTma E;
      E.tmY = &tmY; E.tmY2 = &tmY2; E.tmR = &tmR;
      E.obuf = smem_u32(smem + G.epi_base) + (uint32_t)(g * 2 * EPI_TMA_TILE_BYTES);
      E.tile_bytes = (uint32_t)EPI_TMA_TILE_BYTES;
      E.scale_smem = G.scale_off ? smem_u32(smem + G.bias_base + G.scale_off) : 0u;
      E.rfull = &rfull[2 * g];
      E.k = 0;
      E.se_sum = (P.flags & VCM_CONV_SE_SUM) != 0;
      E.se0 = E.se1 = 0.f;
      E.leader = (q == 0 && lane == 0);
      E.bar = 1 + g;
      const bool has_res = P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE);
      // this warpgroup's k-th tile: item local = 2 * (k / (mh * nsub)) + g, then (half h, column sub-tile)
      const int per_item = G.mh * G.nsub;
      // -> tensor-map coordinates (sx, sy, cs) of the tile, its first GEMM column cb and TMEM column offset tc
      auto tile = [&](int k, int& n, int& sy, int& sx, int& cs, int& cb, int& tc, bool& first, bool& last) -> bool {
        const int loc = 2 * (k / per_item) + g;
        const int rem = k % per_item;
        const int h = rem / G.nsub, sub = rem - h * G.nsub;
        const int item = blockIdx.x + loc * gridDim.x;
        if (item >= G.num_items) return false;
        const int nt = item % G.n_tiles, sp = item / G.n_tiles;
        n = sp / sp_tiles;
        const int r2 = sp - n * sp_tiles;
        const int ox0 = (r2 % G.tiles_x) * HTW;
        const int oy0 = (r2 / G.tiles_x) * (HTH * G.mh) + h * HTH;
        cb = nt * G.bn + sub * G.sw;
        tc = h * G.bn + sub * G.sw;
        if (G.ps_q4) {     // GEMM column = phase * q4 + channel: phase (dy, dx) of the 2x upsampled image
          const int pos = cb / G.ps_q4;
          cs = cb - pos * G.ps_q4;
          sx = 2 * ox0 + (pos & 1);
          sy = 2 * oy0 + (pos >> 1);
        } else {
          cs = cb; sx = ox0; sy = oy0;
        }
        first = rem == 0;
        last = rem == per_item - 1;
        return true;
      };
      int n, sy, sx, cs, cb, tc;
      bool first, last;
      if (has_res && E.leader && tile(0, n, sy, sx, cs, cb, tc, first, last)) {
        if (G.sw == 64) epi_tma_prefetch_res<64>(E, 0, cs, sx, sy, n);
        else epi_tma_prefetch_res<32>(E, 0, cs, sx, sy, n);
      }
      for (int k = 0; tile(k, n, sy, sx, cs, cb, tc, first, last); ++k) {
        int n2, sy2, sx2, cs2, cb2, tc2;
        bool f2, l2;
        const bool ok2 = has_res && tile(k + 1, n2, sy2, sx2, cs2, cb2, tc2, f2, l2);   // whose residual is requested next
        const int loc = 2 * (k / per_item) + g;
        const int ab = loc % G.nacc;
        const uint32_t acc_phase = (uint32_t)(loc / G.nacc) & 1u;
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(ab * G.mh * G.bn + tc);
#define EPI_TMA(BN)                                                                                             \
  epilogue_warp_tma<BN>(P, E, n, sy, sx, cs, cb, taddr, q, lane, first ? &tfull[ab] : nullptr, acc_phase,         \
                        last ? &tempty[ab] : nullptr, bias_smem, ok2, n2, sy2, sx2, cs2)
        if (G.sw == 64) EPI_TMA(64);
        else EPI_TMA(32);
#undef EPI_TMA
      }
      if (E.se_sum) {
        if (G.sw == 64) epi_tma_write_se_sums<64>(E, reinterpret_cast<float*>(P.y2), warp - 4, lane);
        else epi_tma_write_se_sums<32>(E, reinterpret_cast<float*>(P.y2), warp - 4, lane);
      }
      if (E.leader) tma_store_wait_all();
    } else {
    int local = 0;
    for (int item = blockIdx.x; item < G.num_items; item += gridDim.x, ++local) {
      if ((local & 1) != g) continue;
      const int ab = local % G.nacc;                          // accumulator buffer (its parity is g)
      const uint32_t acc_phase = (uint32_t)(local / G.nacc) & 1u;
      const int nt = item % G.n_tiles, sp = item / G.n_tiles;
      const int n = sp / sp_tiles, rem = sp - n * sp_tiles;
      const int ox0 = (rem % G.tiles_x) * HTW;
      const int cbase = nt * G.bn;
      for (int h = 0; h < G.mh; ++h) {
        const int oy0 = (rem / G.tiles_x) * (HTH * G.mh) + h * HTH;
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) +
                               (uint32_t)(G.split ? ab * G.nsacc * G.bn : (ab * G.mh + h) * G.bn);
        uint64_t* wait = h == 0 ? &tfull[ab] : nullptr;
        uint64_t* release = h == G.mh - 1 ? &tempty[ab] : nullptr;
        if (G.split) {
          const int m = q * 32 + lane;
          epilogue_thread_scalar<float>(P, G.bn, n, oy0 + m / HTW, ox0 + m % HTW, cbase, taddr, wait, acc_phase,
                                        release, G.nsacc);
        } else if (G.scalar_epi == 0) {
          if (G.epi_bw == 32)
            epilogue_warp_vec<32, HTW>(P, G.bn, n, oy0, ox0, cbase, taddr, stage, q, lane, wait, acc_phase, release, bias_smem);
          else
            epilogue_warp_vec<16, HTW>(P, G.bn, n, oy0, ox0, cbase, taddr, stage, q, lane, wait, acc_phase, release, bias_smem);
        } else {
          const int m = q * 32 + lane;
          const int oy = oy0 + m / HTW, ox = ox0 + m % HTW;
          if (G.scalar_epi == 2)
            epilogue_thread_scalar<float>(P, G.bn, n, oy, ox, cbase, taddr, wait, acc_phase, release);
          else
            epilogue_thread_scalar<bf>(P, G.bn, n, oy, ox, cbase, taddr, wait, acc_phase, release);
        }
      }
    }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)G.tmem_cols);
  }
}

}  // namespace

// ---------------------------------------------------------------------------------- host
int vcm_conv_halo_launch(const vcm_conv_desc* d, const ConvParams& P, const void* x, const void* w_umma,
                         cudaStream_t stream, int* se_rows) {
  if (P.stride != 1) return 3;
  {   // narrow 3x3 layers on large images: output rows stacked along N (conv_rows.cu)
    const int rc = vcm_conv_rows_launch(d, P, x, w_umma, stream, se_rows);
    if (rc != 3) return rc;
  }
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    vcm_set_error("vcm_conv_halo_launch: cuTensorMapEncodeTiled unavailable");
    return 1;
  }
  HaloGeom G;
  memset(&G, 0, sizeof(G));
  G.split = (P.flags & VCM_CONV_SPLIT3) ? 1 : 0;
  G.pstride = d->in_pstride;
  const int chunk = G.split ? pick_chunk_split(P.Cin, P.K) : pick_chunk(P.Cin);
  pick_n_tiling(P.Cout, G.cout_p, G.bn, G.n_tiles, G.split ? 128 : 256);
  // Small images (the 1/16-scale prior / hyper nets: 75 spatial tiles at 1080p) cannot fill 148 SMs with
  // one N tile: cut Cout into more, narrower N tiles (the activation tile is re-read from L2 instead).
  const int sp1 = cdiv(P.Wo, HTW) * cdiv(P.Ho, HTH) * P.N;
  if (sp1 * G.n_tiles < num_sms()) {
    for (int nt = G.n_tiles + 1; nt <= G.cout_p / 32; ++nt) {
      if (G.cout_p % nt || (G.cout_p / nt) % 16) continue;
      G.n_tiles = nt;
      G.bn = G.cout_p / nt;
      if (sp1 * nt >= num_sms()) break;
    }
  }
  G.nchunks = P.Cin / chunk;
  G.ntaps = P.K * P.K;
  G.hw = HTW + P.K - 1;
  G.b_bytes = G.bn * chunk * 2;
  G.b_slot_bytes = round_up(G.b_bytes, 1024);
  const int slabs = G.ntaps * G.nchunks * (G.split ? 3 : 1);
  int bias_bytes = round_up((G.cout_p < EPI_BIAS_MAX ? G.cout_p : EPI_BIAS_MAX) * 4, 128);
  if ((P.flags & VCM_CONV_RES_SCALE) && !(P.flags & VCM_CONV_PDL) && P.N * P.Cout <= 2048) {
    G.scale_off = bias_bytes;
    bias_bytes += round_up(P.N * P.Cout * 4, 128);
  }
  G.scalar_epi = tc_epilogue_mode(d, P);
  static const bool no_tma_epi = getenv("VCM_TC_NO_TMA_EPI") && getenv("VCM_TC_NO_TMA_EPI")[0] == '1';
  // TMA-store epilogue: the N tile leaves as sub-tiles of 64 (or 32) channels.  PixelShuffle layers qualify when
  // a sub-pixel phase (Cout / 4 GEMM columns) is a whole number of sub-tiles: a phase is then an ordinary tile
  // of a tensor map that walks the 2x upsampled output with element strides 2.
  G.sw = G.bn % 64 == 0 ? 64 : 32;
  const bool ps = P.flags & VCM_CONV_PIXEL_SHUFFLE;
  if (ps && (P.Cout / 4) % G.sw) G.sw = 32;
  G.nsub = G.bn / G.sw;
  static const bool no_wide_tma = getenv("VCM_TC_NO_WIDE_TMA_EPI") && getenv("VCM_TC_NO_WIDE_TMA_EPI")[0] == '1';
  G.tma_epi = (!no_tma_epi && !G.split && G.scalar_epi == 0 && G.bn % G.sw == 0 && P.Cout % G.sw == 0 &&
               G.cout_p == P.Cout && (!ps || ((P.Cout / 4) % G.sw == 0 && P.Cout % 4 == 0 && !(P.flags & VCM_CONV_RES_SCALE))) &&
               (!no_wide_tma || (!ps && G.nsub == 1)) &&
               !(P.flags & VCM_CONV_RES2) &&
               !((P.flags & VCM_CONV_OUT2) && (P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE)))) ? 1 : 0;
  G.ps_q4 = (G.tma_epi && ps) ? P.Cout / 4 : 0;
  if (!no_tma_epi && G.split && d->out_dtype == VCM_F32 && P.Cout % 32 == 0 && G.bn % 32 == 0 &&
      !(P.flags & (VCM_CONV_PIXEL_SHUFFLE | VCM_CONV_RES2 | VCM_CONV_OUT2)) && P.out_cs % 4 == 0 &&
      ((uintptr_t)P.y & 15) == 0 &&
      (!(P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE)) || (P.res_cs % 4 == 0 && ((uintptr_t)P.res & 15) == 0)) &&
      (!(P.flags & VCM_CONV_RES_SCALE) || ((uintptr_t)P.res_scale & 15) == 0))
    G.tma_epi = 2;   // fp32 sub-tiles of 32 channels (epilogue_warp_tma_f32)
  auto a_stage = [&](int mh) { return (G.split ? 3 : 1) * round_up(G.hw * (HTH * mh + P.K - 1) * chunk * 2, 1024); };
  const int epi_bytes = G.tma_epi == 1 ? EPI_TMA_BYTES : EPI_STAGE_BYTES;
  const int budget = SMEM_LIMIT - 1024 - 512 - bias_bytes - epi_bytes;   // alignment slack, barriers, bias, staging

  // resident weights if they leave room for >= 3 activation stages
  G.resident = (G.n_tiles == 1) && (slabs * G.b_slot_bytes + (G.split ? 2 : 3) * a_stage(1) <= budget);
  if (G.resident) {
    G.mh = 1;
    G.b_stages = 0;
    G.a_stage_bytes = a_stage(1);
    G.a_stages = (budget - slabs * G.b_slot_bytes) / G.a_stage_bytes;
  } else {
    // 256-pixel items halve the weight traffic per pixel, but only when they still give every SM an item
    G.mh = (!G.split && G.bn <= 128 && P.Ho > HTH &&
            cdiv(P.Wo, HTW) * cdiv(P.Ho, 2 * HTH) * P.N * G.n_tiles >= num_sms()) ? 2 : 1;
    G.a_stage_bytes = a_stage(G.mh);
    G.a_stages = G.split ? 2 : 3;
    if (G.a_stages * G.a_stage_bytes > budget / 2) G.a_stages = 2;
    if (G.a_stages * G.a_stage_bytes + 2 * G.b_slot_bytes > budget) return 3;
    G.b_stages = (budget - G.a_stages * G.a_stage_bytes) / G.b_slot_bytes;
    if (G.b_stages > H_MAX_B) G.b_stages = H_MAX_B;
    // spend what is left on more activation stages
    const int left = budget - G.b_stages * G.b_slot_bytes - G.a_stages * G.a_stage_bytes;
    G.a_stages += left / G.a_stage_bytes;
  }
  if (G.a_stages > H_MAX_A) G.a_stages = H_MAX_A;
  if (G.a_stages < 2) return 3;
  G.hh = HTH * G.mh + P.K - 1;
  G.a_bytes = G.hw * G.hh * chunk * 2;
  G.b_base = G.a_stages * G.a_stage_bytes;
  G.tiles_x = cdiv(P.Wo, HTW);
  G.tiles_y = cdiv(P.Ho, HTH * G.mh);
  G.num_items = G.tiles_x * G.tiles_y * P.N * G.n_tiles;
  G.nacc = 4 * G.mh * G.bn <= 512 ? 4 : 2;
  G.a_plane_bytes = G.a_stage_bytes / 3;
  if (G.split) {
    // accumulators per item: as many as two items' worth fit the 512 TMEM columns (bn <= 128 -> >= 2),
    // at most 8, and never more main accumulators than there are main-term slabs
    G.nsacc = 512 / (2 * G.bn);
    if (G.nsacc > 8) G.nsacc = 8;
    if (G.nsacc - 1 > G.ntaps * G.nchunks) G.nsacc = 1 + G.ntaps * G.nchunks;
    G.nacc = 2;
  }
  int cols = 32;
  while (cols < G.nacc * (G.split ? G.nsacc : G.mh) * G.bn) cols <<= 1;
  if (cols > 512) return 3;
  G.tmem_cols = cols;
  G.idesc = make_idesc_bf16(128, G.bn);
  G.epi_bw = tc_epilogue_bw(d, G.bn);

  {   // VCM_CONV_SE_SUM: one tile = one whole 128-pixel x Cout block of one image
    const bool se_ok = G.tma_epi == 1 && G.nsub == 1 && G.n_tiles == 1 && P.N == 1 && !G.ps_q4 &&
                       !(P.flags & VCM_CONV_OUT2) && P.Wo % HTW == 0 && P.Ho % (HTH * G.mh) == 0;
    const int grid_q = G.num_items < num_sms() ? G.num_items : num_sms();
    if (se_rows) {
      *se_rows = se_ok ? 8 * grid_q : 0;
      return 0;
    }
    if ((P.flags & VCM_CONV_SE_SUM) && !se_ok) {
      vcm_set_error("vcm_conv_halo_launch: VCM_CONV_SE_SUM is not available for this launch (see vcm_conv2d_se_sum_rows)");
      return 1;
    }
  }
  CUtensorMap tmA, tmB;
  {
    cuuint64_t dims[4] = {(cuuint64_t)(P.Cin + (G.split ? 2 * G.pstride : 0)), (cuuint64_t)P.W, (cuuint64_t)P.H,
                          (cuuint64_t)P.N};
    cuuint64_t strides[3] = {(cuuint64_t)P.in_cs * 2, (cuuint64_t)P.W * P.in_cs * 2,
                             (cuuint64_t)P.H * P.W * P.in_cs * 2};
    cuuint32_t box[4] = {(cuuint32_t)chunk, (cuuint32_t)G.hw, (cuuint32_t)G.hh, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_of(chunk), CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      vcm_set_error("vcm_conv_halo_launch: cuTensorMapEncodeTiled(A) failed (%d)", (int)r);
      return 1;
    }
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)chunk, (cuuint64_t)slabs * G.cout_p};
    cuuint64_t strides[1] = {(cuuint64_t)chunk * 2};
    cuuint32_t box[2] = {(cuuint32_t)chunk, (cuuint32_t)G.bn};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(w_umma), dims, strides, box,
                     estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_of(chunk), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      vcm_set_error("vcm_conv_halo_launch: cuTensorMapEncodeTiled(B) failed (%d)", (int)r);
      return 1;
    }
  }
  CUtensorMap tmY, tmY2, tmR;
  memset(&tmY, 0, sizeof(tmY));
  memset(&tmY2, 0, sizeof(tmY2));
  memset(&tmR, 0, sizeof(tmR));
  if (G.tma_epi == 2) {
    cuuint64_t dims[4] = {(cuuint64_t)P.Cout, (cuuint64_t)P.Wo, (cuuint64_t)P.Ho, (cuuint64_t)P.N};
    cuuint64_t strides[3] = {(cuuint64_t)P.out_cs * 4, (cuuint64_t)P.Wo * P.out_cs * 4,
                             (cuuint64_t)P.Ho * P.Wo * P.out_cs * 4};
    cuuint32_t box[4] = {32, (cuuint32_t)HTW, (cuuint32_t)HTH, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    if (enc(&tmY, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, P.y, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
      vcm_set_error("vcm_conv_halo_launch: cuTensorMapEncodeTiled(fp32 output) failed");
      return 1;
    }
  } else if (G.tma_epi) {
    auto out_map = [&](CUtensorMap* tm, const void* base, int cs) -> bool {
      const int up = G.ps_q4 ? 2 : 1;                      // PixelShuffle: the 2x image, every other pixel per phase
      const int ch = G.ps_q4 ? G.ps_q4 : P.Cout;
      cuuint64_t dims[4] = {(cuuint64_t)ch, (cuuint64_t)(up * P.Wo), (cuuint64_t)(up * P.Ho), (cuuint64_t)P.N};
      cuuint64_t strides[3] = {(cuuint64_t)cs * 2, (cuuint64_t)(up * P.Wo) * cs * 2,
                               (cuuint64_t)(up * P.Ho) * (up * P.Wo) * cs * 2};
      cuuint32_t box[4] = {(cuuint32_t)G.sw, (cuuint32_t)(up * HTW), (cuuint32_t)(up * HTH), 1};
      cuuint32_t estr[4] = {1, (cuuint32_t)up, (cuuint32_t)up, 1};
      return enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                 CU_TENSOR_MAP_INTERLEAVE_NONE, G.sw == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                 CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
    };
    bool ok = out_map(&tmY, P.y, P.out_cs);
    if (ok && (P.flags & VCM_CONV_OUT2)) ok = out_map(&tmY2, P.y2, P.out2_cs);
    if (ok && (P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE))) ok = out_map(&tmR, P.res, P.res_cs);
    if (!ok) {
      vcm_set_error("vcm_conv_halo_launch: cuTensorMapEncodeTiled(output) failed");
      return 1;
    }
  }
  const int b_slots = G.resident ? slabs : G.b_stages;
  G.epi_base = G.b_base + b_slots * G.b_slot_bytes;
  G.bar_base = G.epi_base + epi_bytes;
  G.bias_base = G.bar_base + 512;
  int smem_bytes = G.bias_base + bias_bytes + 1024;
  if (smem_bytes < 120 * 1024) smem_bytes = 120 * 1024;   // one CTA per SM (TMEM is per CTA)
  const int grid = G.num_items < num_sms() ? G.num_items : num_sms();
#define LAUNCH_HALO(CH)                                                                                  \
  do {                                                                                                   \
    static bool attr_set[MAX_DEVICES] = {false};                                                         \
    cudaError_t e = ensure_max_smem(conv_halo_kernel<CH>, attr_set, SMEM_LIMIT);                         \
    if (e != cudaSuccess) {                                                                              \
      vcm_set_error("vcm_conv_halo_launch: cudaFuncSetAttribute: %s", cudaGetErrorString(e));            \
      return 1;                                                                                          \
    }                                                                                                    \
    cudaLaunchConfig_t cfg = {};                                                                         \
    cfg.gridDim = dim3(grid);                                                                            \
    cfg.blockDim = dim3(H_THREADS);                                                                      \
    cfg.dynamicSmemBytes = smem_bytes;                                                                   \
    cfg.stream = stream;                                                                                 \
    cudaLaunchAttribute attr[1];                                                                         \
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                                     \
    attr[0].val.programmaticStreamSerializationAllowed = 1;                                              \
    cfg.attrs = attr;                                                                                    \
    cfg.numAttrs = (P.flags & VCM_CONV_PDL) ? 1 : 0;                                                     \
    cudaLaunchKernelEx(&cfg, conv_halo_kernel<CH>, tmA, tmB, tmY, tmY2, tmR, P, G);                      \
  } while (0)
  if (chunk == 64) LAUNCH_HALO(64);
  else if (chunk == 32) LAUNCH_HALO(32);
  else LAUNCH_HALO(16);
#undef LAUNCH_HALO
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    vcm_set_error("vcm_conv_halo_launch: launch failed: %s", cudaGetErrorString(e));
    return 1;
  }
  return 0;
}

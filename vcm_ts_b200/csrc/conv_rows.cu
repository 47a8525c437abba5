// vcm_ts_b200/csrc/conv_rows.cu -- 3x3 stride-1 tensor-core convolution for NARROW outputs (Cout = 64 / 32)
// on sm_100a: output rows stacked along the MMA's N dimension.
//
// Why.  A 128 x N x 16 tcgen05.mma with both operands in shared memory takes ~64 SM cycles for EVERY N <= 128
// (measured: scripts/umma_rate.py, profiles/umma_rate_r02.json -- 66 cycles at N = 16 ... 128, 128 at N = 256,
// independent of how many accumulators the chain rotates over): the 4 KB A slab is read at 64 B/clk.  A layer
// with 64 output channels therefore runs the tensor pipe at half rate in the halo engine (conv_halo.cu),
// one with 32 channels at a quarter, whatever the tiling.  Widening N needs more columns that share the A
// operand.  A conv has them: the input row r is an operand of output rows r-1, r, r+1 (taps dy = +1, 0, -1).
//
// Here an item is R = 128 / Cout consecutive output rows of a 128-pixel-wide strip.  The A operand of an MMA
// is ONE input row (128 pixels shifted by the tap's dx); its B operand is the weights of up to three taps
// (dy = +1, 0, -1, stored in that order, so consecutive output rows are consecutive weight slabs) side by
// side, N = Cout * (rows fed): the accumulator holds the R rows next to each other in TMEM columns.  Per item
// of R = 2 rows: 4 input rows x 3 dx = 12 MMA groups instead of 18 (R = 4: 18 instead of 36), all at the same
// 64 cycles.  Input rows stream through a shared-memory ring (one TMA box {Cin, 130, 1} per row, zero-filled
// outside the image = the padding) and are read by every item that needs them: a persistent CTA owns a
// contiguous run of items down a strip, so each item loads only its R new rows.  Weights are resident.
// The epilogue is tc_epilogue.cuh's TMA-store path, one [128 px][Cout] tile per output row.
//
// The same holds for a K x K filter: input row r feeds K output rows, an item of R rows reads R + K - 1 input
// rows, and up to min(R, K) rows share one MMA (N = rows fed * BN <= 128).  The narrower the layer the more
// this buys: SpyNet's 7x7 32->16 and 16->2 convs (BN = 16: one MMA of the halo engine fills an eighth of the
// tensor rate) stack 8 rows, 14 x 7 MMA groups per 8 output rows instead of 8 x 49.
//
// Taken for: bf16 input, stride 1, one K chunk (Cin in {16, 32, 64}), resident weights, and
//   K = 3: Cout 64 (R = 2), 32 (R = 4) with the TMA-store epilogue; Cout <= 16 (R = 4, e.g. the 64->3 output conv)
//   K = 7: 16->32 (R = 4), 32->16 (R = 8), 16->2 (R = 8; fp32 output + fp32 residual through the scalar epilogue)
// on images at least 1024 wide.  Everything else stays on conv_halo.cu / conv_tc.cu.  Same products and the same
// fp32 accumulation per output as those engines, in a different order (rows of taps instead of taps of rows).
#include "conv_common.cuh"
#include "tc_epilogue.cuh"
#include "tc_host.cuh"
using namespace tc;

namespace {

constexpr int R_THREADS = 384;
constexpr int R_STRIP = 128;          // output pixels per strip row (= MMA M)
constexpr int R_MAX_SLOTS = 20;       // input-row ring
constexpr int R_NACC = 4;             // TMEM accumulators of 128 columns

struct RowsGeom {
  int rows;            // R: output rows per item; rows * bn <= 128
  int bn;              // N columns per output row = Cout padded to 16
  int scalar_epi;      // 0: TMA-store epilogue (bf16, bn in {64, 32, 16}); 1 / 2: scalar epilogue, bf16 / fp32 output
  int tile_bytes;      // one staging tile of the TMA epilogue (128 * bn * 2, rounded up to 1 KB)
  int slots, slot_bytes, a_bytes;
  int b_base, b_slot_bytes;
  int strips, groups, num_items;      // groups = Ho / R per strip; items = N * strips * groups
  int epi_base, bar_base, bias_base;
  int scale_off;       // byte offset (from bias_base) of the shared-memory copy of res_scale, 0 = none
};

// item -> (n, strip, group); consecutive items walk DOWN a strip
__device__ __forceinline__ void rows_decode(const RowsGeom& G, int item, int& n, int& strip, int& grp) {
  grp = item % G.groups;
  const int t = item / G.groups;
  strip = t % G.strips;
  n = t / G.strips;
}

struct RowsBars {
  uint64_t *afull, *aempty, *bfull, *tfull, *tempty;
};

// MMA-issuing thread.  As in conv_halo.cu the issuing thread, not the tensor pipe, paces the kernel unless every
// operand descriptor is `base + compile-time constant`: a first version with runtime loops over everything
// spent ~135 cycles per MMA (6700 cycles per item against 3100 of tensor-pipe time).  So K, R and BN are
// template parameters, the ring position is kept as (slot, phase) counters instead of divisions, descriptors
// are 32-bit low words + one shared high word, and the K * KSTEPS MMAs of an input row are straight-line code.
// The loop over the R + K - 1 input rows of an item is unrolled too where that stays small (3x3); the 7x7
// variants keep it a loop (up to 200 MMAs per item would not fit the instruction cache beside the epilogue).
template <int CHUNK, int R, int BN, int K>
__device__ __forceinline__ void rows_mma_role(const RowsGeom& G, const RowsBars& B, uint32_t a_smem, uint32_t b_smem,
                                              uint32_t tmem_base, int i_begin, int i_end) {
  constexpr uint32_t LAYOUT = CHUNK == 64 ? 2u : (CHUNK == 32 ? 4u : 6u);
  constexpr uint32_t ROWB = CHUNK * 2u;
  constexpr int KSTEPS = CHUNK / 16;
  constexpr int NR = R + K - 1;                      // input rows of an item; the last K - 1 are the next item's first
  constexpr uint32_t BSLOT16 = ((uint32_t)BN * ROWB) >> 4;     // one tap's weight slab, in 16-byte units
  constexpr bool UNROLL_ROWS = NR * K * KSTEPS <= 80;
  // descriptor = (start address >> 4 | ...) low word, (SBO | version | layout) high word: the high word is the
  // same for every A and B operand here (same swizzle, same 8-row group stride)
  const uint64_t bdesc0 = make_smem_desc(b_smem, 8u * ROWB, LAYOUT);
  const uint32_t b_lo0 = (uint32_t)bdesc0, d_hi = (uint32_t)(bdesc0 >> 32);
  const int slots = G.slots;
  const uint32_t slot_bytes = (uint32_t)G.slot_bytes;
  mbar_wait(B.bfull, 0);
  tc_fence_after();
  int sb = 0, acc = 0;              // ring slot of the item's input row j = 0
  uint32_t ph = 0, acc_phase = 0;   // its phase
  int grp = i_begin % G.groups;
  for (int item = i_begin; item < i_end; ++item) {
    const bool first = (item == i_begin) || grp == 0;
    const bool last = (item == i_end - 1) || grp == G.groups - 1;
    mbar_wait(&B.tempty[acc], acc_phase ^ 1u);
    tc_fence_after();
    const uint32_t d_tmem = tmem_base + (uint32_t)acc * 128u;
#pragma unroll(UNROLL_ROWS ? NR : 1)
    for (int j = 0; j < NR; ++j) {
      int slot = sb + j;
      uint32_t phase = ph;
      if (slot >= slots) { slot -= slots; phase ^= 1u; }
      if (first || j >= K - 1) {
        mbar_wait(&B.afull[slot], phase);
        tc_fence_after();
      }
      // input row y0 - K/2 + j feeds output rows s = s_lo .. s_hi with filter rows ky = j - s
      const int s_lo = j - (K - 1) > 0 ? j - (K - 1) : 0;
      const int s_hi = j < R - 1 ? j : R - 1;
      const int cnt = s_hi - s_lo + 1;
      const bool fresh = j < R;          // row s_hi = j gets its first product in this step
      const uint32_t a_lo0 = (uint32_t)make_smem_desc(a_smem + (uint32_t)slot * slot_bytes, 8u * ROWB, LAYOUT);
      const uint32_t d0 = d_tmem + (uint32_t)(s_lo * BN);
      const uint32_t bslab = (uint32_t)(K - 1 - j + s_lo);      // slabs of one kx are stored ky = K-1 .. 0
      const uint32_t idesc_all = make_idesc_bf16(128, cnt * BN);
#pragma unroll
      for (int kx = 0; kx < K; ++kx) {
        const uint32_t a_lo = a_lo0 + (uint32_t)kx * (ROWB >> 4);
        const uint32_t b_lo = b_lo0 + ((uint32_t)(kx * K) + bslab) * BSLOT16;
#pragma unroll
        for (int k = 0; k < KSTEPS; ++k) {
          if (kx == 0 && k == 0 && fresh) {
            // one instruction has one accumulate flag: the rows already begun and the fresh row separately
            if (cnt > 1) umma_bf16_lh(d0, a_lo, d_hi, b_lo, d_hi, make_idesc_bf16(128, (cnt - 1) * BN), 1u);
            umma_bf16_lh(d0 + (uint32_t)((cnt - 1) * BN), a_lo, d_hi, b_lo + (uint32_t)(cnt - 1) * BSLOT16, d_hi,
                         make_idesc_bf16(128, BN), 0u);
          } else {
            umma_bf16_lh(d0, a_lo + (uint32_t)(2 * k), d_hi, b_lo + (uint32_t)(2 * k), d_hi, idesc_all, 1u);
          }
        }
      }
      if (j < R || last) umma_commit(&B.aempty[slot]);   // later rows are the next item's first K - 1
    }
    umma_commit(&B.tfull[acc]);
    sb += last ? NR : R;
    if (sb >= slots) { sb -= slots; ph ^= 1u; }
    if (++grp == G.groups) grp = 0;
    if (++acc == R_NACC) { acc = 0; acc_phase ^= 1u; }
  }
}

template <int CHUNK, int R, int BN, int K, bool SCALAR>
__global__ void __launch_bounds__(R_THREADS, 1)
conv_rows_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmY, const __grid_constant__ CUtensorMap tmY2,
                 const __grid_constant__ CUtensorMap tmR, const ConvParams P, const RowsGeom G) {
  constexpr uint32_t LAYOUT = CHUNK == 64 ? 2u : (CHUNK == 32 ? 4u : 6u);
  constexpr uint32_t ROWB = CHUNK * 2u;      // bytes per pixel of an input row / per weight row
  constexpr int KSTEPS = CHUNK / 16;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + G.bar_base);
  uint64_t* afull = bars;
  uint64_t* aempty = afull + R_MAX_SLOTS;
  uint64_t* bfull = aempty + R_MAX_SLOTS;    // [1]
  uint64_t* tfull = bfull + 1;
  uint64_t* tempty = tfull + R_NACC;
  uint64_t* rfull = tempty + R_NACC;         // [2][2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(rfull + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // this CTA's contiguous run of items
  const int i_begin = (int)((long long)blockIdx.x * G.num_items / gridDim.x);
  const int i_end = (int)((long long)(blockIdx.x + 1) * G.num_items / gridDim.x);

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    prefetch_tmap(&tmY);
    if (P.flags & VCM_CONV_OUT2) prefetch_tmap(&tmY2);   // (unused maps are zero-filled: prefetching them is harmless)
    if (P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE)) prefetch_tmap(&tmR);
    for (int s = 0; s < R_MAX_SLOTS; ++s) { mbar_init(&afull[s], 1); mbar_init(&aempty[s], 1); }
    mbar_init(bfull, 1);
    for (int a = 0; a < R_NACC; ++a) { mbar_init(&tfull[a], 1); mbar_init(&tempty[a], 4); }
    for (int r = 0; r < 4; ++r) mbar_init(&rfull[r], 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512u);
  {
    float* bs_ = reinterpret_cast<float*>(smem + G.bias_base);
    for (int i = threadIdx.x; i < G.bn; i += blockDim.x) bs_[i] = i < P.Cout ? __ldg(P.bias + i) : 0.f;
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
  pdl_trigger();

  if (warp < 4) {
    reg_dec<72>();
    if (warp == 0) {
      // ============================ input-row producer ===========================================
      if (lane == 0) {
        pdl_wait();
        int q = 0;   // sequence number of the next row in this CTA's row stream
        for (int item = i_begin; item < i_end; ++item) {
          int n, strip, grp;
          rows_decode(G, item, n, strip, grp);
          const bool first = (item == i_begin) || grp == 0;
          for (int j = first ? 0 : K - 1; j < R + K - 1; ++j, ++q) {
            const int slot = q % G.slots;
            mbar_wait(&aempty[slot], (((uint32_t)(q / G.slots)) & 1u) ^ 1u);
            mbar_expect_tx(&afull[slot], (uint32_t)G.a_bytes);
            tma_load_4d(smem_u32(smem + (size_t)slot * G.slot_bytes), &tmA, &afull[slot], 0,
                        strip * R_STRIP - K / 2, grp * R - K / 2 + j, n);
          }
        }
      }
    } else if (warp == 2) {
      // ============================ weights: resident, slab (ky, kx) -> slot kx * K + (K - 1 - ky) ==
      if (lane == 0) {
        if (!(P.flags & VCM_CONV_PDL)) pdl_wait();
        mbar_expect_tx(bfull, (uint32_t)(K * K * BN * CHUNK * 2));
        for (int ky = 0; ky < K; ++ky)
          for (int kx = 0; kx < K; ++kx)
            tma_load_2d(smem_u32(smem + G.b_base + (size_t)(kx * K + (K - 1 - ky)) * G.b_slot_bytes), &tmB, bfull, 0,
                        (ky * K + kx) * BN);
      }
    } else if (warp == 1) {
      // ============================ MMA issue ====================================================
      if (lane == 0) {
        RowsBars B{afull, aempty, bfull, tfull, tempty};
        const uint32_t a_smem = smem_u32(smem), b_smem = smem_u32(smem + G.b_base);
        rows_mma_role<CHUNK, R, BN, K>(G, B, a_smem, b_smem, tmem_base, i_begin, i_end);
      }
    }
  } else {
    reg_inc<216>();
    // ============ epilogue: warpgroup g takes every other item; one [128 px][bn] tile per output row ======
    const int g = (warp - 4) >> 2;
    const int q = warp & 3;
    pdl_wait();
    if constexpr (SCALAR) {
      // Cout below 16 and / or fp32 output (SpyNet conv5 + flow residual, the 64->3 output conv): thread = pixel,
      // global loads / stores (tc_epilogue.cuh epilogue_thread_scalar); few channels, so few bytes
      int local = 0;
      for (int item = i_begin; item < i_end; ++item, ++local) {
        if ((local & 1) != g) continue;
        const int ab = local % R_NACC;
        const uint32_t acc_phase = (uint32_t)(local / R_NACC) & 1u;
        int n, strip, grp;
        rows_decode(G, item, n, strip, grp);
        const int ox = strip * R_STRIP + q * 32 + lane;
        for (int h = 0; h < R; ++h) {
          const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(ab * 128 + h * BN);
          uint64_t* wait = h == 0 ? &tfull[ab] : nullptr;
          uint64_t* release = h == R - 1 ? &tempty[ab] : nullptr;
          if (G.scalar_epi == 2)
            epilogue_thread_scalar<float>(P, BN, n, grp * R + h, ox, 0, taddr, wait, acc_phase, release);
          else
            epilogue_thread_scalar<__nv_bfloat16>(P, BN, n, grp * R + h, ox, 0, taddr, wait, acc_phase, release);
        }
      }
    } else {
    EpiTma E;
    E.tmY = &tmY; E.tmY2 = &tmY2; E.tmR = &tmR;
    E.obuf = smem_u32(smem + G.epi_base) + (uint32_t)(g * 2 * G.tile_bytes);
    E.tile_bytes = (uint32_t)G.tile_bytes;
    E.scale_smem = G.scale_off ? smem_u32(smem + G.bias_base + G.scale_off) : 0u;
    E.rfull = &rfull[2 * g];
    E.k = 0;
    E.se_sum = (P.flags & VCM_CONV_SE_SUM) != 0;
    E.se0 = E.se1 = 0.f;
    E.leader = (q == 0 && lane == 0);
    E.bar = 1 + g;
    const bool has_res = P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE);
    auto tile = [&](int k, int& h, int& n, int& oy0, int& ox0) -> bool {
      const int loc = 2 * (k / R) + g;
      h = k % R;
      const int item = i_begin + loc;
      if (item >= i_end) return false;
      int strip, grp;
      rows_decode(G, item, n, strip, grp);
      ox0 = strip * R_STRIP;
      oy0 = grp * R + h;
      return true;
    };
    int h, n, oy0, ox0;
    bool ok = tile(0, h, n, oy0, ox0);
    if (ok && has_res && E.leader) epi_tma_prefetch_res<BN>(E, 0, 0, ox0, oy0, n);
    for (int k = 0; ok; ++k) {
      int h2, n2, oy2, ox2;
      const bool ok2 = tile(k + 1, h2, n2, oy2, ox2);
      const int loc = 2 * (k / R) + g;
      const int ab = loc % R_NACC;
      const uint32_t acc_phase = (uint32_t)(loc / R_NACC) & 1u;
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(ab * 128 + h * BN);
      epilogue_warp_tma<BN>(P, E, n, oy0, ox0, 0, 0, taddr, q, lane, h == 0 ? &tfull[ab] : nullptr, acc_phase,
                            h == R - 1 ? &tempty[ab] : nullptr, bias_smem, ok2, n2, oy2, ox2, 0);
      ok = ok2; h = h2; n = n2; oy0 = oy2; ox0 = ox2;
    }
    if (E.se_sum) epi_tma_write_se_sums<BN>(E, reinterpret_cast<float*>(P.y2), warp - 4, lane);
    if (E.leader) tma_store_wait_all();
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512u);
  }
}

}  // namespace

// ---------------------------------------------------------------------------------- host
// returns 0 ok, 1 error, 3 = shape not taken
int vcm_conv_rows_launch(const vcm_conv_desc* d, const ConvParams& P, const void* x, const void* w_umma,
                         cudaStream_t stream, int* se_rows) {
  const char* off_env = getenv("VCM_TC_NO_ROWS");   // read per launch: tests compare the engines in one process
  const bool off = off_env && off_env[0] == '1';
  static const bool no_tma_epi = getenv("VCM_TC_NO_TMA_EPI") && getenv("VCM_TC_NO_TMA_EPI")[0] == '1';
  if (off || no_tma_epi) return 3;
  if (!(P.K == 3 || P.K == 7) || P.stride != 1 || P.pad != P.K / 2 || (P.flags & VCM_CONV_SPLIT3)) return 3;
  if (d->in_dtype != VCM_BF16) return 3;
  if (!(P.Cin == 64 || P.Cin == 32 || P.Cin == 16)) return 3;
  if (P.flags & (VCM_CONV_PIXEL_SHUFFLE | VCM_CONV_PRE_ACT)) return 3;
  RowsGeom G;
  memset(&G, 0, sizeof(G));
  const int chunk = P.Cin;
  G.bn = round_up(P.Cout, 16);
  G.scalar_epi = tc_epilogue_mode(d, P);             // 0 vector-capable bf16, 1 scalar bf16, 2 fp32 output
  if (G.bn > 16 && G.scalar_epi != 0) return 3;      // wide tiles only through the TMA-store epilogue
  if (G.bn == 16 && P.Cout != 16 && G.scalar_epi == 0) G.scalar_epi = 1;   // padded channels: no TMA tile
  // instantiated: scalar epilogue for (K 3, Cin 64, bn 16) and (K 7, Cin 16, bn 16); TMA epilogue for the rest
  if (G.bn == 16) {
    const bool scalar_inst = (P.K == 3 && chunk == 64) || (P.K == 7 && chunk == 16);
    if (scalar_inst != (G.scalar_epi != 0)) return 3;
  }
  if (G.scalar_epi == 0) {
    if (P.flags & VCM_CONV_RES2) return 3;
    if ((P.flags & VCM_CONV_OUT2) && (P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE))) return 3;
  }
  // the instantiated (K, Cin chunk, rows per item, N columns per row) combinations
  if (P.K == 3) {
    if (G.bn == 64) G.rows = 2;
    else if (G.bn == 32) G.rows = 4;
    else if (G.bn == 16 && chunk == 64) G.rows = 4;   // 8 rows of 64-channel input do not fit beside the ring
    else return 3;
  } else {
    if (chunk == 16 && G.bn == 32) G.rows = 4;        // SpyNet conv1 (8 -> 32, input padded to 16)
    else if (chunk == 32 && G.bn == 16) G.rows = 8;   // conv4 32 -> 16
    else if (chunk == 16 && G.bn == 16) G.rows = 8;   // conv5 16 -> 2 (+ flow residual, fp32)
    else return 3;
  }
  if (P.Ho % G.rows) return 3;
  G.strips = cdiv(P.Wo, R_STRIP);
  G.groups = P.Ho / G.rows;
  G.num_items = P.N * G.strips * G.groups;
  // Worth it only where every SM gets a long run of items: measured on B200 (scripts/conv_bench.py, same box,
  // halo engine -> this one) 1088x1920: 64->32 134 -> 112 us, 32->32 91 -> 71, 16->64 106 -> 94, 64->64 155 -> 145
  // (with residual 172 -> 167: HBM-bound either way); 544x960 (15 items per SM, 7.5 strips): 64->64 + residual
  // 49 -> 55 us, so the halo engine keeps the smaller images.  VCM_TC_ROWS_MIN_W overrides the width
  // threshold (tests use it to reach this engine with small images).
  const char* mw_env = getenv("VCM_TC_ROWS_MIN_W");
  const int min_w = mw_env ? atoi(mw_env) : 1024;
  if (P.Wo < min_w || G.num_items < 4 * num_sms()) return 3;
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    vcm_set_error("vcm_conv_rows_launch: cuTensorMapEncodeTiled unavailable");
    return 1;
  }
  const int taps = P.K * P.K;
  G.a_bytes = (R_STRIP + P.K - 1) * chunk * 2;
  G.slot_bytes = round_up(G.a_bytes, 1024);
  G.b_slot_bytes = G.bn * chunk * 2;                 // 64x64: 8 KB ... 16x16: 512 B: multiples of the swizzle atom
  if (G.b_slot_bytes % (8 * chunk * 2)) return 3;
  int bias_bytes = 512;
  if ((P.flags & VCM_CONV_RES_SCALE) && !(P.flags & VCM_CONV_PDL) && P.N * P.Cout <= 2048) {
    G.scale_off = bias_bytes;
    bias_bytes += round_up(P.N * P.Cout * 4, 128);
  }
  G.tile_bytes = round_up(128 * G.bn * 2, 1024);
  const int epi_bytes = G.scalar_epi ? 0 : 4 * G.tile_bytes;   // two staging tiles per epilogue warpgroup
  const int budget = SMEM_LIMIT - 1024 - 512 - bias_bytes - epi_bytes - round_up(taps * G.b_slot_bytes, 1024);
  G.slots = budget / G.slot_bytes;
  if (G.slots > R_MAX_SLOTS) G.slots = R_MAX_SLOTS;
  if (G.slots < G.rows + P.K) return 3;              // an item's R + K - 1 rows and at least one in flight
  G.b_base = round_up(G.slots * G.slot_bytes, 1024);
  G.epi_base = round_up(G.b_base + taps * G.b_slot_bytes, 1024);
  G.bar_base = G.epi_base + epi_bytes;
  G.bias_base = G.bar_base + 512;
  int smem_bytes = G.bias_base + bias_bytes + 1024;
  if (smem_bytes > SMEM_LIMIT) return 3;
  if (smem_bytes < 120 * 1024) smem_bytes = 120 * 1024;   // one CTA per SM (TMEM is per CTA)

  {   // VCM_CONV_SE_SUM: whole strips of one image through the TMA epilogue
    const bool se_ok = !G.scalar_epi && P.N == 1 && P.Wo % R_STRIP == 0 && !(P.flags & VCM_CONV_OUT2) &&
                       (G.bn == 64 || G.bn == 32 || G.bn == 16);
    if (se_rows) {
      *se_rows = se_ok ? 8 * num_sms() : 0;
      return 0;
    }
    if ((P.flags & VCM_CONV_SE_SUM) && !se_ok) {
      vcm_set_error("vcm_conv_rows_launch: VCM_CONV_SE_SUM is not available for this launch (see vcm_conv2d_se_sum_rows)");
      return 1;
    }
  }
  CUtensorMap tmA, tmB, tmY, tmY2, tmR;
  memset(&tmY, 0, sizeof(tmY));
  memset(&tmY2, 0, sizeof(tmY2));
  memset(&tmR, 0, sizeof(tmR));
  {
    cuuint64_t dims[4] = {(cuuint64_t)P.Cin, (cuuint64_t)P.W, (cuuint64_t)P.H, (cuuint64_t)P.N};
    cuuint64_t strides[3] = {(cuuint64_t)P.in_cs * 2, (cuuint64_t)P.W * P.in_cs * 2,
                             (cuuint64_t)P.H * P.W * P.in_cs * 2};
    cuuint32_t box[4] = {(cuuint32_t)chunk, (cuuint32_t)(R_STRIP + P.K - 1), 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_of(chunk), CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      vcm_set_error("vcm_conv_rows_launch: cuTensorMapEncodeTiled(A) failed (%d)", (int)r);
      return 1;
    }
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)chunk, (cuuint64_t)taps * G.bn};
    cuuint64_t strides[1] = {(cuuint64_t)chunk * 2};
    cuuint32_t box[2] = {(cuuint32_t)chunk, (cuuint32_t)G.bn};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(w_umma), dims, strides, box,
                     estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_of(chunk), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      vcm_set_error("vcm_conv_rows_launch: cuTensorMapEncodeTiled(B) failed (%d)", (int)r);
      return 1;
    }
  }
  if (!G.scalar_epi) {
    auto out_map = [&](CUtensorMap* tm, const void* base, int cs) -> bool {
      cuuint64_t dims[4] = {(cuuint64_t)P.Cout, (cuuint64_t)P.Wo, (cuuint64_t)P.Ho, (cuuint64_t)P.N};
      cuuint64_t strides[3] = {(cuuint64_t)cs * 2, (cuuint64_t)P.Wo * cs * 2, (cuuint64_t)P.Ho * P.Wo * cs * 2};
      cuuint32_t box[4] = {(cuuint32_t)G.bn, (cuuint32_t)R_STRIP, 1, 1};
      cuuint32_t estr[4] = {1, 1, 1, 1};
      return enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                 CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_of(G.bn), CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
    };
    bool ok = out_map(&tmY, P.y, P.out_cs);
    if (ok && (P.flags & VCM_CONV_OUT2)) ok = out_map(&tmY2, P.y2, P.out2_cs);
    if (ok && (P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE))) ok = out_map(&tmR, P.res, P.res_cs);
    if (!ok) {
      vcm_set_error("vcm_conv_rows_launch: cuTensorMapEncodeTiled(output) failed");
      return 1;
    }
  }
  const int grid = num_sms();
#define LAUNCH_ROWS(CH, RR, BNN, KK, SC)                                                                   \
  do {                                                                                                   \
    static bool attr_set[MAX_DEVICES] = {false};                                                         \
    cudaError_t e = ensure_max_smem(conv_rows_kernel<CH, RR, BNN, KK, SC>, attr_set, SMEM_LIMIT);            \
    if (e != cudaSuccess) {                                                                              \
      vcm_set_error("vcm_conv_rows_launch: cudaFuncSetAttribute: %s", cudaGetErrorString(e));            \
      return 1;                                                                                          \
    }                                                                                                    \
    cudaLaunchConfig_t cfg = {};                                                                         \
    cfg.gridDim = dim3(grid);                                                                            \
    cfg.blockDim = dim3(R_THREADS);                                                                      \
    cfg.dynamicSmemBytes = smem_bytes;                                                                   \
    cfg.stream = stream;                                                                                 \
    cudaLaunchAttribute attr[1];                                                                         \
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                                     \
    attr[0].val.programmaticStreamSerializationAllowed = 1;                                              \
    cfg.attrs = attr;                                                                                    \
    cfg.numAttrs = (P.flags & VCM_CONV_PDL) ? 1 : 0;                                                     \
    cudaLaunchKernelEx(&cfg, conv_rows_kernel<CH, RR, BNN, KK, SC>, tmA, tmB, tmY, tmY2, tmR, P, G);         \
  } while (0)
  if (P.K == 3) {
    if (G.bn == 64) {
      if (chunk == 64) LAUNCH_ROWS(64, 2, 64, 3, false);
      else if (chunk == 32) LAUNCH_ROWS(32, 2, 64, 3, false);
      else LAUNCH_ROWS(16, 2, 64, 3, false);
    } else if (G.bn == 32) {
      if (chunk == 64) LAUNCH_ROWS(64, 4, 32, 3, false);
      else if (chunk == 32) LAUNCH_ROWS(32, 4, 32, 3, false);
      else LAUNCH_ROWS(16, 4, 32, 3, false);
    } else {
      LAUNCH_ROWS(64, 4, 16, 3, true);
    }
  } else {
    if (chunk == 16 && G.bn == 32) LAUNCH_ROWS(16, 4, 32, 7, false);
    else if (chunk == 32) LAUNCH_ROWS(32, 8, 16, 7, false);
    else LAUNCH_ROWS(16, 8, 16, 7, true);
  }
#undef LAUNCH_ROWS
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    vcm_set_error("vcm_conv_rows_launch: launch failed: %s", cudaGetErrorString(e));
    return 1;
  }
  return 0;
}

// vcm_ts_b200/csrc/wgrad_tc.cu -- convolution weight gradient on tcgen05 (bf16 operands, fp32 TMEM
// accumulate): the implicit GEMM over pixels of the training path (the reference gets
// it from torch autograd, core/engine/train_multi.py:203-240 `loss_to_opt.backward()`).
//
//   dW[tap][co][ci] = sum over output pixels p of dY[p][co] * X[p*stride + tap - pad][ci]
//
// GEMM view per tap: D[M = co][N = ci] += A[M][K = pixel] * B[N][K = pixel].  Both operands live in memory
// as NHWC, i.e. [pixel][channel] with the CHANNEL contiguous: that is an "MN-major" UMMA operand (the
// instruction descriptor's a_major / b_major bits), whose canonical 128B-swizzled shared-memory image --
// rows of 64 channels (128 B), one row per k = pixel, 8-row swizzle atoms 1024 B apart, 64-channel blocks
// LBO apart -- is exactly what a plain TMA box {64 ch, bw, bh} of the NHWC tensor writes.  So neither
// operand is transposed anywhere: the A box is a 64-pixel patch of dY, the B box of tap (ky, kx) the same
// patch of X shifted by the tap (TMA zero-fills outside the image = the conv padding; element strides
// give stride-2 convs).
//
// Work item = (pixel slab, tap group, 128-row co tile, ci tile): all taps of the group accumulate in their
// own TMEM columns while the slab's pixel boxes stream through a shared-memory ring; the epilogue writes the
// fp32 partials [slab][tap][co][ci], which wgrad_reduce (train_kernels.cu) sums in fixed order.
// Roles (192 threads): warp 0 TMA producer, warp 1 MMA issue + TMEM, warps 2-5 epilogue.
#include "common.cuh"
#include "tc_host.cuh"
#include "tc_ptx.cuh"
using namespace tc;

namespace {

constexpr int WT_THREADS = 192;
constexpr int WT_PIX = 64;           // pixels (GEMM K) per box
constexpr int WT_BLK = 8192;         // bytes of one [64 px][64 ch] bf16 block
constexpr int WT_MAX_STAGES = 6;

struct WtGeom {
  int K, stride, pad, Cin, Cout;
  int bw, bh, tiles_x, tiles_y, n_boxes;
  int taps, tpg, ngroups, m_tiles, n_tiles, nn, nblk;
  int nsplit, boxes_per_split, num_items;
  int stages, stage_bytes, bar_base, tmem_cols;
  uint32_t idesc;
};

// M x N x 16 MMA, bf16 operands, BOTH MN-major (bits 15 / 16), fp32 accumulate
__host__ __device__ constexpr uint32_t make_idesc_bf16_mn(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(N >> 3) << 17) |
         ((uint32_t)(M >> 4) << 24);
}
// MN-major 128B-swizzled operand: lbo = bytes between 64-channel blocks, sbo = bytes between 8-pixel groups
__device__ __forceinline__ uint64_t make_desc_mn(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= 1ull << 46;
  d |= 2ull << 61;
  return d;
}

__global__ void __launch_bounds__(WT_THREADS, 1)
wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                float* __restrict__ partial, const WtGeom G) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + G.bar_base);
  uint64_t* empty = full + WT_MAX_STAGES;
  uint64_t* tfull = empty + WT_MAX_STAGES;
  uint64_t* tempty = tfull + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int s = 0; s < G.stages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(tfull, 1);
    mbar_init(tempty, 4);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, (uint32_t)G.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int per_slab = G.ngroups * G.m_tiles * G.n_tiles;
  const int b_tap_bytes = G.nblk * WT_BLK;

  if (warp == 0) {
    // ===================================== TMA producer =====================================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = blockIdx.x; item < G.num_items; item += gridDim.x) {
        const int sl = item / per_slab;
        int r = item - sl * per_slab;
        const int g = r / (G.m_tiles * G.n_tiles);
        r -= g * (G.m_tiles * G.n_tiles);
        const int mt = r / G.n_tiles, nt = r - mt * G.n_tiles;
        const int tap0 = g * G.tpg;
        const int ntap = min(G.tpg, G.taps - tap0);
        const int b_end = min(G.n_boxes, (sl + 1) * G.boxes_per_split);
        for (int b = sl * G.boxes_per_split; b < b_end; ++b) {
          const int n = b / (G.tiles_x * G.tiles_y);
          const int rem = b - n * (G.tiles_x * G.tiles_y);
          const int oy0 = (rem / G.tiles_x) * G.bh, ox0 = (rem % G.tiles_x) * G.bw;
          mbar_wait(&empty[stage], phase ^ 1u);
          const uint32_t sa = smem_u32(smem + (size_t)stage * G.stage_bytes);
          mbar_expect_tx(&full[stage], (uint32_t)(2 * WT_BLK + ntap * b_tap_bytes));
          tma_load_4d(sa, &tmA, &full[stage], mt * 128, ox0, oy0, n);
          tma_load_4d(sa + WT_BLK, &tmA, &full[stage], mt * 128 + 64, ox0, oy0, n);
          for (int t = 0; t < ntap; ++t) {
            const int tap = tap0 + t;
            const int ky = tap / G.K, kx = tap - ky * G.K;
            for (int blk = 0; blk < G.nblk; ++blk)
              tma_load_4d(sa + 2 * WT_BLK + t * b_tap_bytes + blk * WT_BLK, &tmB, &full[stage],
                          nt * G.nn + blk * 64, ox0 * G.stride + kx - G.pad, oy0 * G.stride + ky - G.pad, n);
          }
          if (++stage == G.stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================== MMA issue ========================================
    int stage = 0;
    uint32_t phase = 0, acc_phase = 0;
    for (int item = blockIdx.x; item < G.num_items; item += gridDim.x) {
      const int sl = item / per_slab;
      const int g = (item - sl * per_slab) / (G.m_tiles * G.n_tiles);
      const int ntap = min(G.tpg, G.taps - g * G.tpg);
      const int b_begin = sl * G.boxes_per_split, b_end = min(G.n_boxes, (sl + 1) * G.boxes_per_split);
      mbar_wait(tempty, acc_phase ^ 1u);
      tc_fence_after();
      for (int b = b_begin; b < b_end; ++b) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(smem + (size_t)stage * G.stage_bytes);
          const uint64_t adesc = make_desc_mn(sa, WT_BLK, 1024u);
          for (int t = 0; t < ntap; ++t) {
            const uint64_t bdesc = make_desc_mn(sa + 2 * WT_BLK + t * b_tap_bytes, WT_BLK, 1024u);
            const uint32_t d = tmem_base + (uint32_t)(t * G.nn);
#pragma unroll
            for (int k = 0; k < WT_PIX / 16; ++k)   // 16 pixels = 2 swizzle atoms = 2048 B along K
              umma_bf16(d, adesc + (uint64_t)(128 * k), bdesc + (uint64_t)(128 * k), G.idesc,
                        (b > b_begin || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty[stage]);
          if (b == b_end - 1) umma_commit(tfull);
        }
        __syncwarp();
        if (++stage == G.stages) { stage = 0; phase ^= 1u; }
      }
      acc_phase ^= 1u;
    }
  } else {
    // ===================================== epilogue: TMEM -> partials ==========================
    const int q = warp & 3;                    // TMEM lane quarter this warp may access
    uint32_t acc_phase = 0;
    for (int item = blockIdx.x; item < G.num_items; item += gridDim.x) {
      const int sl = item / per_slab;
      int r = item - sl * per_slab;
      const int g = r / (G.m_tiles * G.n_tiles);
      r -= g * (G.m_tiles * G.n_tiles);
      const int mt = r / G.n_tiles, nt = r - mt * G.n_tiles;
      const int tap0 = g * G.tpg;
      const int ntap = min(G.tpg, G.taps - tap0);
      const int co = mt * 128 + q * 32 + lane;
      mbar_wait(tfull, acc_phase);
      tc_fence_after();
      for (int t = 0; t < ntap; ++t) {
        float* out = partial + (((size_t)sl * G.taps + tap0 + t) * G.Cout + co) * G.Cin + nt * G.nn;
        for (int c0 = 0; c0 < G.nn; c0 += 16) {
          uint32_t v[16];
          tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(t * G.nn + c0), v);
          if (co < G.Cout) {
#pragma unroll
            for (int i = 0; i < 16; i += 4)
              *reinterpret_cast<float4*>(out + c0 + i) =
                  make_float4(__uint_as_float(v[i]), __uint_as_float(v[i + 1]), __uint_as_float(v[i + 2]),
                              __uint_as_float(v[i + 3]));
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty);
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

bool wt_geometry(const vcm_conv_desc* d, WtGeom& G) {
  if (d->in_dtype != VCM_BF16 || d->out_dtype != VCM_BF16) return false;
  if (!(d->K == 1 || d->K == 3 || d->K == 7) || !(d->stride == 1 || d->stride == 2)) return false;
  if (d->Cin % 16 || d->in_cstride % 8 || d->out_cstride % 8) return false;
  if (d->stride == 2 && ((d->H & 1) || (d->W & 1))) return false;
  memset(&G, 0, sizeof(G));
  G.K = d->K; G.stride = d->stride; G.pad = d->K / 2; G.Cin = d->Cin; G.Cout = d->Cout;
  const int Ho = (d->H + 2 * G.pad - d->K) / d->stride + 1, Wo = (d->W + 2 * G.pad - d->K) / d->stride + 1;
  G.bw = 64;
  while (G.bw > 1 && G.bw / 2 >= Wo) G.bw >>= 1;
  G.bh = WT_PIX / G.bw;
  G.tiles_x = cdiv(Wo, G.bw);
  G.tiles_y = cdiv(Ho, G.bh);
  G.n_boxes = d->N * G.tiles_x * G.tiles_y;
  G.taps = d->K * d->K;
  G.n_tiles = 1;
  while (d->Cin / G.n_tiles > 256 || d->Cin % G.n_tiles || (d->Cin / G.n_tiles) % 16) ++G.n_tiles;
  G.nn = d->Cin / G.n_tiles;
  G.nblk = cdiv(G.nn, 64);
  G.m_tiles = cdiv(d->Cout, 128);
  int tpg_max = 512 / G.nn;
  if (tpg_max > 7 / G.nblk) tpg_max = 7 / G.nblk;
  if (tpg_max < 1) tpg_max = 1;
  G.ngroups = cdiv(G.taps, tpg_max);
  G.tpg = cdiv(G.taps, G.ngroups);
  G.ngroups = cdiv(G.taps, G.tpg);
  const int per_slab = G.ngroups * G.m_tiles * G.n_tiles;
  G.nsplit = cdiv(2 * num_sms(), per_slab);
  if (G.nsplit > G.n_boxes) G.nsplit = G.n_boxes;
  if (G.nsplit < 1) G.nsplit = 1;
  G.boxes_per_split = cdiv(G.n_boxes, G.nsplit);
  G.nsplit = cdiv(G.n_boxes, G.boxes_per_split);
  G.num_items = G.nsplit * per_slab;
  G.stage_bytes = 2 * WT_BLK + G.tpg * G.nblk * WT_BLK;
  G.stages = (SMEM_LIMIT - 2048) / G.stage_bytes;
  if (G.stages > WT_MAX_STAGES) G.stages = WT_MAX_STAGES;
  if (G.stages < 2) return false;
  G.bar_base = G.stages * G.stage_bytes;
  int cols = 32;
  while (cols < G.tpg * G.nn) cols <<= 1;
  if (cols > 512) return false;
  G.tmem_cols = cols;
  G.idesc = make_idesc_bf16_mn(128, G.nn);
  return true;
}

}  // namespace

// 0 = not taken (caller uses the SIMT kernel), else the number of scratch floats the tensor-core path needs
int64_t vcm_wgrad_tc_scratch_floats(const vcm_conv_desc* d) {
  WtGeom G;
  if (!wt_geometry(d, G)) return 0;
  return (int64_t)G.nsplit * G.taps * d->Cout * d->Cin;
}

// returns 0 ok, 1 error, 3 = shape not taken; *nsplit_out = slabs written to `scratch`
int vcm_wgrad_tc_launch(const vcm_conv_desc* d, const void* x, const void* dy, float* scratch, int* nsplit_out,
                        cudaStream_t stream) {
  WtGeom G;
  if (!wt_geometry(d, G)) return 3;
  if (((uintptr_t)x & 15) || ((uintptr_t)dy & 15) || ((uintptr_t)scratch & 15)) return 3;
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return 3;
  const int Ho = (d->H + 2 * G.pad - d->K) / d->stride + 1, Wo = (d->W + 2 * G.pad - d->K) / d->stride + 1;
  CUtensorMap tmA, tmB;
  {
    cuuint64_t dims[4] = {(cuuint64_t)d->Cout, (cuuint64_t)Wo, (cuuint64_t)Ho, (cuuint64_t)d->N};
    cuuint64_t strides[3] = {(cuuint64_t)d->out_cstride * 2, (cuuint64_t)Wo * d->out_cstride * 2,
                             (cuuint64_t)Ho * Wo * d->out_cstride * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)G.bw, (cuuint32_t)G.bh, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    if (enc(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(dy), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 3;
  }
  {
    cuuint64_t dims[4] = {(cuuint64_t)d->Cin, (cuuint64_t)d->W, (cuuint64_t)d->H, (cuuint64_t)d->N};
    cuuint64_t strides[3] = {(cuuint64_t)d->in_cstride * 2, (cuuint64_t)d->W * d->in_cstride * 2,
                             (cuuint64_t)d->H * d->W * d->in_cstride * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)(G.bw * d->stride), (cuuint32_t)(G.bh * d->stride), 1};
    cuuint32_t estr[4] = {1, (cuuint32_t)d->stride, (cuuint32_t)d->stride, 1};
    if (enc(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 3;
  }
  static bool attr_set[MAX_DEVICES] = {false};
  if (ensure_max_smem(wgrad_tc_kernel, attr_set, SMEM_LIMIT) != cudaSuccess) {
    vcm_set_error("vcm_wgrad_tc_launch: cudaFuncSetAttribute failed");
    return 1;
  }
  int smem_bytes = G.bar_base + 1024 + 256;
  if (smem_bytes < 120 * 1024) smem_bytes = 120 * 1024;      // one CTA per SM (TMEM is per CTA)
  const int grid = G.num_items < num_sms() ? G.num_items : num_sms();
  wgrad_tc_kernel<<<grid, WT_THREADS, smem_bytes, stream>>>(tmA, tmB, scratch, G);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    vcm_set_error("vcm_wgrad_tc_launch: launch failed: %s", cudaGetErrorString(e));
    return 1;
  }
  *nsplit_out = G.nsplit;
  return 0;
}

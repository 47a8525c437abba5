// vcm_ts_b200/csrc/probe/tc_probe.cu (libvcmts_probe.so, diagnostics only) -- probe of the UMMA shared-memory descriptor semantics (see
// include/vcmts_b200.h vcm_umma_probe): measured on B200, the 128B swizzle is a function of the
// absolute shared-memory address, so an A tile may start any number of 128-byte rows into a
// TMA-written buffer and use any multiple of 128 B as the stride between 8-row groups, with
// base_offset = 0 (profiles/umma_probe_r01.json).  The halo-tile conv variant relies on this.
#include <stdarg.h>

#include "../tc_ptx.cuh"
#include "../../../include/vcmts_b200_probe.h"
using namespace tc;

// the probe library is self-contained: its own error slot
void vcm_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vfprintf(stderr, fmt, ap);
  va_end(ap);
  fputc('\n', stderr);
}

namespace {
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

}  // namespace

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

// ---- MMA issue-rate probe: `iters` back-to-back 128 x N x 16 bf16 MMAs on fixed smem operands, one
// commit at the end; out[0] = SM cycles from first issue to completion, out[1] = iters.
namespace {
__global__ void __launch_bounds__(128, 1) umma_rate_kernel(int N, int iters, int start_row, int sbo, int nacc,
                                                           long long* __restrict__ out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;                 // 384 x 128 B (room for halo-style start rows / group strides)
  uint8_t* sB = smem + 384 * 128;     // 256 x 128 B
  uint64_t* bar = reinterpret_cast<uint64_t*>(sB + 256 * 128);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
  for (int i = threadIdx.x; i < (384 + 256) * 128 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    fence_barrier_init();
  }
  fence_proxy_async();
  if ((threadIdx.x >> 5) == 0) tmem_alloc(slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *slot;
  if (threadIdx.x == 0) {
    const uint64_t adesc = make_smem_desc(smem_u32(sA) + (uint32_t)start_row * 128u, (uint32_t)sbo, 2u);
    const uint64_t bdesc = make_smem_desc(smem_u32(sB), 1024u, 2u);
    const uint32_t idesc = make_idesc_bf16(128, N);
    const long long t0 = clock64();
    if (nacc <= 0) {
      for (int i = 0; i < iters; ++i) umma_bf16(tmem_base + (uint32_t)((i & 1) * 256), adesc + (uint64_t)(2 * (i & 3)),
                                                bdesc + (uint64_t)(2 * (i & 3)), idesc, 1u);
    } else {
      // rotate over `nacc` accumulators of N columns each: separates the issue rate from the latency of a
      // dependent accumulate into the same TMEM columns
      int a = 0;
      for (int i = 0; i < iters; ++i) {
        umma_bf16(tmem_base + (uint32_t)(a * N), adesc + (uint64_t)(2 * (i & 3)), bdesc + (uint64_t)(2 * (i & 3)),
                  idesc, 1u);
        if (++a == nacc) a = 0;
      }
    }
    umma_commit(bar);
    mbar_wait(bar, 0);
    const long long t1 = clock64();
    out[0] = t1 - t0;
    out[1] = iters;
  }
  tc_fence_before();
  __syncthreads();
  if ((threadIdx.x >> 5) == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}
}  // namespace

extern "C" int vcm_umma_rate(int32_t N, int32_t iters, long long* out, void* stream) {
  return vcm_umma_rate2(N, iters, 0, 1024, out, stream);
}

extern "C" int vcm_umma_rate2(int32_t N, int32_t iters, int32_t start_row, int32_t sbo_bytes, long long* out,
                              void* stream) {
  VCM_CHECK_ARG(out && N >= 16 && N <= 256 && N % 16 == 0 && iters > 0, "bad argument");
  VCM_CHECK_ARG(start_row >= 0 && sbo_bytes % 128 == 0 && start_row * 128 + 15 * sbo_bytes + 1024 <= 384 * 128,
                "A tile outside the probe buffer");
  const int smem_bytes = (384 + 256) * 128 + 64 + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(umma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    attr_set = true;
  }
  umma_rate_kernel<<<1, 128, smem_bytes, (cudaStream_t)stream>>>(N, iters, start_row, sbo_bytes, 0, out);
  VCM_CHECK_LAUNCH();
  return 0;
}

extern "C" int vcm_umma_rate3(int32_t N, int32_t iters, int32_t nacc, long long* out, void* stream) {
  VCM_CHECK_ARG(out && N >= 16 && N <= 256 && N % 16 == 0 && iters > 0 && nacc >= 1 && nacc * N <= 512,
                "bad argument");
  const int smem_bytes = (384 + 256) * 128 + 64 + 1024;
  cudaFuncSetAttribute(umma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  umma_rate_kernel<<<1, 128, smem_bytes, (cudaStream_t)stream>>>(N, iters, 0, 1024, nacc, out);
  VCM_CHECK_LAUNCH();
  return 0;
}

// ---------------------------------------------------------------------------------- TMA box probe
// Every CTA streams `iters` boxes {64 ch, bw, bh} (128B swizzle) of an NHWC bf16 image through a
// `stages`-deep shared-memory ring, walking the image like the conv engines do (CTA-interleaved
// tiles).  nsplit > 1 issues each box as nsplit TMA operations of bh/nsplit rows.  out[0] = cycles of
// CTA 0, out[1] = boxes it loaded.
#include "../tc_host.cuh"
namespace {
__global__ void __launch_bounds__(64, 1)
tma_probe_kernel(const __grid_constant__ CUtensorMap tm, int bw, int bh, int tiles_x, int n_tiles, int stages,
                 int stage_bytes, int nsplit, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)stages * stage_bytes);
  uint64_t* empty = full + 16;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    fence_barrier_init();
  }
  __syncthreads();
  const long long t0 = clock64();
  const int sub = bh / nsplit;
  if (threadIdx.x == 0) {
    int stage = 0; uint32_t phase = 0;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
      const int y0 = (t / tiles_x) * bh, x0 = (t % tiles_x) * bw;
      mbar_wait(&empty[stage], phase ^ 1u);
      mbar_expect_tx(&full[stage], (uint32_t)(bw * bh * 128));
      for (int s = 0; s < nsplit; ++s)
        tma_load_4d(smem_u32(smem + (size_t)stage * stage_bytes + (size_t)s * sub * bw * 128), &tm, &full[stage], 0,
                    x0, y0 + s * sub, 0);
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  } else if (threadIdx.x == 32) {
    int stage = 0; uint32_t phase = 0; long long n = 0;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++n) {
      mbar_wait(&full[stage], phase);
      mbar_arrive(&empty[stage]);
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
    if (blockIdx.x == 0) { out[0] = clock64() - t0; out[1] = n; }
  }
}
}  // namespace

extern "C" int vcm_tma_probe(const void* img, int32_t H, int32_t W, int32_t bw, int32_t bh, int32_t stages,
                             int32_t nsplit, long long* out, void* stream) {
  EncodeTiledFn enc = get_encode_fn();
  VCM_CHECK_ARG(enc && img && out && bh % nsplit == 0 && stages >= 1 && stages <= 16, "bad argument");
  CUtensorMap tm;
  cuuint64_t dims[4] = {64, (cuuint64_t)W, (cuuint64_t)H, 1};
  cuuint64_t strides[3] = {128, (cuuint64_t)W * 128, (cuuint64_t)H * W * 128};
  cuuint32_t box[4] = {64, (cuuint32_t)bw, (cuuint32_t)(bh / nsplit), 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(img), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  VCM_CHECK_ARG(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed");
  const int stage_bytes = round_up(bw * bh * 128, 1024);
  const int smem_bytes = stages * stage_bytes + 1024 + 512;
  VCM_CHECK_ARG(smem_bytes <= SMEM_LIMIT, "ring does not fit shared memory");
  cudaFuncSetAttribute(tma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  const int tiles_x = W / bw, n_tiles = tiles_x * (H / bh);
  tma_probe_kernel<<<num_sms(), 64, smem_bytes, (cudaStream_t)stream>>>(tm, bw, bh, tiles_x, n_tiles, stages, stage_bytes,
                                                                       nsplit, out);
  VCM_CHECK_LAUNCH();
  return 0;
}

// ---------------------------------------------------------------------------------- accumulator rounding
namespace {
__global__ void __launch_bounds__(128, 1) umma_acc_probe_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                                float* __restrict__ out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;                      // 128 x 128 B
  uint8_t* sB = smem + 128 * 128;          // 64 x 128 B
  uint64_t* bar = reinterpret_cast<uint64_t*>(sB + 64 * 128);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
  for (int i = threadIdx.x; i < 128 * 64; i += blockDim.x) {
    const int p = i / 64, c = i % 64;
    const uint32_t off = p * 128 + (((c >> 3) ^ (p & 7)) << 4) + (c & 7) * 2;
    *reinterpret_cast<__nv_bfloat16*>(sA + off) = __float2bfloat16_rn(a[i]);
  }
  for (int i = threadIdx.x; i < 64 * 64; i += blockDim.x) {
    const int nrow = i / 64, k = i % 64;
    const uint32_t off = nrow * 128 + (((k >> 3) ^ (nrow & 7)) << 4) + (k & 7) * 2;
    *reinterpret_cast<__nv_bfloat16*>(sB + off) = __float2bfloat16_rn(b[k]);
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
    const uint64_t adesc = make_smem_desc(smem_u32(sA), 1024u, 2u);
    const uint64_t bdesc = make_smem_desc(smem_u32(sB), 1024u, 2u);
    const uint32_t idesc = make_idesc_bf16(128, 64);
#pragma unroll
    for (int k = 0; k < 4; ++k) umma_bf16(tmem_base, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, k > 0);
    umma_commit(bar);
  }
  mbar_wait(bar, 0);
  tc_fence_after();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t r[16];
  tmem_ld16(tmem_base + ((uint32_t)(warp * 32) << 16), r);
  out[warp * 32 + lane] = __uint_as_float(r[0]);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 64);
  }
}
}  // namespace

extern "C" int vcm_umma_acc_probe(const float* a, const float* b, float* out, void* stream) {
  VCM_CHECK_ARG(a && b && out, "null argument");
  const int smem_bytes = 128 * 128 + 64 * 128 + 64 + 1024;
  umma_acc_probe_kernel<<<1, 128, smem_bytes, (cudaStream_t)stream>>>(a, b, out);
  VCM_CHECK_LAUNCH();
  return 0;
}

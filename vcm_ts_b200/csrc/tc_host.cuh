// vcm_ts_b200/csrc/tc_host.cuh -- host-side helpers shared by the tcgen05 conv engines:
// driver entry point for cuTensorMapEncodeTiled, K-chunk / swizzle / N-tiling choices.
#pragma once
#include <cuda.h>
#include <mutex>

#include "common.cuh"

namespace tc {

// ---------------------------------------------------------------------------------- host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

inline EncodeTiledFn get_encode_fn() {
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

inline int pick_chunk(int cin) { return cin % 64 == 0 ? 64 : (cin % 32 == 0 ? 32 : 16); }
// split-bf16 parity mode: a stage holds three planes, so 7x7 halo tiles use 32-channel chunks at most
inline int pick_chunk_split(int cin, int K) {
  const int c = pick_chunk(cin);
  return (K == 7 && c > 32) ? 32 : c;
}
inline CUtensorMapSwizzle swizzle_of(int chunk) {
  return chunk == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : (chunk == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
}
inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

// N tiling: Cout padded to 16; above 256 split into the fewest equal tiles that are multiples of 16
inline void pick_n_tiling(int cout, int& cout_p, int& bn, int& n_tiles, int max_bn = 256) {
  cout_p = round_up(cout, 16);
  n_tiles = 1;
  while (cout_p / n_tiles > max_bn || cout_p % n_tiles || (cout_p / n_tiles) % 16) ++n_tiles;
  bn = cout_p / n_tiles;
}

// per-device state (a process may drive several GPUs; writes are idempotent, so unsynchronised
// updates from the worker threads are harmless)
constexpr int MAX_DEVICES = 64;
inline int current_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev >= 0 && dev < MAX_DEVICES ? dev : 0;
}
inline int num_sms() {
  static int n[MAX_DEVICES] = {0};
  const int dev = current_device();
  if (!n[dev]) {
    int v = 0;
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    n[dev] = v > 0 ? v : 148;
  }
  return n[dev];
}
// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device attribute of the function
template <class Kernel>
inline cudaError_t ensure_max_smem(Kernel k, bool (&done)[MAX_DEVICES], int bytes) {
  const int dev = current_device();
  if (done[dev]) return cudaSuccess;
  const cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) done[dev] = true;
  return e;
}

constexpr int SMEM_LIMIT = 227 * 1024;

}  // namespace tc

#include "conv_common.cuh"

namespace tc {

// 0 = vector epilogue (bf16 out, Cout % 16 == 0, every view 16-byte aligned), 1 = scalar bf16,
// 2 = scalar fp32 (tc_epilogue.cuh)
inline int tc_epilogue_mode(const vcm_conv_desc* d, const ConvParams& P) {
  if (d->out_dtype == VCM_F32) return 2;
  auto al = [](const void* p) { return ((uintptr_t)p & 15) == 0; };
  bool vec = d->Cout % 16 == 0 && d->out_cstride % 8 == 0 && al(P.y) && al(P.bias);
  if (d->flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE)) vec = vec && d->res_cstride % 8 == 0 && al(P.res);
  if (d->flags & VCM_CONV_RES_SCALE) vec = vec && al(P.res_scale);
  if (d->flags & VCM_CONV_RES2) vec = vec && d->res2_cstride % 8 == 0 && al(P.res2);
  if (d->flags & VCM_CONV_OUT2) vec = vec && d->out2_cstride % 8 == 0 && al(P.y2);
  if (d->flags & VCM_CONV_PIXEL_SHUFFLE) vec = vec && (d->Cout / 4) % 16 == 0;
  return vec ? 0 : 1;
}
// channels per epilogue block: 32 when the N tile (and a PixelShuffle position group) is a multiple
inline int tc_epilogue_bw(const vcm_conv_desc* d, int bn) {
  if (bn % 32) return 16;
  if ((d->flags & VCM_CONV_PIXEL_SHUFFLE) && (d->Cout / 4) % 32) return 16;
  return 32;
}

}  // namespace tc

// stride-1 halo engine (conv_halo.cu); returns 0 ok, 1 error, 3 = shape not taken (use per-tap)
struct ConvParams;
// se_rows != nullptr: query only -- *se_rows = rows of channel sums a VCM_CONV_SE_SUM launch of this shape would
// write (0 = it cannot), nothing is launched
int vcm_conv_halo_launch(const vcm_conv_desc* d, const ConvParams& P, const void* x, const void* w_umma,
                         cudaStream_t stream, int* se_rows = nullptr);
// row-stacked 3x3 engine for Cout = 64 / 32 (conv_rows.cu); same return convention
int vcm_conv_rows_launch(const vcm_conv_desc* d, const ConvParams& P, const void* x, const void* w_umma,
                         cudaStream_t stream, int* se_rows = nullptr);

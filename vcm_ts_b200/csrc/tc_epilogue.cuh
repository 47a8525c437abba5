// vcm_ts_b200/csrc/tc_epilogue.cuh -- TMEM -> registers -> fused conv epilogue, shared by the two
// tcgen05 kernels (conv_tc.cu: per-tap TMA, conv_halo.cu: halo tile + resident weights).
#pragma once
#include "conv_common.cuh"
#include "tc_ptx.cuh"

namespace tc {

// ---------------------------------------------------------------------------------- epilogue
using bf = __nv_bfloat16;

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

// Address (in bf16 elements from a view base) of the 4-channel unit `u` of the BW-wide channel
// block starting at c0, for this thread's pixel.  Non-PS: unit u = channels c0+4u..c0+4u+3 of pixel
// (oy,ox).  PS: units are grouped by sub-pixel position pos = u / (BW/16): 4 consecutive output
// channels oc0 + 4*(u % (BW/16)) .. of pixel (2oy+dy, 2ox+dx).
template <int BW>
struct UnitAddr {
  size_t pix[4];   // pixel index per sub-position (non-PS uses [0])
  int ch0;         // first output channel of the block (oc0 for PS)
  bool ps;
  __device__ __forceinline__ void init(const ConvParams& P, int n, int oy, int ox, int c0) {
    ps = P.flags & VCM_CONV_PIXEL_SHUFFLE;
    if (!ps) {
      pix[0] = ((size_t)n * P.Ho + oy) * P.Wo + ox;
      ch0 = c0;
    } else {
#pragma unroll
      for (int pos = 0; pos < 4; ++pos)
        pix[pos] = ((size_t)n * (2 * P.Ho) + 2 * oy + (pos >> 1)) * (2 * P.Wo) + 2 * ox + (pos & 1);
      ch0 = c0 >> 2;
    }
  }
  __device__ __forceinline__ size_t off(int u, int cs) const {
    if (!ps) return pix[0] * cs + ch0 + 4 * u;
    constexpr int UPP = BW / 16;   // units per position
    return pix[u / UPP] * cs + ch0 + 4 * (u % UPP);
  }
};

// Loads the residual operands of one BW block (issued BEFORE the accumulator is awaited).
template <int BW>
__device__ __forceinline__ void epi_prefetch(const ConvParams& P, const UnitAddr<BW>& A, uint2 (&r1)[BW / 4],
                                             uint2 (&r2)[BW / 4]) {
  if (P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE)) {
#pragma unroll
    for (int u = 0; u < BW / 4; ++u)
      r1[u] = __ldg(reinterpret_cast<const uint2*>(reinterpret_cast<const bf*>(P.res) + A.off(u, P.res_cs)));
  }
  if (P.flags & VCM_CONV_RES2) {
#pragma unroll
    for (int u = 0; u < BW / 4; ++u)
      r2[u] = __ldg(reinterpret_cast<const uint2*>(reinterpret_cast<const bf*>(P.res2) + A.off(u, P.res2_cs)));
  }
}

// v[BW]: accumulators of channels c0..c0+BW-1 (GEMM column order).  Same op order as
// conv_epilogue_store: bias -> act -> (+res | + res*scale) -> (+res2) -> store (-> twin).
template <int BW>
__device__ __forceinline__ void epi_finish(const ConvParams& P, const UnitAddr<BW>& A, int n, int c0, float* v,
                                           const uint2 (&r1)[BW / 4], const uint2 (&r2)[BW / 4]) {
#pragma unroll
  for (int q = 0; q < BW / 4; ++q) {
    const float4 b = __ldg(reinterpret_cast<const float4*>(P.bias + c0) + q);
    v[4 * q] += b.x; v[4 * q + 1] += b.y; v[4 * q + 2] += b.z; v[4 * q + 3] += b.w;
  }
  if (P.flags & VCM_CONV_POST_ACT) {
#pragma unroll
    for (int i = 0; i < BW; ++i) v[i] = leaky(v[i], P.post_slope);
  }
  constexpr int UPP = BW / 16;
#pragma unroll
  for (int u = 0; u < BW / 4; ++u) {
    float w[4];
    int chan;   // first output channel of this unit (for the SE scale lookup)
    if (!A.ps) {
#pragma unroll
      for (int i = 0; i < 4; ++i) w[i] = v[4 * u + i];
      chan = c0 + 4 * u;
    } else {
      const int pos = u / UPP, k0 = 4 * (u % UPP);   // oc = ch0 + k0 + i  <-  co = c0 + 4*(k0+i) + pos
#pragma unroll
      for (int i = 0; i < 4; ++i) w[i] = v[4 * (k0 + i) + pos];
      chan = A.ch0 + k0;
    }
    if (P.flags & (VCM_CONV_RES | VCM_CONV_RES_SCALE)) {
      float r[4];
      unpack4(r1[u], r);
      if (P.flags & VCM_CONV_RES) {
#pragma unroll
        for (int i = 0; i < 4; ++i) w[i] = r[i] + w[i];
      } else {
        const int Cres = A.ps ? (P.Cout >> 2) : P.Cout;
        const float4 s = __ldg(reinterpret_cast<const float4*>(P.res_scale + (size_t)n * Cres + chan));
        w[0] = w[0] + __fmul_rn(r[0], s.x); w[1] = w[1] + __fmul_rn(r[1], s.y);
        w[2] = w[2] + __fmul_rn(r[2], s.z); w[3] = w[3] + __fmul_rn(r[3], s.w);
      }
    }
    if (P.flags & VCM_CONV_RES2) {
      float r[4];
      unpack4(r2[u], r);
#pragma unroll
      for (int i = 0; i < 4; ++i) w[i] = r[i] + w[i];
    }
    *reinterpret_cast<uint2*>(reinterpret_cast<bf*>(P.y) + A.off(u, P.out_cs)) = pack4(w);
    if (P.flags & VCM_CONV_OUT2) {
#pragma unroll
      for (int i = 0; i < 4; ++i) w[i] = leaky(w[i], P.out2_slope);
      *reinterpret_cast<uint2*>(reinterpret_cast<bf*>(P.y2) + A.off(u, P.out2_cs)) = pack4(w);
    }
  }
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

template <int BW, class GEO>
__device__ __forceinline__ void epilogue_tile(const ConvParams& P, const GEO& G, int n, int oy, int ox,
                                              bool valid, int cbase, uint32_t taddr, uint64_t* tfull,
                                              uint32_t parity) {
  // block 0 residuals go out before we wait for the MMAs of this tile
  UnitAddr<BW> A;
  uint2 r1[BW / 4], r2[BW / 4];
  const bool vec = !G.scalar_epi;
  if (valid && vec) {
    A.init(P, n, oy, ox, cbase);
    epi_prefetch<BW>(P, A, r1, r2);
  }
  mbar_wait(tfull, parity);
  tc_fence_after();
  for (int cb = 0; cb < G.bn; cb += BW) {
    float v[BW];
    tmem_load_block<BW>(taddr + (uint32_t)cb, v);
    if (!valid) continue;
    if (vec) {
      UnitAddr<BW> An;
      uint2 n1[BW / 4], n2[BW / 4];
      const bool more = cb + BW < G.bn;
      if (more) {   // next block's residuals in flight while this block is finished
        An.init(P, n, oy, ox, cbase + cb + BW);
        epi_prefetch<BW>(P, An, n1, n2);
      }
      epi_finish<BW>(P, A, n, cbase + cb, v, r1, r2);
      if (more) {
        A = An;
#pragma unroll
        for (int u = 0; u < BW / 4; ++u) { r1[u] = n1[u]; r2[u] = n2[u]; }
      }
    } else {
#pragma unroll
      for (int i = 0; i < BW; ++i) {
        const int co = cbase + cb + i;
        if (co < P.Cout) {
          if (G.scalar_epi == 2) conv_epilogue_store<float>(P, n, oy, ox, co, v[i]);
          else conv_epilogue_store<bf>(P, n, oy, ox, co, v[i]);
        }
      }
    }
  }
}


}  // namespace tc

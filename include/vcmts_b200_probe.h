/*
 * include/vcmts_b200_probe.h -- C-ABI of libvcmts_probe.so: hardware probes used while designing the
 * tcgen05 conv engines (UMMA descriptor semantics, MMA issue rate, TMA box throughput, accumulator
 * rounding).  Diagnostics only: NOT part of the product library (libvcmts_b200.so) and not an interface
 * of the reference.  Built by vcm_ts_b200/build.py from vcm_ts_b200/csrc/probe/.
 */
#ifndef VCMTS_B200_PROBE_H
#define VCMTS_B200_PROBE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
/* probe of the UMMA shared-memory descriptor semantics used by the halo-tile variant: runs
 * one 128x64x64 MMA whose A rows start `start_row` rows into a 128B-swizzled buffer with the
 * given stride between 8-row groups; out[128*64] fp32 = the A rows the tensor core read.     */
int vcm_umma_probe(int32_t start_row, int32_t sbo_bytes, int32_t use_base_offset, float* out,
                   void* stream);
/* issue-rate probe: `iters` back-to-back 128xNx16 bf16 MMAs; out[0] = SM cycles, out[1] = iters */
int vcm_umma_rate(int32_t N, int32_t iters, long long* out, void* stream);
/* same, with the A operand starting `start_row` 128-byte rows into the buffer and 8-row groups sbo_bytes apart */
int vcm_umma_rate2(int32_t N, int32_t iters, int32_t start_row, int32_t sbo_bytes, long long* out,
                   void* stream);
/* same, rotating over `nacc` accumulators of N TMEM columns each (nacc * N <= 512): issue rate vs the latency
 * of a dependent accumulate */
int vcm_umma_rate3(int32_t N, int32_t iters, int32_t nacc, long long* out, void* stream);
/* TMA box-shape throughput probe: all SMs stream {64 ch, bw, bh} boxes of an NHWC bf16 [H][W][64] image
 * through a `stages`-deep ring (each box as nsplit TMA ops); out[0] = cycles of CTA 0, out[1] = its boxes */
int vcm_tma_probe(const void* img, int32_t H, int32_t W, int32_t bw, int32_t bh, int32_t stages,
                  int32_t nsplit, long long* out, void* stream);

/* accumulator-rounding probe: one 128x64x64 bf16 MMA chain (4 instructions of K = 16, accumulating
 * in TMEM) with A[m][k] = bf16(a[m*64 + k]) and B[n][k] = bf16(b[k]) for every n, so
 * out[m] = sum_k a[m][k] * b[k] as the tensor core computes it: values placed in one 16-wide k group
 * are added inside one instruction, values in different groups by the fp32 accumulate.           */
int vcm_umma_acc_probe(const float* a, const float* b, float* out, void* stream);
#ifdef __cplusplus
}
#endif
#endif

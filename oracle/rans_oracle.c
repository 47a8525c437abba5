/*
 * oracle/rans_oracle.c -- TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C restatement of the reference's native entropy coder, used only by
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg as the
 * checker for the product coder (vcm_ts_b200/csrc/rans_coder.cpp).
 *
 * Follows:
 *   BufferedRansEncoder::encode_with_indexes  DCVC_HEM/src/cpp/rans/rans_interface.cpp:85-145
 *   BufferedRansEncoder::flush                rans_interface.cpp:147-172
 *   RansDecoder::set_stream / decode_stream   rans_interface.cpp:176-244
 *   Rans64EncPutBits / Rans64DecGetBits       rans_interface.cpp:46-82
 *   pmf_to_quantized_cdf                      DCVC_HEM/src/cpp/ops/ops.cpp:24-82
 *   rans64.h (ryg_rans c9d162d, un-vendored)  restated in oracle/rans64.h
 *
 * Parity: pinned against oracle/_ref (the reference C++ compiled unmodified)
 * and the big-int model oracle/rans_py.py by tests/test_rans_oracle.py.  The
 * reference has no golden bitstreams of its own ("parity unpinned" upstream).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define Rans64Assert(x) ((void)0)
#include "rans64.h"

enum { PRECISION = 16, BYPASS_BITS = 4, BYPASS_MAX = (1 << BYPASS_BITS) - 1 };

typedef struct {
  uint16_t start, range;
  uint8_t bypass;
} osym_t;

typedef struct {
  osym_t *syms;
  size_t n, cap;
  uint32_t *out;
  size_t out_words;
} oenc_t;

typedef struct {
  uint32_t *words;
  size_t nwords;
  uint32_t *ptr;
  Rans64State state;
} odec_t;

static void push(oenc_t *e, uint16_t start, uint16_t range, uint8_t bypass) {
  if (e->n == e->cap) {
    e->cap = e->cap ? e->cap * 2 : 1024;
    e->syms = (osym_t *)realloc(e->syms, e->cap * sizeof(osym_t));
  }
  e->syms[e->n].start = start;
  e->syms[e->n].range = range;
  e->syms[e->n].bypass = bypass;
  e->n++;
}

void *oracle_enc_new(void) { return calloc(1, sizeof(oenc_t)); }
void oracle_enc_reset(void *h) { ((oenc_t *)h)->n = 0; }
void oracle_enc_free(void *h) {
  oenc_t *e = (oenc_t *)h;
  free(e->syms);
  free(e->out);
  free(e);
}

/* rans_interface.cpp:85-145 -- symbols are mapped to [0, max_value] through the
 * per-CDF offset; out-of-range values become the sentinel max_value followed
 * by an Exp-Golomb-like 4-bit bypass sequence. */
void oracle_enc_push(void *h, const int32_t *symbols, const int32_t *indexes,
                     int64_t n, const int32_t *cdfs, int32_t cdf_stride,
                     const int32_t *cdf_sizes, const int32_t *offsets) {
  oenc_t *e = (oenc_t *)h;
  for (int64_t i = 0; i < n; ++i) {
    const int32_t ci = indexes[i];
    const int32_t *cdf = cdfs + (int64_t)ci * cdf_stride;
    const int32_t max_value = cdf_sizes[ci] - 2;
    int32_t value = symbols[i] - offsets[ci];
    uint32_t raw = 0;
    if (value < 0) {
      raw = (uint32_t)(-2 * value - 1);
      value = max_value;
    } else if (value >= max_value) {
      raw = (uint32_t)(2 * (value - max_value));
      value = max_value;
    }
    push(e, (uint16_t)cdf[value], (uint16_t)(cdf[value + 1] - cdf[value]), 0);
    if (value == max_value) {
      int32_t nb = 0;
      while ((raw >> (nb * BYPASS_BITS)) != 0) ++nb;
      int32_t v = nb;
      while (v >= BYPASS_MAX) {
        push(e, BYPASS_MAX, BYPASS_MAX + 1, 1);
        v -= BYPASS_MAX;
      }
      push(e, (uint16_t)v, (uint16_t)(v + 1), 1);
      for (int32_t j = 0; j < nb; ++j) {
        const int32_t d = (raw >> (j * BYPASS_BITS)) & BYPASS_MAX;
        push(e, (uint16_t)d, (uint16_t)(d + 1), 1);
      }
    }
  }
}

/* rans_interface.cpp:46-64 */
static void put_bits(Rans64State *r, uint32_t **pptr, uint32_t val, uint32_t nbits) {
  uint64_t x = *r;
  const uint32_t freq = 1u << (16 - nbits);
  const uint64_t limit = ((RANS64_L >> 16) << 32) * (uint64_t)freq;
  if (x >= limit) {
    *pptr -= 1;
    **pptr = (uint32_t)x;
    x >>= 32;
  }
  *r = (x << nbits) | val;
}

/* rans_interface.cpp:66-82 */
static uint32_t get_bits(Rans64State *r, uint32_t **pptr, uint32_t nbits) {
  uint64_t x = *r;
  const uint32_t val = (uint32_t)(x & ((1u << nbits) - 1));
  x >>= nbits;
  if (x < RANS64_L) {
    x = (x << 32) | (uint64_t)(**pptr);
    *pptr += 1;
  }
  *r = x;
  return val;
}

/* rans_interface.cpp:147-172 -- symbols are emitted last-to-first; returns the
 * byte count and a pointer valid until the next flush/free. */
int64_t oracle_enc_flush(void *h, const uint8_t **bytes) {
  oenc_t *e = (oenc_t *)h;
  free(e->out);
  e->out_words = e->n + 2;
  e->out = (uint32_t *)malloc(e->out_words * sizeof(uint32_t));
  uint32_t *end = e->out + e->out_words;
  uint32_t *ptr = end;
  Rans64State r;
  Rans64EncInit(&r);
  for (size_t i = e->n; i-- > 0;) {
    const osym_t s = e->syms[i];
    if (!s.bypass)
      Rans64EncPut(&r, &ptr, s.start, s.range, PRECISION);
    else
      put_bits(&r, &ptr, s.start, BYPASS_BITS);
  }
  e->n = 0;
  Rans64EncFlush(&r, &ptr);
  *bytes = (const uint8_t *)ptr;
  return (int64_t)((end - ptr) * sizeof(uint32_t));
}

void *oracle_dec_new(void) { return calloc(1, sizeof(odec_t)); }
void oracle_dec_free(void *h) {
  odec_t *d = (odec_t *)h;
  free(d->words);
  free(d);
}

/* rans_interface.cpp:176-182 */
void oracle_dec_set_stream(void *h, const uint8_t *bytes, int64_t nbytes) {
  odec_t *d = (odec_t *)h;
  free(d->words);
  d->nwords = (size_t)(nbytes / 4);
  d->words = (uint32_t *)malloc((d->nwords + 4) * sizeof(uint32_t));
  memset(d->words, 0, (d->nwords + 4) * sizeof(uint32_t));
  memcpy(d->words, bytes, (size_t)nbytes);
  d->ptr = d->words;
  Rans64DecInit(&d->state, &d->ptr);
}

/* rans_interface.cpp:184-244 -- linear CDF search, sentinel -> bypass decode. */
void oracle_dec_decode(void *h, const int32_t *indexes, int64_t n,
                       const int32_t *cdfs, int32_t cdf_stride,
                       const int32_t *cdf_sizes, const int32_t *offsets,
                       int32_t *out) {
  odec_t *d = (odec_t *)h;
  for (int64_t i = 0; i < n; ++i) {
    const int32_t ci = indexes[i];
    const int32_t *cdf = cdfs + (int64_t)ci * cdf_stride;
    const int32_t size = cdf_sizes[ci];
    const int32_t max_value = size - 2;
    const uint32_t cum = Rans64DecGet(&d->state, PRECISION);
    int32_t k = 0;
    while (k < size && !((uint32_t)cdf[k] > cum)) ++k;
    const int32_t s = k - 1;
    Rans64DecAdvance(&d->state, &d->ptr, (uint32_t)cdf[s],
                     (uint32_t)(cdf[s + 1] - cdf[s]), PRECISION);
    int32_t value = s;
    if (value == max_value) {
      int32_t v = (int32_t)get_bits(&d->state, &d->ptr, BYPASS_BITS);
      int32_t nb = v;
      while (v == BYPASS_MAX) {
        v = (int32_t)get_bits(&d->state, &d->ptr, BYPASS_BITS);
        nb += v;
      }
      int32_t raw = 0;
      for (int32_t j = 0; j < nb; ++j) {
        v = (int32_t)get_bits(&d->state, &d->ptr, BYPASS_BITS);
        raw |= v << (j * BYPASS_BITS);
      }
      value = raw >> 1;
      if (raw & 1)
        value = -value - 1;
      else
        value += max_value;
    }
    out[i] = value + offsets[ci];
  }
}

/* ops.cpp:24-82 -- 16-bit CDF quantisation with frequency stealing.
 * cdf_out has n+1 entries.  Returns 0 on success, -1 if no donor was found. */
int oracle_pmf_to_quantized_cdf(const float *pmf, int n, int precision, uint32_t *cdf) {
  const int m = n + 1;
  cdf[0] = 0;
  for (int i = 0; i < n; ++i)
    cdf[i + 1] = (uint32_t)(roundf(pmf[i] * (float)(1 << precision)) + 0.5f);
  uint32_t total = 0;
  for (int i = 0; i < m; ++i) total += cdf[i];
  for (int i = 0; i < m; ++i)
    cdf[i] = (uint32_t)((((uint64_t)1 << precision) * cdf[i]) / total);
  for (int i = 1; i < m; ++i) cdf[i] += cdf[i - 1];
  cdf[m - 1] = 1u << precision;
  for (int i = 0; i < m - 1; ++i) {
    if (cdf[i] == cdf[i + 1]) {
      uint32_t best_freq = ~0u;
      int best = -1;
      for (int j = 0; j < m - 1; ++j) {
        const uint32_t f = cdf[j + 1] - cdf[j];
        if (f > 1 && f < best_freq) {
          best_freq = f;
          best = j;
        }
      }
      if (best < 0) return -1;
      if (best < i) {
        for (int j = best + 1; j <= i; ++j) cdf[j]--;
      } else {
        for (int j = i + 1; j <= best; ++j) cdf[j]++;
      }
    }
  }
  return 0;
}

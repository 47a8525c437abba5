// vcm_ts_b200/csrc/rans_coder.cpp -- host entropy coder of libvcmts_b200.so (CPU, GIL-free).
//
// Produces and consumes exactly the reference's bitstream:
//   BufferedRansEncoder::encode_with_indexes / flush   DCVC_HEM/src/cpp/rans/rans_interface.cpp:85-172
//   RansDecoder::set_stream / decode_stream            rans_interface.cpp:176-244
//   bypass put/get bits                                rans_interface.cpp:46-82
//   pmf_to_quantized_cdf                               DCVC_HEM/src/cpp/ops/ops.cpp:24-82
// on top of the 64-bit rANS of ryg_rans `rans64.h` (state in [2^31, 2^63), 32-bit renormalisation
// words written backwards, 16-bit probabilities, 4-bit bypass digits for out-of-table values).
//
// What is different from the reference is only the plumbing: symbols arrive as int16 and table
// indexes as uint8 straight from pinned buffers the GPU epilogue wrote (no int32 numpy round
// trip, no Python list), the encoder replaces the per-symbol division by a precomputed round-up
// reciprocal (exact for every state < 2^63), the decoder tries the most probable value first and else finds the symbol with a
// 256-bucket lookup + short scan instead of a linear search from 0, and nothing touches Python.
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

#include <vector>

#include "../../include/vcmts_b200.h"

static thread_local char g_err[512] = "";

void vcm_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* vcm_last_error(void) { return g_err; }
extern "C" int vcm_version(void) { return 100; }

namespace {
constexpr uint32_t kPrecision = 16;
constexpr uint32_t kBypassBits = 4;
constexpr int32_t kBypassMax = (1 << kBypassBits) - 1;
constexpr uint64_t kRansL = 1ull << 31;
}  // namespace

// Encoder record of one (table, value): x -> C(s, x) without a division.  With l = ceil(log2 freq)
// and rcp = ceil(2^(63+l) / freq) (< 2^64), floor(x / freq) = mulhi(x, rcp) >> (l - 1) for every
// x < 2^63 (Granlund-Montgomery / Alverson round-up reciprocal, the construction ryg_rans'
// Rans64EncSymbolInit uses), so
//     ((x / freq) << 16) + x % freq + start  ==  x + start + q * (2^16 - freq),  q = floor(x / freq).
// freq == 1: rcp = 2^64 - 1, shift 0 gives q = x - 1, compensated in `bias`.
struct EncSym {
  uint64_t rcp;
  uint32_t bias;       // start (+ 2^16 - 1 when freq == 1)
  uint16_t freq;       // 1 .. 65535 (a 2-symbol table cannot hold 65536)
  uint16_t cmpl;       // 2^16 - freq
  uint32_t shift;
  uint32_t pad;
};

struct vcm_cdf_tables {
  int32_t ntab = 0, stride = 0;
  std::vector<uint32_t> cdf;      // [ntab][stride]
  std::vector<int32_t> sizes, offsets;
  std::vector<uint8_t> lut;       // [ntab][256]: first symbol whose upper bound exceeds bucket<<8
  std::vector<EncSym> enc;        // [ntab][stride]
  std::vector<uint32_t> mode;     // [ntab][4]: most probable value, its start, its freq, 0
};

struct vcm_rans_enc {
  // address of the symbol's EncSym (8-byte aligned, bit 0 clear) or (4-bit bypass digit << 1) | 1.
  // The tables pushed from must stay alive until the flush (the models keep them for their lifetime).
  std::vector<uint64_t> pending;
  std::vector<uint32_t> out;
};

struct vcm_rans_dec {
  std::vector<uint32_t> words;
  size_t pos = 0;
  uint64_t state = 0;
};

extern "C" vcm_cdf_tables* vcm_cdf_tables_new(const int32_t* cdf, const int32_t* sizes,
                                              const int32_t* offsets, int32_t ntab, int32_t stride) {
  if (!cdf || !sizes || !offsets || ntab <= 0 || ntab > 256 || stride < 3) {
    vcm_set_error("vcm_cdf_tables_new: bad argument (ntab must be 1..256)");
    return nullptr;
  }
  auto* t = new vcm_cdf_tables;
  t->ntab = ntab;
  t->stride = stride;
  t->cdf.resize((size_t)ntab * stride);
  t->sizes.assign(sizes, sizes + ntab);
  t->offsets.assign(offsets, offsets + ntab);
  t->lut.resize((size_t)ntab * 256);
  t->enc.assign((size_t)ntab * stride, EncSym{});
  t->mode.assign((size_t)ntab * 4, 0u);
  for (int32_t i = 0; i < ntab; ++i) {
    const int32_t sz = sizes[i];
    if (sz < 3 || sz > stride) {
      vcm_set_error("vcm_cdf_tables_new: table %d has size %d (stride %d)", i, sz, stride);
      delete t;
      return nullptr;
    }
    uint32_t* row = &t->cdf[(size_t)i * stride];
    for (int32_t k = 0; k < stride; ++k) row[k] = (uint32_t)cdf[(size_t)i * stride + k];
    // the decoder's scan `while (cdf[s + 1] <= cum)` terminates inside the row only if the row spans
    // [0, 2^16] exactly (ops.cpp:75-79 asserts the same of the tables it builds)
    if (row[0] != 0u || row[sz - 1] != (1u << kPrecision)) {
      vcm_set_error("vcm_cdf_tables_new: table %d does not span [0, 65536] (first %u, last %u)", i, row[0],
                    row[sz - 1]);
      delete t;
      return nullptr;
    }
    int32_t s = 0;
    for (int32_t b = 0; b < 256; ++b) {
      const uint32_t cum = (uint32_t)b << 8;
      while (s + 1 < sz - 1 && row[s + 1] <= cum) ++s;
      t->lut[(size_t)i * 256 + b] = (uint8_t)s;
    }
    uint32_t best = 0;
    for (int32_t v = 0; v < sz - 1; ++v) {
      const uint32_t start = row[v], freq = row[v + 1] - row[v];
      if (row[v + 1] <= row[v] || freq > 65535u || row[v + 1] > (1u << kPrecision)) {
        vcm_set_error("vcm_cdf_tables_new: table %d is not a strictly increasing 16-bit cdf", i);
        delete t;
        return nullptr;
      }
      EncSym& e = t->enc[(size_t)i * stride + v];
      e.freq = (uint16_t)freq;
      e.cmpl = (uint16_t)((1u << kPrecision) - freq);
      if (freq == 1) {
        e.rcp = ~0ull;
        e.shift = 0;
        e.bias = start + (1u << kPrecision) - 1;
      } else {
        uint32_t l = 1;
        while ((1u << l) < freq) ++l;
        e.rcp = (uint64_t)((((unsigned __int128)1 << (63 + l)) + freq - 1) / freq);
        e.shift = l - 1;
        e.bias = start;
      }
      if (v < sz - 2 && freq > best) {   // decoder fast path: most probable in-table value (never the escape)
        best = freq;
        uint32_t* m = &t->mode[(size_t)i * 4];
        m[0] = (uint32_t)v;
        m[1] = start;
        m[2] = freq;
      }
    }
  }
  return t;
}
extern "C" void vcm_cdf_tables_free(vcm_cdf_tables* t) { delete t; }

extern "C" vcm_rans_enc* vcm_rans_enc_new(void) { return new vcm_rans_enc; }
extern "C" void vcm_rans_enc_free(vcm_rans_enc* e) { delete e; }
extern "C" void vcm_rans_enc_reset(vcm_rans_enc* e) {
  if (e) e->pending.clear();
}

static inline void push_symbol(std::vector<uint64_t>& pend, const vcm_cdf_tables* t, int32_t ti,
                               int32_t symbol) {
  const int32_t max_value = t->sizes[ti] - 2;
  int32_t value = symbol - t->offsets[ti];
  uint32_t raw = 0;
  if (value < 0) {
    raw = (uint32_t)(-2 * value - 1);
    value = max_value;
  } else if (value >= max_value) {
    raw = (uint32_t)(2 * (value - max_value));
    value = max_value;
  }
  pend.push_back((uint64_t)(uintptr_t)&t->enc[(size_t)ti * t->stride + value]);
  if (value == max_value) {
    int32_t nb = 0;
    while ((raw >> (nb * kBypassBits)) != 0) ++nb;
    int32_t v = nb;
    while (v >= kBypassMax) {
      pend.push_back(((uint64_t)kBypassMax << 1) | 1u);
      v -= kBypassMax;
    }
    pend.push_back(((uint64_t)v << 1) | 1u);
    for (int32_t j = 0; j < nb; ++j) pend.push_back(((uint64_t)((raw >> (j * kBypassBits)) & kBypassMax) << 1) | 1u);
  }
}

extern "C" int vcm_rans_enc_push(vcm_rans_enc* e, const int16_t* sym, const uint8_t* idx, int64_t n,
                                 const vcm_cdf_tables* t) {
  if (!e || !sym || !idx || !t || n < 0) {
    vcm_set_error("vcm_rans_enc_push: bad argument");
    return 1;
  }
  e->pending.reserve(e->pending.size() + (size_t)n + 64);
  for (int64_t i = 0; i < n; ++i) {
    const int32_t ti = idx[i];
    if (ti >= t->ntab) {
      vcm_set_error("vcm_rans_enc_push: table index %d out of range", ti);
      return 1;
    }
    push_symbol(e->pending, t, ti, sym[i]);
  }
  return 0;
}

extern "C" int vcm_rans_enc_push_channels(vcm_rans_enc* e, const int16_t* sym, int32_t C, int64_t hw,
                                          const vcm_cdf_tables* t) {
  if (!e || !sym || !t || C <= 0 || hw < 0 || C > t->ntab) {
    vcm_set_error("vcm_rans_enc_push_channels: bad argument");
    return 1;
  }
  e->pending.reserve(e->pending.size() + (size_t)C * hw + 64);
  for (int32_t c = 0; c < C; ++c)
    for (int64_t i = 0; i < hw; ++i) push_symbol(e->pending, t, c, sym[(int64_t)c * hw + i]);
  return 0;
}

extern "C" int64_t vcm_rans_enc_flush(vcm_rans_enc* e, const uint8_t** bytes) {
  if (!e || !bytes) {
    vcm_set_error("vcm_rans_enc_flush: bad argument");
    return -1;
  }
  const size_t n = e->pending.size();
  e->out.resize(n + 2);
  uint32_t* const end = e->out.data() + e->out.size();
  uint32_t* ptr = end;
  uint64_t x = kRansL;
  const uint64_t* pend = e->pending.data();
  for (size_t i = n; i-- > 0;) {
    const uint64_t p = pend[i];
    if (!(p & 1u)) {
      const EncSym* s = reinterpret_cast<const EncSym*>((uintptr_t)p);
      if (i >= 8) __builtin_prefetch(reinterpret_cast<const void*>((uintptr_t)(pend[i - 8] & ~(uint64_t)1)));
      const uint64_t limit = (uint64_t)s->freq << (31 - kPrecision + 32);   // ((L >> 16) << 32) * freq
      if (x >= limit) {
        *--ptr = (uint32_t)x;
        x >>= 32;
      }
      const uint64_t q = (uint64_t)(((unsigned __int128)x * s->rcp) >> 64) >> s->shift;
      x = x + s->bias + q * s->cmpl;
    } else {
      const uint64_t limit = 1ull << (31 - kPrecision + 32 + 16 - kBypassBits);   // freq = 2^(16-4)
      if (x >= limit) {
        *--ptr = (uint32_t)x;
        x >>= 32;
      }
      x = (x << kBypassBits) | (p >> 1);
    }
  }
  ptr -= 2;
  ptr[0] = (uint32_t)(x & 0xffffffffu);
  ptr[1] = (uint32_t)(x >> 32);
  e->pending.clear();
  *bytes = reinterpret_cast<const uint8_t*>(ptr);
  return (int64_t)((end - ptr) * sizeof(uint32_t));
}

extern "C" vcm_rans_dec* vcm_rans_dec_new(void) { return new vcm_rans_dec; }
extern "C" void vcm_rans_dec_free(vcm_rans_dec* d) { delete d; }

extern "C" int vcm_rans_dec_set_stream(vcm_rans_dec* d, const uint8_t* bytes, int64_t nbytes) {
  if (!d || !bytes || nbytes < 8 || (nbytes & 3)) {
    vcm_set_error("vcm_rans_dec_set_stream: stream must be >= 8 bytes and a multiple of 4");
    return 1;
  }
  const size_t nw = (size_t)nbytes / 4;
  d->words.assign(nw + 8, 0u);  // zero tail: a truncated stream decodes garbage, never reads OOB
  memcpy(d->words.data(), bytes, (size_t)nbytes);
  d->state = (uint64_t)d->words[0] | ((uint64_t)d->words[1] << 32);
  d->pos = 2;
  return 0;
}

namespace {
struct DecCursor {
  uint64_t x;
  const uint32_t* w;
  size_t pos, nw;
  inline uint32_t next() { return pos < nw ? w[pos++] : 0u; }
  inline uint32_t get_bits() {
    const uint32_t val = (uint32_t)(x & ((1u << kBypassBits) - 1));
    x >>= kBypassBits;
    if (x < kRansL) x = (x << 32) | next();
    return val;
  }
  inline int32_t symbol(const vcm_cdf_tables* t, int32_t ti) {
    const uint32_t* cdf = &t->cdf[(size_t)ti * t->stride];
    const int32_t max_value = t->sizes[ti] - 2;
    const uint32_t cum = (uint32_t)(x & 0xffffu);
    {   // most probable value first: its record depends only on the table index, not on the state
      const uint32_t* m = &t->mode[(size_t)ti * 4];
      const uint32_t d = cum - m[1];
      if (d < m[2]) {
        x = (uint64_t)m[2] * (x >> kPrecision) + d;
        if (x < kRansL) x = (x << 32) | next();
        return (int32_t)m[0] + t->offsets[ti];
      }
    }
    int32_t s = t->lut[(size_t)ti * 256 + (cum >> 8)];
    while (cdf[s + 1] <= cum) ++s;
    const uint32_t start = cdf[s], freq = cdf[s + 1] - start;
    x = (uint64_t)freq * (x >> kPrecision) + cum - start;
    if (x < kRansL) x = (x << 32) | next();
    int32_t value = s;
    if (value == max_value) {
      int32_t v = (int32_t)get_bits();
      int32_t nb = v;
      while (v == kBypassMax) {
        v = (int32_t)get_bits();
        nb += v;
      }
      int32_t raw = 0;
      for (int32_t j = 0; j < nb; ++j) raw |= (int32_t)get_bits() << (j * kBypassBits);
      value = raw >> 1;
      if (raw & 1)
        value = -value - 1;
      else
        value += max_value;
    }
    return value + t->offsets[ti];
  }
};
inline int16_t sat16(int32_t v) { return (int16_t)(v < -32768 ? -32768 : (v > 32767 ? 32767 : v)); }
}  // namespace

extern "C" int vcm_rans_dec_decode(vcm_rans_dec* d, const uint8_t* idx, int64_t n,
                                   const vcm_cdf_tables* t, int16_t* out) {
  if (!d || !idx || !t || !out || n < 0 || d->words.empty()) {
    vcm_set_error("vcm_rans_dec_decode: bad argument / no stream");
    return 1;
  }
  DecCursor c{d->state, d->words.data(), d->pos, d->words.size()};
  for (int64_t i = 0; i < n; ++i) {
    const int32_t ti = idx[i];
    if (ti >= t->ntab) {
      vcm_set_error("vcm_rans_dec_decode: table index %d out of range", ti);
      return 1;
    }
    out[i] = sat16(c.symbol(t, ti));
  }
  d->state = c.x;
  d->pos = c.pos;
  return 0;
}

extern "C" int vcm_rans_dec_decode_channels(vcm_rans_dec* d, int32_t C, int64_t hw,
                                            const vcm_cdf_tables* t, int16_t* out) {
  if (!d || !t || !out || C <= 0 || C > t->ntab || hw < 0 || d->words.empty()) {
    vcm_set_error("vcm_rans_dec_decode_channels: bad argument / no stream");
    return 1;
  }
  DecCursor c{d->state, d->words.data(), d->pos, d->words.size()};
  for (int32_t ch = 0; ch < C; ++ch)
    for (int64_t i = 0; i < hw; ++i) out[(int64_t)ch * hw + i] = sat16(c.symbol(t, ch));
  d->state = c.x;
  d->pos = c.pos;
  return 0;
}

extern "C" int vcm_pmf_to_quantized_cdf(const float* pmf, int32_t n, int32_t precision, int32_t* cdf_out) {
  if (!pmf || !cdf_out || n <= 0 || precision <= 0 || precision > 30) {
    vcm_set_error("vcm_pmf_to_quantized_cdf: bad argument");
    return 2;
  }
  const int m = n + 1;
  std::vector<uint32_t> cdf((size_t)m);
  cdf[0] = 0;
  for (int i = 0; i < n; ++i) cdf[i + 1] = (uint32_t)(roundf(pmf[i] * (float)(1 << precision)) + 0.5f);
  uint32_t total = 0;
  for (int i = 0; i < m; ++i) total += cdf[i];
  if (total == 0) {
    vcm_set_error("vcm_pmf_to_quantized_cdf: empty pmf");
    return 2;
  }
  for (int i = 0; i < m; ++i) cdf[i] = (uint32_t)((((uint64_t)1 << precision) * cdf[i]) / total);
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
      if (best < 0) {
        vcm_set_error("vcm_pmf_to_quantized_cdf: no frequency donor");
        return 1;
      }
      if (best < i) {
        for (int j = best + 1; j <= i; ++j) cdf[j]--;
      } else {
        for (int j = i + 1; j <= best; ++j) cdf[j]++;
      }
    }
  }
  for (int i = 0; i < m; ++i) cdf_out[i] = (int32_t)cdf[i];
  return 0;
}

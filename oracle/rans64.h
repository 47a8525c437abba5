/*
 * oracle/rans64.h -- TEST INFRASTRUCTURE, not product code.
 *
 * Restatement of the six entry points of `rans64.h` from rygorous/ryg_rans
 * (pinned by the reference at commit c9d162d996fd600315af9ae8eb89d832576cb32d,
 * DCVC_HEM/src/cpp/3rdparty/ryg_rans/CMakeLists.txt.in:8-9).  The header is
 * NOT vendored in /root/reference and there is no network, so the published
 * algorithm (64-bit state, lower bound 2^31, 32-bit renormalisation words) is
 * restated here from its description.  It exists so that the reference's own
 * C++ (DCVC_HEM/src/cpp/rans/rans_interface.cpp) can be compiled UNMODIFIED
 * into oracle/_ref/ and used as the byte-exact checker for the product coder.
 *
 * Call sites in the reference that fix the semantics:
 *   Rans64EncInit     rans_interface.cpp:149
 *   Rans64EncPut      rans_interface.cpp:159
 *   Rans64EncFlush    rans_interface.cpp:167
 *   Rans64DecInit     rans_interface.cpp:181
 *   Rans64DecGet      rans_interface.cpp:205
 *   Rans64DecAdvance  rans_interface.cpp:213
 *   in-tree twins that mirror the renormalisation rule:
 *   Rans64EncPutBits  rans_interface.cpp:46-64, Rans64DecGetBits :66-82
 *
 * Parity note: no upstream golden bytes exist (the reference has no tests);
 * the restatement is pinned by (i) round trips through the reference C++,
 * (ii) byte equality with the independent big-int model oracle/rans_py.py and
 * the plain-C restatement oracle/rans_oracle.c.
 */
#ifndef ORACLE_RANS64_H
#define ORACLE_RANS64_H

#include <stdint.h>
#include <assert.h>

#ifndef Rans64Assert
#define Rans64Assert(x) assert(x)
#endif

/* Lower bound of the normalised interval [L, L * 2^32). */
#define RANS64_L (1ull << 31)

typedef uint64_t Rans64State;

/* Encoder starts at the bottom of the normalised interval. */
static inline void Rans64EncInit(Rans64State *r) { *r = RANS64_L; }

/* Push one symbol [start, start+freq) out of 2^scale_bits.  Output words are
 * written BACKWARDS (the pointer is pre-decremented), at most one 32-bit word
 * per symbol. */
static inline void Rans64EncPut(Rans64State *r, uint32_t **pptr, uint32_t start,
                                uint32_t freq, uint32_t scale_bits) {
  Rans64Assert(freq != 0);
  uint64_t x = *r;
  const uint64_t limit = ((RANS64_L >> scale_bits) << 32) * (uint64_t)freq;
  if (x >= limit) {
    *pptr -= 1;
    **pptr = (uint32_t)x;
    x >>= 32;
    Rans64Assert(x < limit);
  }
  *r = ((x / freq) << scale_bits) + (x % freq) + start;
}

/* Final state: two words, low word at the lower address. */
static inline void Rans64EncFlush(Rans64State *r, uint32_t **pptr) {
  const uint64_t x = *r;
  *pptr -= 2;
  (*pptr)[0] = (uint32_t)(x & 0xffffffffu);
  (*pptr)[1] = (uint32_t)(x >> 32);
}

static inline void Rans64DecInit(Rans64State *r, uint32_t **pptr) {
  uint64_t x = (uint64_t)(*pptr)[0];
  x |= ((uint64_t)(*pptr)[1]) << 32;
  *pptr += 2;
  *r = x;
}

/* Cumulative frequency of the next symbol. */
static inline uint32_t Rans64DecGet(Rans64State *r, uint32_t scale_bits) {
  return (uint32_t)(*r & ((1u << scale_bits) - 1));
}

/* Pop the symbol [start, start+freq) and renormalise (reads forwards). */
static inline void Rans64DecAdvance(Rans64State *r, uint32_t **pptr,
                                    uint32_t start, uint32_t freq,
                                    uint32_t scale_bits) {
  const uint64_t mask = (1ull << scale_bits) - 1;
  uint64_t x = *r;
  x = (uint64_t)freq * (x >> scale_bits) + (x & mask) - start;
  if (x < RANS64_L) {
    x = (x << 32) | (uint64_t)(**pptr);
    *pptr += 1;
    Rans64Assert(x >= RANS64_L);
  }
  *r = x;
}

#endif /* ORACLE_RANS64_H */

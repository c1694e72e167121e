// Device-side helpers shared by all kernels of libgtsfx_b200: packed-text access, the special
// bitmap, k-mer codes and the 64-bit sort key.  sm_100a only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define GTB_FULL 0xffffffffu
#define GTB_POSMODE 0xffffffffu   // depth marker: group is ordered by text position

// ---------------------------------------------------------------------------------------------
// Text layout in HBM
//   text : uint64_t words, symbols packed big-endian, BITS per symbol, SPW = 64/BITS per word:
//          symbol i sits in bits [64-BITS*(i%SPW+1), 64-BITS*(i%SPW)) of word i/SPW.  For DNA
//          (BITS=2) this IS GtEncseq's twobitencoding (src/core/encseq.c:73-81), uploaded as is.
//          Other alphabets use BITS=8.  Special positions hold an arbitrary filler for BITS=2 and
//          0xff for BITS=8 (as do all bytes from position n to the end of the padding), so byte
//          text can be keyed without the bitmap.
//   mask : bitmap, bit (i&63) of word i>>6 (LSB first) is 1 iff position i holds a wildcard or
//          separator; all bits >= totallength are 1 (the virtual end is a special, GT_UNIQUEINT,
//          src/core/encseq.h:640).
//   gather: one 16-byte record per text word, {word i, word i+1}: the key window of any suffix (at most
//          SPW-1 symbols of offset + W symbols) lies inside the record of the word it starts in, so the
//          sorters fetch it with ONE aligned 128-bit load instead of two 64-bit ones.  For BITS=2 bit 0
//          of the second word (its last base is never part of a key) says whether any of the record's
//          first 59 positions is special; if not, the key needs no bitmap access either.  Costs a
//          second copy of the text (n/2 bytes for DNA), pays for itself in the sort kernels, whose
//          limit is the L1 tag rate of scattered loads.  Optional (g == nullptr: not built).
// Both arrays are padded with two extra words so a 64-bit window may start at any position <= n.
// ---------------------------------------------------------------------------------------------
template <int BITS>
struct Text {
  const uint64_t *__restrict__ w;
  const uint64_t *__restrict__ m;
  const ulonglong2 *__restrict__ g;
  static constexpr int SPW = 64 / BITS;
  static constexpr int W = 56 / BITS;   // symbols per sort key (low byte of the key is a tag)

  // 64 bits of text starting at symbol p (first symbol in the most significant bits)
  __device__ __forceinline__ uint64_t window(uint64_t p) const {
    const uint64_t i = p / SPW;
    const unsigned off = (unsigned)(p % SPW) * BITS;
    const uint64_t hi = __ldg(w + i), lo = __ldg(w + i + 1);
    return off ? (hi << off) | (lo >> (64 - off)) : hi;
  }
  // the same window through the gather table (only its leading W symbols are exact); special: the record
  // holds a special somewhere, the caller consults the bitmap (BITS=2 only; byte text shows specials itself)
  // g == nullptr (table not built: it would push the text out of L2, see open_impl): two 64-bit loads, and
  // for BITS=2 the bitmap decides.
  __device__ __forceinline__ uint64_t keywin(uint64_t p, bool &special) const {
    const unsigned off = (unsigned)(p % SPW) * BITS;
    if (g == nullptr) {
      const uint64_t hi = __ldg(w + p / SPW), lo = __ldg(w + p / SPW + 1);
      special = BITS == 2;
      return off ? (hi << off) | (lo >> (64 - off)) : hi;
    }
    const ulonglong2 r = __ldg(g + p / SPW);
    special = BITS == 2 && (r.y & 1ull);
    return off ? (r.x << off) | (r.y >> (64 - off)) : r.x;
  }
  // bit j = 1 iff position p+j is special, j = 0..63
  __device__ __forceinline__ uint64_t maskwin(uint64_t p) const {
    const uint64_t i = p >> 6;
    const unsigned s = (unsigned)(p & 63);
    const uint64_t a = __ldg(m + i), b = __ldg(m + i + 1);
    return s ? (a >> s) | (b << (64 - s)) : a;
  }
  // number of real symbols from p up to the next special, capped at 64
  __device__ __forceinline__ unsigned reallen(uint64_t p) const {
    const uint64_t mw = maskwin(p);
    return mw ? (unsigned)(__ffsll((long long)mw) - 1) : 64u;
  }
};

// Sort key of the suffix whose next unread symbol is at position q (= start + depth):
//   bits 63..8 : the next W symbols, positions at/after the first special padded with ones
//   bits  7..0 : W - (number of real symbols among them)   (0 = full window, order still open)
// Comparing (key, start position) reproduces the reference order: a key cut
// short by a special is larger than every real continuation, and two suffixes cut at the same depth
// with equal real prefixes are ordered by position.
template <int BITS>
__device__ __forceinline__ uint64_t make_key(const Text<BITS> &T, uint64_t q) {
  constexpr int W = Text<BITS>::W;
  unsigned avail;
  bool special;
  const uint64_t win = T.keywin(q, special);
  if (BITS == 2 && !special) return (win >> 8) << 8;      // W real symbols, tag 0
  if (BITS == 8) {
    // byte text: a special (and everything from position n on) is stored as 0xff/0xfe, real symbols are < 254,
    // so the window itself tells where the first special is - no bitmap access.  z has a zero 7-bit field
    // exactly where the byte is >= 0xfe; adding 0x7f sets bit 7 of every other field without carries.
    const uint64_t z = ((~win) >> 1) & 0x7f7f7f7f7f7f7f7full;
    const uint64_t real = (z + 0x7f7f7f7f7f7f7f7full) & 0x8080808080808080ull;
    const uint64_t spm = real ^ 0x8080808080808080ull;
    avail = spm ? (unsigned)__clzll((long long)spm) >> 3 : 8u;
  } else {
    const uint64_t mw = T.maskwin(q);
    avail = mw ? (unsigned)(__ffsll((long long)mw) - 1) : 64u;
  }
  if (avail > (unsigned)W) avail = W;
  uint64_t x = win >> 8;                     // W symbols in bits 55..0
  if (avail < (unsigned)W) {
    const unsigned padbits = (W - avail) * BITS;   // 1..56
    x |= (~0ull) >> (64 - padbits);
  }
  return (x << 8) | (uint64_t)(W - avail);
}

// same key when the number of real symbols ahead of q is already known (no bitmap access)
template <int BITS>
__device__ __forceinline__ uint64_t make_key_known(const Text<BITS> &T, uint64_t q, unsigned avail) {
  constexpr int W = Text<BITS>::W;
  if (avail > (unsigned)W) avail = W;
  bool special;
  uint64_t x = T.keywin(q, special) >> 8;
  if (avail < (unsigned)W) {
    const unsigned padbits = (W - avail) * BITS;
    x |= (~0ull) >> (64 - padbits);
  }
  return (x << 8) | (uint64_t)(W - avail);
}

// lcp contribution of two keys taken at the same depth: number of equal leading real symbols
template <int BITS>
__device__ __forceinline__ unsigned key_common(uint64_t ka, uint64_t kb) {
  constexpr int W = Text<BITS>::W;
  const uint64_t x = ka ^ kb;
  unsigned common = x ? (unsigned)__clzll((long long)x) / BITS : (unsigned)W;
  if (common > (unsigned)W) common = W;
  const unsigned aa = W - (unsigned)(ka & 0xff), ab = W - (unsigned)(kb & 0xff);
  common = common < aa ? common : aa;
  return common < ab ? common : ab;
}

// Code arithmetic (src/match/kmercodes.h:24-46 for sigma = 4; src/match/initbasepower.c:71-106 for
// the general base-sigma form).  pw[i] = sigma^i.
struct CodeParams {
  unsigned sigma, pl;
  int pow2;                 // sigma == 1 << BITS: codes are plain bit fields
  uint64_t pw[34];
};

template <int BITS>
__device__ __forceinline__ uint64_t prefix_value(uint64_t win, unsigned len, const CodeParams &cp) {
  // value of the first len symbols of the window read as a base-sigma number
  if (len == 0) return 0;
  if (cp.pow2) return win >> (64 - BITS * len);
  uint64_t v = 0;
  for (unsigned k = 0; k < len; ++k) v = v * cp.sigma + ((win >> (64 - BITS * (k + 1))) & ((1u << BITS) - 1));
  return v;
}

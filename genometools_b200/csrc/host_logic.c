/* Host-side pieces of the Sfxiterator logic that the device layer's callers need and that do not
 * touch the GPU: the automatic prefix length and the split of the bucket table into parts.
 * Written from the behaviour of the reference functions cited in include/gtsfx_b200.h. */
#include <stdint.h>
#include <stdlib.h>

#include "../../include/gtsfx_b200.h"

static uint64_t ipow(unsigned base, unsigned e) {
  uint64_t r = 1;
  while (e--) r *= base;
  return r;
}

static uint64_t bcktab_bytes(unsigned sigma, unsigned pl, uint64_t maxvalue) {
  const uint64_t w = maxvalue <= 0xffffffffULL ? 4 : 8;
  uint64_t dist = 0;
  unsigned i;
  for (i = 1; i + 1 < pl; i++) dist += ipow(sigma, i);
  return w * (ipow(sigma, pl) + 1) + w * ipow(sigma, pl - 1) + w * dist;
}

unsigned gtb200_recommendedprefixlength(unsigned numofchars, uint64_t totallength) {
  /* largest pl whose table, divided by the multiplier 0.25, does not exceed totallength bytes;
   * capped by the largest exponent for which sigma^pl still fits a code word */
  unsigned pl, cap = 0;
  uint64_t p = 1;
  const uint64_t minfailure = UINT64_MAX / numofchars;
  while (p < minfailure) {
    p *= numofchars;
    cap++;
  }
  for (pl = 1;; pl++)
    if ((double) bcktab_bytes(numofchars, pl, totallength + 1) / 0.25 > (double) totallength) break;
  pl--;
  if (pl == 0) return 1;
  return pl < cap ? pl : cap;
}

/* end offset of bucket c = start offset of bucket c+1 */
static uint64_t bucket_end(const uint32_t *leftborder, uint64_t c) { return leftborder[c + 1]; }

/* the reference's binary search over bucket END offsets (gt_bcktab_findfirstlarger): a code whose
 * end offset is >= target; stops at the first exact hit the bisection meets */
static uint64_t find_first_larger(const uint32_t *leftborder, uint64_t numofallcodes, uint64_t target) {
  uint64_t left = 0, right = numofallcodes - 1, found = numofallcodes;
  while (left <= right) {
    const uint64_t mid = (left + right) / 2;
    const uint64_t midval = bucket_end(leftborder, mid);
    if (target <= midval) {
      found = mid;
      if (target < midval) {
        if (mid == 0) break;
        right = mid - 1;
      } else {
        break;
      }
    } else {
      left = mid + 1;
    }
  }
  return found;
}

int gtb200_suftabparts(unsigned numofparts, const uint32_t *leftborder, uint64_t numofallcodes,
                       unsigned prefixlength, uint64_t *mincodes, uint64_t *maxcodes, uint64_t *widths) {
  const uint64_t total = leftborder[numofallcodes];
  unsigned parts, part, kept = 0, remainder;
  uint64_t target = 0, firstbound = 0, per;
  if (total == 0 || numofparts == 0) return 0;
  parts = (total < numofparts || prefixlength == 1) ? 1 : numofparts;
  per = total / parts;
  remainder = (unsigned) (total % parts);
  for (part = 0; part < parts; part++) {
    uint64_t lastcode, secondbound;
    target += per;
    if (remainder > 0) {
      target++;
      remainder--;
    }
    lastcode = (part == parts - 1) ? numofallcodes - 1 : find_first_larger(leftborder, numofallcodes, target);
    secondbound = bucket_end(leftborder, lastcode);
    if (secondbound > firstbound) {              /* empty parts are dropped */
      mincodes[kept] = kept ? maxcodes[kept - 1] + 1 : 0;
      maxcodes[kept] = lastcode;
      widths[kept] = secondbound - firstbound;
      kept++;
    } else if (kept) {
      maxcodes[kept - 1] = lastcode;             /* keep the code ranges contiguous */
    }
    firstbound = secondbound;
  }
  if (kept) maxcodes[kept - 1] = numofallcodes - 1;
  return (int) kept;
}

/* ---------------------------------------------------------------------------------------------
 * Host half of the narrow transfer (gtb200_sort_part with a GtUword suftab): positions cross PCIe as the 32-bit
 * values the device sorts (totallength < 2^32 - 1) and are widened here into the GtUword entries
 * gt_suffixsortspace_to_file writes (src/match/sfx-suffixgetset.c:462-478) - half the bytes on the bus.
 * Streaming stores: the .suf image is written once and not read again by this code.
 * --------------------------------------------------------------------------------------------- */
#include <stddef.h>
#if defined(__x86_64__)
#include <immintrin.h>
__attribute__((target("avx2"))) static void widen_avx2(const uint32_t *src, uint64_t *dst, size_t n)
{
  size_t i = 0;
  while (i < n && ((uintptr_t) (dst + i) & 31) != 0) { dst[i] = src[i]; i++; }
  for (; i + 8 <= n; i += 8)
  {
    const __m128i lo = _mm_loadu_si128((const __m128i *) (src + i));
    const __m128i hi = _mm_loadu_si128((const __m128i *) (src + i + 4));
    _mm256_stream_si256((__m256i *) (dst + i), _mm256_cvtepu32_epi64(lo));
    _mm256_stream_si256((__m256i *) (dst + i + 4), _mm256_cvtepu32_epi64(hi));
  }
  for (; i < n; i++) dst[i] = src[i];
  _mm_sfence();
}
#endif

void gtb200_widen_u32_u64(const uint32_t *src, uint64_t *dst, size_t n)
{
#if defined(__x86_64__)
  if (__builtin_cpu_supports("avx2")) { widen_avx2(src, dst, n); return; }
#endif
  for (size_t i = 0; i < n; i++) dst[i] = src[i];
}

/* ---------------------------------------------------------------------------------------------
 * The per-GPU cut of one part (gtb200_multi_sort_part): [mincode, maxcode] into at most nshards code ranges of about
 * equal suffix count.  Same rule as gt_suftabparts_new applies to the whole table (src/match/sfx-partssuf.c:172-343,
 * gt_bcktab_findfirstlarger src/match/bcktab.c:1322-1381): shard d ends at the first bucket whose END offset reaches
 * base + width * d / nshards; ranges without suffixes are merged into their successor.  leftborder = bucket START
 * offsets (numofallcodes + 1 entries).  Returns the number of ranges written to lo[] / hi[] (inclusive codes).
 * --------------------------------------------------------------------------------------------- */
int gtb200_split_range(const uint32_t *leftborder, uint64_t mincode, uint64_t maxcode, int nshards,
                       uint64_t *lo, uint64_t *hi)
{
  const uint64_t base = leftborder[mincode], width = (uint64_t) leftborder[maxcode + 1] - base;
  uint64_t start = mincode;
  int k = 0, d;

  for (d = 1; d <= nshards && start <= maxcode; d++)
  {
    uint64_t end;
    if (d == nshards)
    {
      end = maxcode;
    } else
    {
      const uint64_t target = base + width * (uint64_t) d / (uint64_t) nshards;
      uint64_t a = start + 1, b = maxcode + 2;          /* first index i in [a,b) with leftborder[i] >= target */
      while (a < b)
      {
        const uint64_t mid = a + (b - a) / 2;
        if ((uint64_t) leftborder[mid] < target) a = mid + 1; else b = mid;
      }
      end = a - 1;
      if (end > maxcode) end = maxcode;
      if (leftborder[end + 1] == leftborder[start]) continue;     /* nothing to sort yet: extend with the next target */
    }
    lo[k] = start;
    hi[k] = end;
    k++;
    start = end + 1;
  }
  if (k > 0 && hi[k - 1] != maxcode) hi[k - 1] = maxcode;
  return k;
}

/* ---------------------------------------------------------------------------------------------
 * "0,1,4-7" -> {0,1,4,5,6,7}: the device list of GTB200_DEVICES (integration/sfx_suffixer_b200.c).  Stops at the
 * first character that is neither a number, ',' nor '-'; returns the number of devices written (<= capacity), or 0.
 * --------------------------------------------------------------------------------------------- */
#include <stdlib.h>
int gtb200_parse_devices(const char *text, int *devices, int capacity)
{
  int n = 0;
  if (text == NULL) return 0;
  while (*text != '\0' && n < capacity)
  {
    char *end;
    long a = strtol(text, &end, 10), b;
    if (end == text || a < 0) break;
    b = a;
    if (*end == '-')
    {
      text = end + 1;
      b = strtol(text, &end, 10);
      if (end == text || b < a) break;
    }
    for (; a <= b && n < capacity; a++) devices[n++] = (int) a;
    if (*end != ',') break;
    text = end + 1;
  }
  return n;
}

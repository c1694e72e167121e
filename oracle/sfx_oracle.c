/*
 * oracle/sfx_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C, single-threaded CPU restatement of what GenomeTools' enhanced-suffix-array path
 * (`gt suffixerator -suf -lcp -bck`) *outputs*, written from the definitions the reference
 * implements, not from its code.  It is deliberately naive (comparison sort + direct LCP) so that
 * it can be read in one sitting and trusted as a checker.
 *
 * Pinned: tests/test_oracle.py compares it byte for byte against (a) the golden .suf/.lcp/.bck
 * files in tests/golden/ that tests/golden/make_golden.py produced with the UNMODIFIED reference
 * (oracle/_ref/gt_ref, built by oracle/Makefile from /root/reference), and (b) against gt_ref
 * itself on seeded random inputs whenever the binary is present.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library; the
 * product (genometools_b200/) must never call it.
 *
 * Reference semantics restated here (paths relative to /root/reference):
 *   order of suffixes ........ src/core/encseq.c:7381-7473 (gt_encseq_check_comparetwosuffixes),
 *                              src/core/encseq.h:640 (GT_UNIQUEINT): a special symbol or the end of
 *                              the text is larger than every real symbol and two specials are
 *                              ordered by their text position.
 *   lcp ...................... src/match/sfx-linlcp.c:31-72 (Kasai) with the rule that a special
 *                              never matches (src/match/sfx-lcpvalues.c:91-111,125-215): the result
 *                              is the number of equal leading *real* symbols; tail entries are 0.
 *   bucket codes / .bck ...... src/match/kmercodes.h:24-46 (code of a pl-mer),
 *                              src/match/sfx-suffixer.c:1081-1102 (padding with sigma-1),
 *                              src/match/bcktab.c:876-901 (countspecialcodes / distpfxidx indices),
 *                              src/match/bcktab.c:519-570 (table order in the file).
 *   tail of the suffix array . src/match/sfx-suffixgetset.c:586-690: positions of all special
 *                              symbols in ascending order, then totallength.
 *   statistics for .prj ...... src/match/sfx-lcpvalues.c:414 (sum only over suffixes whose first pl
 *                              symbols are real), :153-156 (maxbranchdepth over all entries).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define SFXO_WILDCARD 254

typedef struct {
  uint64_t nonspecial;       /* number of suffixes that start with a real symbol */
  uint64_t longest;          /* index r with suf[r] == 0 */
  uint64_t maxbranchdepth;   /* max lcp */
  uint64_t numlargelcp;      /* #entries with lcp >= 255 */
  double   lcpsum;           /* see header comment */
} sfxo_stats;

static const uint8_t *g_seq;
static const uint32_t *g_dist;  /* g_dist[p] = number of real symbols from p up to the next special */

static int sfxo_compare(const void *pa, const void *pb)
{
  uint64_t a = *(const uint64_t *) pa, b = *(const uint64_t *) pb;
  uint32_t la = g_dist[a], lb = g_dist[b], common = la < lb ? la : lb;
  int c = memcmp(g_seq + a, g_seq + b, common);
  if (c != 0) return c;
  if (la != lb) return la < lb ? 1 : -1;  /* the one that hits a special first is larger */
  return a < b ? -1 : 1;                  /* two specials at the same depth: by position */
}

static uint64_t sfxo_pow(unsigned base, unsigned e)
{
  uint64_t r = 1;
  while (e--) r *= base;
  return r;
}

/* number of counters in distpfxidx: sum_{i=1}^{pl-2} sigma^i (src/match/bcktab.c:240-256) */
uint64_t sfxo_numofdistpfxidxcounters(unsigned sigma, unsigned pl)
{
  uint64_t s = 0;
  unsigned i;
  for (i = 1; i + 1 < pl; i++) s += sfxo_pow(sigma, i);
  return s;
}

/* size in bytes of the bucket table for a given pl (src/match/bcktab.c:263-324), 4-byte entries
 * while maxvalue <= UINT_MAX */
static uint64_t sfxo_sizeofbcktab(unsigned sigma, unsigned pl, uint64_t maxvalue)
{
  uint64_t w = maxvalue <= 0xffffffffULL ? 4 : 8;
  return w * (sfxo_pow(sigma, pl) + 1) + w * sfxo_pow(sigma, pl - 1) +
         w * sfxo_numofdistpfxidxcounters(sigma, pl);
}

/* automatic prefix length (src/match/sfx-apfxlen.c:50-107 with multiplier 0.25,
 * src/match/initbasepower.c:23-34 for the cap) */
unsigned sfxo_recommendedprefixlength(unsigned sigma, uint64_t totallength)
{
  unsigned pl, cap = 0;
  uint64_t p = 1, minfailure = UINT64_MAX / sigma;
  while (p < minfailure) { p *= sigma; cap++; }
  for (pl = 1;; pl++) {
    if ((double) sfxo_sizeofbcktab(sigma, pl, totallength + 1) / 0.25 > (double) totallength)
      break;
  }
  pl -= 1;
  if (pl == 0) return 1;
  return pl < cap ? pl : cap;
}

/*
 * seq[0..n): symbol codes 0..sigma-1, 254 wildcard, 255 separator.
 * suf[n+1], lcp[n+1] (full-width values), leftborder[sigma^pl+1] (bucket START offsets, last =
 * number of non-special suffixes), countspecialcodes[sigma^(pl-1)], distpfxidx[sum sigma^i] are
 * caller-allocated.  Returns 0 on success.
 */
int sfxo_build(const uint8_t *seq, uint64_t n, unsigned sigma, unsigned pl,
               uint64_t *suf, uint32_t *lcp, uint32_t *leftborder,
               uint32_t *countspecialcodes, uint32_t *distpfxidx, sfxo_stats *stats)
{
  uint32_t *dist = malloc(sizeof *dist * (n + 1));
  uint64_t p, m = 0, r, numcodes = sfxo_pow(sigma, pl), t;
  uint64_t *fill;
  unsigned i;

  if (dist == NULL) return -1;
  dist[n] = 0;
  for (p = n; p-- > 0;) dist[p] = seq[p] >= SFXO_WILDCARD ? 0 : dist[p + 1] + 1;

  /* --- suffix array: sorted part, then the tail --- */
  for (p = 0; p < n; p++) if (dist[p] > 0) suf[m++] = p;
  g_seq = seq; g_dist = dist;
  qsort(suf, m, sizeof *suf, sfxo_compare);
  t = m;
  for (p = 0; p < n; p++) if (dist[p] == 0) suf[t++] = p;
  suf[t++] = n;

  /* --- lcp: equal leading real symbols of neighbours; 0 in the tail and at index 0 --- */
  memset(lcp, 0, sizeof *lcp * (n + 1));
  for (r = 1; r < m; r++) {
    uint64_t a = suf[r - 1], b = suf[r];
    uint32_t lim = dist[a] < dist[b] ? dist[a] : dist[b], l = 0;
    while (l < lim && seq[a + l] == seq[b + l]) l++;
    lcp[r] = l;
  }

  /* --- bucket table --- */
  memset(leftborder, 0, sizeof *leftborder * (numcodes + 1));
  memset(countspecialcodes, 0, sizeof *countspecialcodes * sfxo_pow(sigma, pl - 1));
  if (sfxo_numofdistpfxidxcounters(sigma, pl) > 0)
    memset(distpfxidx, 0, sizeof *distpfxidx * sfxo_numofdistpfxidxcounters(sigma, pl));
  fill = malloc(sizeof *fill * (pl + 1));
  for (i = 0; i <= pl; i++) fill[i] = sfxo_pow(sigma, pl - i) - 1; /* i real symbols, rest sigma-1 */
  for (p = 0; p < n; p++) {
    uint32_t real = dist[p] < pl ? dist[p] : pl;
    uint64_t prefix = 0, code;
    if (real == 0) continue;
    for (i = 0; i < real; i++) prefix = prefix * sigma + seq[p + i];
    code = prefix * (fill[real] + 1) + fill[real];
    leftborder[code + 1]++;      /* shifted by one: becomes the start offset after the scan */
    if (real < pl) {
      countspecialcodes[(code - (sigma - 1)) / sigma]++;
      if (real + 1 < pl) {
        uint64_t off = 0;
        for (i = 1; i < real; i++) off += sfxo_pow(sigma, i);
        distpfxidx[off + (code - fill[real]) / (fill[real] + 1)]++;
      }
    }
  }
  for (t = 1; t <= numcodes; t++) leftborder[t] += leftborder[t - 1];

  /* --- statistics --- */
  stats->nonspecial = m;
  stats->longest = 0;
  stats->maxbranchdepth = 0;
  stats->numlargelcp = 0;
  stats->lcpsum = 0.0;
  for (r = 0; r <= n; r++) {
    if (suf[r] == 0) stats->longest = r;
    if (lcp[r] > stats->maxbranchdepth) stats->maxbranchdepth = lcp[r];
    if (lcp[r] >= 255) stats->numlargelcp++;
    if (r < m && dist[suf[r]] >= pl) stats->lcpsum += (double) lcp[r];
  }
  free(fill);
  free(dist);
  return 0;
}

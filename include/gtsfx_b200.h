/*
 * gtsfx_b200.h -- C ABI of libgtsfx_b200.so: the B200 (sm_100a) device layer behind GenomeTools'
 * enhanced-suffix-array construction path (`gt suffixerator -suf -lcp -bck`).
 *
 * The interface this library sits behind is the reference's Sfxiterator API
 * (src/match/sfx-suffixer.h:33-71, unchanged).  The functions below are what a B200 build of
 * src/match/sfx-suffixer.c calls instead of its CPU loops; INTEGRATION.md shows that binding.
 * All pointers are plain host pointers unless a comment says "device"; sizes are explicit; no
 * C++/torch types appear.  Every function returns 0 on success and -1 on error, in which case
 * gtb200_last_error() returns a message to copy into GtError (gt_error_set); there is NO CPU
 * fallback: without a CUDA device gtb200_open fails.
 *
 * Citations are relative to the reference checkout (genometools 1.6.7).
 */
#ifndef GTSFX_B200_H
#define GTSFX_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gtb200_ctx gtb200_ctx;

/* GtRange, half-open [start,end): src/core/range_api.h; produced in ascending order by
 * gt_specialrangeiterator_next (src/core/encseq.c:3745). */
typedef struct {
  uint64_t start, end;
} gtb200_range;

/* Largelcpvalue: src/match/lcpoverflow.h:24-30 -- one 16-byte .llv record. */
typedef struct {
  uint64_t position; /* index into the (global) suffix array */
  uint64_t value;
} gtb200_largelcp;

#define GTB200_MEM_HOST   0
#define GTB200_MEM_DEVICE 1

/*
 * The sequence as the host side already holds it in a GtEncseq.  Give EITHER
 *  - twobitencoding (+ the special ranges): the words gt_encseq_twobitencoding_export returns
 *    (src/core/encseq.c:6697; base i in bits 62-2*(i%32) of word i/32, src/core/intbits.h:78-84),
 *    only for numofchars == 4; OR
 *  - encoded: one byte per symbol as gt_encseq_extract_encoded delivers them (codes 0..numofchars-1,
 *    254 = GT_WILDCARD, 255 = GT_SEPARATOR, src/core/chardef_api.h:34-58), any alphabet; the
 *    special ranges are then derived on the device and the two range fields are ignored.
 * memory says where ALL pointers of the descriptor live (GTB200_MEM_DEVICE: already resident in
 * the HBM of `device`, e.g. after an NCCL broadcast).
 */
typedef struct {
  unsigned numofchars;              /* gt_alphabet_num_of_chars */
  uint64_t totallength;             /* gt_encseq_total_length; must be < 2^32 - 1 */
  const uint64_t *twobitencoding;
  uint64_t numofwords;              /* words readable at twobitencoding */
  const uint8_t *encoded;
  const gtb200_range *specialranges;
  uint64_t numofspecialranges;
  int memory;
  int readmode;                     /* GtReadmode (src/core/readmode_api.h:24-27): 0 forward, 1 reverse, 2 complement,
                                       3 reverse complement; the arrays above always describe the FORWARD sequence */
} gtb200_seqdesc;

/* What gt_Sfxiterator_new logs / needs after the counting phase. */
typedef struct {
  uint64_t numofallcodes;        /* sigma^prefixlength */
  uint64_t numofspecialcodes;    /* sigma^(prefixlength-1) */
  uint64_t numofdistpfxidx;      /* sum_{i=1}^{pl-2} sigma^i (src/match/bcktab.c:240-256) */
  uint64_t numofsuffixestosort;  /* totallength - specialcharacters */
  uint64_t specialcharacters;
  uint64_t largestbucket;        /* return value of gt_bcktab_leftborderpartialsums */
} gtb200_countinfo;

/* Per-part results the host needs for GtSuffixsortspace, GtOutlcpinfo and the .prj file. */
typedef struct {
  uint64_t partoffset;           /* global suffix-array index of the part's first entry */
  uint64_t width;                /* number of suffixes in the part */
  uint64_t numoflargelcpvalues;  /* entries with lcp >= 255 (records delivered in llv) */
  uint64_t maxbranchdepth;       /* max lcp in the part */
  double   lcptabsum;            /* sum the reference accumulates in outlcpvalues
                                    (src/match/sfx-lcpvalues.c:414): special-padded entries excluded */
  uint64_t longest;              /* global index r with suftab[r] == 0, or UINT64_MAX if not in this part */
  uint64_t fallback_suffixes;    /* diagnostics: suffixes that needed the refinement path */
  uint64_t fallback_rounds;
} gtb200_partinfo;

/* -- life cycle ------------------------------------------------------------------------------ */

/* Number of CUDA devices, or -1 (with message) if the CUDA runtime cannot be initialised. */
int gtb200_device_count(void);

/* Thread-local message of the last failing call. */
const char *gtb200_last_error(void);

/*
 * Replaces the sequence set-up at the top of gt_Sfxiterator_new_withadditionalvalues
 * (src/match/sfx-suffixer.c:1363-1480): uploads the packed sequence, builds the special-position
 * bitmap and keeps both resident on `device` until gtb200_close.
 * prefixlength: as chosen by detpfxlen (src/match/sfx-run.c:319-367).
 */
int gtb200_open(gtb200_ctx **ctx, int device, const gtb200_seqdesc *seq, unsigned prefixlength);
void gtb200_close(gtb200_ctx *ctx);

/* Device buffers of closed contexts are kept in a process-wide cache so that the next gtb200_open
 * costs no cudaMalloc/cudaFree (gt_Sfxiterator_delete followed by a new iterator, e.g. in
 * src/match/esa-mmsearch.c:609-636).  This returns the cached memory to the driver. */
void gtb200_release_cached_memory(void);

/* The CUDA stream (cudaStream_t) all kernels of this context are launched on; for callers that
 * time with CUDA events or order their own copies after the kernels. */
void *gtb200_stream(gtb200_ctx *ctx);

/*
 * The same set-up straight from FASTA text (for the index path): the scan that
 * gt_sequence_buffer_fasta_advance (src/core/sequence_buffer_fasta.c:40-170) performs character by character inside
 * gt_encseq_encoder_encode (src/core/encseq.c:8523) - descriptions from '>' to the end of the line, white space
 * skipped, one separator between sequences, symbols mapped as gt_alphabet_new_dna/_protein do
 * (src/core/alphabet.c:84-91) - runs on the device; the text of several files is passed concatenated.
 * prefixlength 0 = gt_recommendedprefixlength.  The on-disk .esq/.ssp/.des files remain the host encoder's job.
 * An illegal character or a text without '>' is an error, as in the reference.
 */
typedef struct {
  uint64_t totallength, numofsequences, specialcharacters;
  unsigned prefixlength;
} gtb200_fastainfo;
int gtb200_open_fasta(gtb200_ctx **ctx, int device, const uint8_t *fasta, uint64_t nbytes, int protein,
                      unsigned prefixlength, gtb200_fastainfo *info);

/* -- phase 1: "counting prefix distribution" ------------------------------------------------- */

/*
 * Replaces updateleftborder_getencseqkmers_twobitencoding (src/match/sfx-mapped4.gen:232-333 as
 * instantiated at src/match/sfx-suffixer.c:1132-1138), gt_updateleftborderforspecialkmer
 * (src/match/sfx-suffixer.c:1081-1102), getencseqkmers+updatekmercount for other alphabets
 * (src/match/sfx-mappedstr.c:728-779, src/match/sfx-suffixer.c:258-337), the counter part of
 * gt_bcktab_updatespecials (src/match/bcktab.c:876-901) and gt_bcktab_leftborderpartialsums
 * (src/match/bcktab.c:1274-1304).
 * Outputs (any may be NULL to skip the copy), 4-byte entries exactly as the .bck file stores them
 * (src/match/bcktab.c:519-570):
 *   leftborder[numofallcodes+1]       start index of each bucket; last = numofsuffixestosort
 *   countspecialcodes[numofspecialcodes]
 *   distpfxidx[numofdistpfxidx]
 */
int gtb200_count(gtb200_ctx *ctx, uint32_t *leftborder, uint32_t *countspecialcodes,
                 uint32_t *distpfxidx, gtb200_countinfo *info);

/*
 * Multi-GPU form of the counting phase ("counting position-sharded + all-reduce of
 * the sigma^pl histogram").  gtb200_count_shard counts only the suffixes that START in [pos_lo, pos_hi)
 * (pos_lo and pos_hi multiples of 64, pos_hi may be totallength) and leaves the raw counters - leftborder
 * counts, countspecialcodes, distpfxidx, contiguous uint32_t[*ntablewords] at *tables_dev (device) - on the
 * device; exactly one rank passes with_padded_counts = 1 (the special-padded suffixes come from the range
 * table, which every rank holds).  The caller sums the tables over all ranks (ncclAllReduce on the
 * context's stream), then every rank calls gtb200_count_finish, which performs the prefix sums and
 * delivers what gtb200_count delivers.  One rank with [0, totallength) and with_padded_counts = 1 followed
 * by gtb200_count_finish is identical to gtb200_count.
 */
int gtb200_count_shard(gtb200_ctx *ctx, uint64_t pos_lo, uint64_t pos_hi, int with_padded_counts,
                       void **tables_dev, uint64_t *ntablewords);
int gtb200_count_finish(gtb200_ctx *ctx, uint32_t *leftborder, uint32_t *countspecialcodes,
                        uint32_t *distpfxidx, gtb200_countinfo *info);

/* -- phase 2+3: insert, sort and LCP-annotate one part --------------------------------------- */

/*
 * Replaces, for the code range [mincode,maxcode] of one part (gt_suftabparts_minindex/maxindex,
 * src/match/sfx-partssuf.c): sfx_derivespecialcodesonthefly (src/match/sfx-suffixer.c:476-516),
 * insertsuffix_getencseqkmers_twobitencoding (:1164-1169, :371-399), gt_sortallbuckets
 * (src/match/sfx-bentsedg.c:1536-1612) with every sorter below it, the in-sort gt_lcptab_update
 * calls and the bucket-boundary / special-suffix LCP logic of src/match/sfx-lcpvalues.c:91-215,
 * 591-795.  The result stays in device buffers owned by the context (see gtb200_part_device).
 * gtb200_count must have been called.
 */
int gtb200_sort_part_device(gtb200_ctx *ctx, uint64_t mincode, uint64_t maxcode, gtb200_partinfo *info);

/* Device pointers of the last sorted part: suftab (uint64_t[width]), lcptab (uint8_t[width], 255 =
 * overflow), llv (gtb200_largelcp[numoflargelcpvalues], ascending position). */
int gtb200_part_device(gtb200_ctx *ctx, void **suftab_dev, void **lcptab_dev, void **llv_dev);

/*
 * Copies the last sorted part to host buffers: suftab is GtUword[width] (or uint32_t[width] when
 * suftab_is_uint32, the reference's -suftabuint, src/match/sfx-suffixgetset.c:48-55), lcptab
 * uint8_t[width], llv up to llv_capacity records (error if more are needed).  Any may be NULL.
 */
int gtb200_fetch_part(gtb200_ctx *ctx, void *suftab, int suftab_is_uint32, uint8_t *lcptab,
                      gtb200_largelcp *llv, uint64_t llv_capacity);

/* gtb200_sort_part_device followed by gtb200_fetch_part: the one call the patched
 * gt_sfxiterator_preparethispart (src/match/sfx-suffixer.c:1947-2160) makes. */
int gtb200_sort_part(gtb200_ctx *ctx, uint64_t mincode, uint64_t maxcode, void *suftab,
                     int suftab_is_uint32, uint8_t *lcptab, gtb200_largelcp *llv,
                     uint64_t llv_capacity, gtb200_partinfo *info);

/*
 * Replaces gt_SSSPbuf_fillspecialnextpage (src/match/sfx-suffixgetset.c:586-690): entries
 * [first, first+count) of the tail of the suffix array (positions of all special symbols in
 * ascending order, then totallength; specialcharacters+1 entries in total).  With suftab == NULL
 * the entries are only produced on the device, stream-ordered on gtb200_stream (no host wait).
 */
int gtb200_tail(gtb200_ctx *ctx, uint64_t first, uint64_t count, void *suftab, int suftab_is_uint32);
/* Device pointer of the entries the last gtb200_tail call produced (GtUword or uint32_t as requested there), valid
 * until the next gtb200_tail or gtb200_close and ordered after that call on gtb200_stream: the device-side
 * counterpart of the page buffer gt_SSSPbuf_fillspecialnextpage fills (src/match/sfx-suffixgetset.c:586-690). */
int gtb200_tail_device(gtb200_ctx *ctx, void **tail_dev);

/* -- host-side helpers that mirror reference logic the iterator needs (pure C, no device) ---- */

/* gt_recommendedprefixlength with GT_RECOMMENDED_MULTIPLIER_DEFAULT (src/match/sfx-apfxlen.c:82-107). */
unsigned gtb200_recommendedprefixlength(unsigned numofchars, uint64_t totallength);

/*
 * gt_suftabparts_new for a whole-range bucket table (src/match/sfx-partssuf.c:172-343 with
 * gt_bcktab_findfirstlarger, src/match/bcktab.c:1322-1381): splits [0,numofallcodes) into at most
 * numofparts code ranges of about equal suffix count, empty parts removed.  leftborder = bucket
 * START offsets as gtb200_count returns them.  mincodes/maxcodes/widths have room for numofparts
 * entries; returns the number of parts (>= 0).
 */
int gtb200_suftabparts(unsigned numofparts, const uint32_t *leftborder, uint64_t numofallcodes,
                       unsigned prefixlength, uint64_t *mincodes, uint64_t *maxcodes,
                       uint64_t *widths);

/*
 * The per-GPU cut of one part, as gtb200_multi_sort_part applies it: [mincode, maxcode] into at most nshards code
 * ranges of about equal suffix count, by the rule gt_suftabparts_new applies to the whole table
 * (src/match/sfx-partssuf.c:172-343).  lo/hi have room for nshards entries (inclusive codes); returns their number.
 */
int gtb200_split_range(const uint32_t *leftborder, uint64_t mincode, uint64_t maxcode, int nshards,
                       uint64_t *lo, uint64_t *hi);

/* "0,1,4-7" -> {0,1,4,5,6,7}: the device list a B200 build of libgenometools reads from GTB200_DEVICES (no reference
 * counterpart: the reference has no devices; its analogue is `gt -j`, src/gtr.c:181).  Returns the count, 0 if none. */
int gtb200_parse_devices(const char *text, int *devices, int capacity);

/* Host half of the narrow transfer: widens 32-bit positions to the GtUword entries gt_suffixsortspace_to_file writes
 * (src/match/sfx-suffixgetset.c:462-478), streaming stores. */
void gtb200_widen_u32_u64(const uint32_t *src, uint64_t *dst, size_t n);

/* Per-kernel-class device time (ms, CUDA events on the context's stream) accumulated since the
 * last gtb200_profile_reset; names[i] are static strings.  Returns the number of classes. */
int gtb200_profile_enable(gtb200_ctx *ctx, int on);
int gtb200_profile_read(gtb200_ctx *ctx, const char **names, double *ms, uint64_t *launches, int capacity);
void gtb200_profile_reset(gtb200_ctx *ctx);

/* -- several GPUs of one box behind one handle ------------------------------------------------ */

/*
 * The multi-GPU form of the calls above ("gtb200_open(devs[], ndev, ...)", 8(e)).  It replaces
 * the same reference code - the parts loop of gt_Sfxiterator_next / gt_sfxiterator_preparethispart
 * (src/match/sfx-suffixer.c:1947-2204) driven by gt_suftabparts_new (src/match/sfx-partssuf.c:172-343) - with
 * every part cut once more into ndev bucket ranges, one per GPU:
 *   gtb200_multi_open       uploads the sequence to devices[0] and copies its device image to the other GPUs peer to
 *                           peer (NVLink); same arguments as gtb200_open otherwise.
 *   gtb200_multi_count      = gtb200_count; the k-mers are counted position-sharded, the tables summed over peer
 *                           memory by the library's own kernel (replicated counting if the GPUs cannot reach each
 *                           other's memory).
 *   gtb200_multi_sort_part  = gtb200_sort_part for [mincode, maxcode]: the ndev ranges are sorted concurrently (one
 *                           host thread per GPU inside the call) and land in bucket order in the caller's buffers;
 *                           info is the merged part info.  Output is byte-identical for any number of devices.
 *   gtb200_multi_tail       = gtb200_tail.
 * With ndev == 1 every call is the single-GPU call.
 */
typedef struct gtb200_multi gtb200_multi;
int gtb200_multi_open(gtb200_multi **m, const int *devices, int ndev, const gtb200_seqdesc *seq,
                      unsigned prefixlength);
void gtb200_multi_close(gtb200_multi *m);
int gtb200_multi_count(gtb200_multi *m, uint32_t *leftborder, uint32_t *countspecialcodes,
                       uint32_t *distpfxidx, gtb200_countinfo *info);
int gtb200_multi_sort_part(gtb200_multi *m, uint64_t mincode, uint64_t maxcode, void *suftab,
                           int suftab_is_uint32, uint8_t *lcptab, gtb200_largelcp *llv,
                           uint64_t llv_capacity, gtb200_partinfo *info);
int gtb200_multi_tail(gtb200_multi *m, uint64_t first, uint64_t count, void *suftab, int suftab_is_uint32);
int gtb200_multi_devices(const gtb200_multi *m);
gtb200_ctx *gtb200_multi_ctx(gtb200_multi *m, int i);        /* the i-th device's context (profiling, streams) */
uint64_t gtb200_multi_peer_bytes(const gtb200_multi *m);     /* bytes moved between the GPUs so far */
/* Smallest free (and total) device memory over the handle's GPUs; the host side derives the number of parts from
 * it, as gt_suftabparts_fit_memlimit does from -memlimit (src/match/sfx-partssuf.c:449-527). */
int gtb200_multi_free_memory(gtb200_multi *m, uint64_t *free_bytes, uint64_t *total_bytes);

/*
 * Page-locks (cudaHostRegister) / releases a host buffer the caller owns - the GtUword array of the
 * GtSuffixsortspace (src/match/sfx-suffixgetset.c:157-167) and the lcp buffer - so that gtb200_sort_part's copies of
 * finished slices overlap the sorting of the next ones; with pageable memory every copy is synchronous.
 */
int gtb200_host_register(void *ptr, size_t bytes);
void gtb200_host_unregister(void *ptr);

/* Diagnostics of the last sorted part, for the benchmark's roofline arithmetic (no reference counterpart; the
 * reference logs the analogous sorter counters at src/match/sfx-bentsedg.c:1520-1525): values[0] = 1 if the tile
 * path ran, [1] = partition levels (1 or 2), [2] = 1 if the records carried their sort keys (16-byte records),
 * [3] = suffixes sorted by the block-level sorter (buckets of 129..1024), [4] = upper bound on the number of
 * special-padded suffixes.  Returns the number of defined values (<= 8). */
int gtb200_part_diag(gtb200_ctx *ctx, uint64_t *values, int capacity);

/* -- entry point that never allocates -------------------------------------------------------- */

/*
 * No counterpart in the reference: this exists for hosts that cannot let a library allocate device
 * memory - an application whose own pool already owns all of the GPU's memory (a library cudaMalloc
 * would fail or fragment it), or one that cannot afford the device-wide synchronisation hidden in
 * cudaMalloc/cudaFree while its other streams are busy.
 *
 * It is the prefix-sum step of gtb200_count on its own: from the bucket sizes counts_dev[0 ..
 * numofallcodes) it writes the bucket start offsets leftborder_dev[0 .. numofallcodes] (last entry =
 * total), both in device memory owned by the caller; leftborder_dev may be the same buffer as
 * counts_dev if it has room for the extra entry.  The partial sums go to a fixed array inside the
 * library (about 64 KB, resident on every GPU that loads the library until the process ends), so
 * no cudaMalloc/cudaFree or other allocating call is made.  numofallcodes above
 * GTB200_NOALLOC_MAX_CODES (4^13, the table of a 3.1 Gbp DNA index) is rejected with an error.
 * The first call on a device resolves the array's address; issue it before starting any stream
 * capture.  The array is shared by all calls on that device and keeps the previous call's
 * contents - what must start at zero is cleared on entry - so calls that run concurrently on one
 * GPU corrupt each other; keep them on one stream or otherwise serialised.
 * Work is enqueued on `stream` (a cudaStream_t, NULL = default stream) of the current device and
 * is asynchronous.  Returns 0, or -1 with a message in errbuf (may be NULL).
 */
#define GTB200_NOALLOC_MAX_CODES 67108864ull
int gtb200_leftborder_partialsums_noalloc(const uint32_t *counts_dev, uint32_t *leftborder_dev,
                                          uint64_t numofallcodes, void *stream, char *errbuf,
                                          size_t errbuf_len);

#ifdef __cplusplus
}
#endif
#endif /* GTSFX_B200_H */

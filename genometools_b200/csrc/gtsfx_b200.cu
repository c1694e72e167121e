// libgtsfx_b200: sm_100a device layer for GenomeTools' enhanced-suffix-array construction path.
// See include/gtsfx_b200.h for the C ABI and DESIGN.md for the data layout and kernel list.
//
// Pipeline per context (one CUDA stream, everything resident in HBM):
//   open   : upload packed text, build special bitmap (k_mask_*, k_pack_bytes)
//   count  : k_count (regular pl-mers) + k_padded (special-padded suffixes) -> histogram,
//            scan -> leftborder; padded list radix-sorted once
//   part   : k_scatter (regular suffixes into buckets), k_place_padded, k_sort_buckets (warp-level
//            register bitonic sort of each bucket on the next 28 bases + fused in-bucket LCP),
//            refinement rounds for buckets that are too large or still tied (segmented LSD radix
//            sort on deeper text windows), k_boundary (LCP across bucket borders), k_widen.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/gtsfx_b200.h"
#include "common.cuh"
#include "prims.cuh"

// ---------------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

static int set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
  return -1;
}

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess)                                                                         \
      return set_error("CUDA error %s at %s:%d (%s)", cudaGetErrorString(e_), __FILE__, __LINE__, #call); \
  } while (0)

extern "C" const char *gtb200_last_error(void) { return g_err; }

extern "C" int gtb200_device_count(void) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) return set_error("cudaGetDeviceCount: %s", cudaGetErrorString(e));
  return n;
}

// ---------------------------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------------------------
enum KClass { KC_PREP = 0, KC_COUNT, KC_SCAN, KC_PADDED, KC_SCATTER, KC_SORT, KC_REFINE, KC_RADIX,
              KC_BOUNDARY, KC_WIDEN, KC_TAIL, KC_INIT, KC_SORTMED, KC_NUM };
static const char *const kclass_names[KC_NUM] = {"prep", "count", "scan", "padded", "scatter", "sort_lcp",
                                                 "refine", "radix", "boundary", "widen", "tail", "part_init",
                                                 "sort_medium"};

struct PartStats {          // lives in device memory, mirrored to pinned host memory at sync points
  unsigned long long lcpsum;
  unsigned long long llv_count;
  unsigned long long fb_elems;
  unsigned long long longest;
  unsigned long long scan_total;
  unsigned long long scan_total2;
  unsigned long long bsum;       // bucket-boundary part of lcpsum
  unsigned long long med_elems;  // suffixes in the buckets queued for the block-level sorter
  unsigned int maxdepth;
  unsigned int fb_buckets;
  unsigned int largest;
  unsigned int bmax;             // bucket-boundary part of maxdepth
  unsigned int med_count;        // buckets queued for the block-level sorter
  unsigned int tile_leftover;    // buckets the tile sorter handed to the global sorters
  unsigned int fb_unsorted;      // buckets that reach refinement without having been sorted at all
  unsigned int posmode_kept;     // open suffixes whose group is cut by a special (next round orders them by position)
};

struct ProfEvent { int cls; cudaEvent_t a, b; };
struct PoolBlock { void *p; size_t sz; };

struct gtb200_ctx {
  int device;
  cudaStream_t stream;
  unsigned bits;
  CodeParams cp;
  uint64_t n;
  uint64_t ntextwords, nmaskwords;
  uint64_t *d_text, *d_mask;
  ulonglong2 *d_gather;      // {word i, word i+1} per text word (common.cuh)
  gtb200_range *d_ranges;
  uint64_t nranges;
  uint32_t *d_rangeoff;      // [nranges+1] exclusive scan of range lengths (tail paging)
  uint64_t numcodes, numspecialcodes, numdist;
  uint32_t *d_lb;            // [numcodes+1] bucket start offsets (.bck leftborder)
  uint32_t *d_cursor;        // [numcodes]
  uint32_t *d_csc, *d_dist;
  uint64_t padcap, npadded;
  unsigned long long *d_persist;   // [0] number of special-padded suffixes, [1] largest bucket, [2] sum of counts
  uint64_t *d_padkey;        // (code << 6) | (pl - reallen), sorted
  uint32_t *d_padpos;
  uint32_t *d_padcnt;
  uint64_t specialchars, nsort, largestbucket;
  bool counted;
  uint32_t sort_maxbucket;   // buckets above this size take the refinement path (testing knob)
  uint64_t twolevel_bytes;   // parts whose 4-byte suffix array exceeds this use the two-level insertion
  bool l2_window;            // persisting-L2 window set on the stream
  bool part_imaged;          // every range of the current part had its 64-bit image written by the sorters
  int fused_widen;           // 1 = direct path writes the 64-bit image from the sorters (GTB200_FUSED_WIDEN, testing knob)
  int pipeline;              // 1 = overlap D2H of finished slices with the sort of the next (GTB200_PIPELINE, testing knob)
  int narrow;                // 1 = a GtUword suftab crosses PCIe as 32-bit positions and is widened by host threads
                             //     (GTB200_NARROW; 0 copies the 64-bit image)
  int widen_threads;
  cudaStream_t copy_stream;
  cudaEvent_t slice_ev[8];
  cudaStream_t side_stream;  // small kernels that are independent of the main chain (special-padded suffixes) run beside it
  cudaEvent_t side_ev[2];
  int overlap;               // GTB200_OVERLAP=0 keeps everything on one stream (testing knob)
  int tile_mode;             // 0 off, 1 automatic, 2 forced (GTB200_TILE, testing knob)
  int tile_levels;           // 0 automatic, 1/2 forced (GTB200_TILE_LEVELS, testing knob)
  uint64_t *d_recB;          // second record buffer of the tile path (two partition levels); 16-byte records when keyed
  int keyed_mode;            // 0 never, 1 automatic (text larger than L2's share), 2 always (GTB200_KEYED, testing knob)
  uint64_t cap_suf64_bytes;
  int subsort;               // byte text, keyed tiles: split buckets by the first key symbol (GTB200_SUBSORT=0: testing knob)
  uint64_t diag[8];          // last part: [0] tile path, [1] partition levels, [2] keyed records, [3] suffixes in medium buckets
  uint32_t *d_isa;           // rank array of the prefix-doubling refinement (n+1 entries)
  uint64_t cap_isa;
  int doubling;              // 1 = prefix doubling when the part is the whole array (GTB200_DOUBLING, testing knob)
  int extend;                // text-window rounds of sharded parts: 2 = pivot rounds, 1 = skip shared windows, 0 = plain (GTB200_EXTEND)
  uint32_t *d_bins;          // bin cursors / starts / chunk table of the tile path
  uint64_t cap_recB, cap_bins;
  // per part
  uint64_t cap_width;
  uint32_t *d_sa;
  uint8_t *d_lcp;
  uint64_t *d_suf64;
  uint32_t *d_medlist;       // codes of buckets handed to the block-level sorter
  uint32_t *d_fbbits;
  uint64_t cap_fbwords;
  gtb200_largelcp *d_llv;    // sorted records of the last part
  uint64_t cap_llv;
  uint64_t part_offset, part_width, part_nllv;
  PartStats *d_stats, *h_stats;
  // scratch
  uint32_t *d_tilesums;
  uint64_t cap_tilesums;
  unsigned long long *d_scanstate;   // look-back words of the single-pass scan, [cap_scanstate] + ticket word
  uint64_t cap_scanstate;
  uint64_t scan_epoch;
  void *d_scratch;
  size_t cap_scratch;
  void *d_tail;              // entries of the last gtb200_tail call (own buffer: no other call recycles it)
  size_t cap_tail;
  std::vector<PoolBlock> blocks;   // every device allocation of this context
  // profiling
  bool prof_on;
  std::vector<ProfEvent> pending;
  std::vector<cudaEvent_t> evpool;
  double prof_ms[KC_NUM];
  uint64_t prof_launches[KC_NUM];
};

static int prof_fold(gtb200_ctx *c) {
  if (c->pending.empty()) return 0;
  CK(cudaEventSynchronize(c->pending.back().b));
  for (auto &p : c->pending) {
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, p.a, p.b));
    c->prof_ms[p.cls] += ms;
    c->evpool.push_back(p.a);
    c->evpool.push_back(p.b);
  }
  c->pending.clear();
  return 0;
}

static cudaEvent_t prof_event(gtb200_ctx *c) {
  if (!c->evpool.empty()) {
    cudaEvent_t e = c->evpool.back();
    c->evpool.pop_back();
    return e;
  }
  cudaEvent_t e;
  cudaEventCreate(&e);
  return e;
}

struct ProfScope {
  gtb200_ctx *c;
  ProfEvent ev;
  bool on;
  ProfScope(gtb200_ctx *c_, int cls, int launches = 1) : c(c_), on(c_->prof_on) {
    c->prof_launches[cls] += launches;
    if (on) {
      ev.cls = cls;
      ev.a = prof_event(c);
      ev.b = prof_event(c);
      cudaEventRecord(ev.a, c->stream);
    }
  }
  ~ProfScope() {
    if (on) {
      cudaEventRecord(ev.b, c->stream);
      c->pending.push_back(ev);
      if (c->pending.size() > 2048) prof_fold(c);
    }
  }
};

// ---------------------------------------------------------------------------------------------
// Device memory: every buffer comes from a process-wide cache of freed blocks, so that opening a
// context for the next sequence (or the next benchmark step) costs no cudaMalloc/cudaFree - both
// are expensive and cudaFree synchronises the device.  gtb200_release_cached_memory() empties it.
// ---------------------------------------------------------------------------------------------
static std::mutex g_pool_mu;
static std::vector<PoolBlock> g_pool[64];
static std::vector<void *> g_pinned_pool;
// page-locked ring of the narrow transfer (narrow_copy_run), cached like the device blocks
#define NR_RING 4
#define NR_CHUNK (16u << 20)          // entries per chunk: 64 MB in, 128 MB out
struct NarrowRing { uint32_t *buf[NR_RING]; };
static std::vector<NarrowRing> g_ring_pool;      // guarded by g_pool_mu
struct StreamSet { cudaStream_t main, copy, side; cudaEvent_t ev[8], sev[2]; };
static std::vector<StreamSet> g_stream_pool[64];   // streams and events of closed contexts, reused by the next open

static void pool_release_device(int dev) {   // caller holds g_pool_mu
  for (auto &b : g_pool[dev]) cudaFree(b.p);
  g_pool[dev].clear();
}

template <typename T>
static cudaError_t dmalloc(gtb200_ctx *c, T **out, size_t bytes) {
  bytes = (bytes + 511) & ~(size_t)511;
  if (bytes == 0) bytes = 512;
  void *p = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_pool_mu);
    auto &fl = g_pool[c->device];
    size_t best = (size_t)-1;
    for (size_t i = 0; i < fl.size(); ++i)
      if (fl[i].sz >= bytes && fl[i].sz <= bytes + bytes / 4 + (1u << 16) &&
          (best == (size_t)-1 || fl[i].sz < fl[best].sz))
        best = i;
    if (best != (size_t)-1) {
      p = fl[best].p;
      bytes = fl[best].sz;
      fl[best] = fl.back();
      fl.pop_back();
    }
  }
  if (!p) {
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) {                  // give cached blocks back to the driver and retry once
      (void)cudaGetLastError();
      {
        std::lock_guard<std::mutex> lock(g_pool_mu);
        pool_release_device(c->device);
      }
      e = cudaMalloc(&p, bytes);
      if (e != cudaSuccess) return e;
    }
  }
  c->blocks.push_back({p, bytes});
  *out = (T *)p;
  return cudaSuccess;
}

static void dfree(gtb200_ctx *c, void *p) {
  if (!p) return;
  if (c->stream) cudaStreamSynchronize(c->stream);   // the block may be handed to another context
  for (size_t i = 0; i < c->blocks.size(); ++i)
    if (c->blocks[i].p == p) {
      std::lock_guard<std::mutex> lock(g_pool_mu);
      g_pool[c->device].push_back(c->blocks[i]);
      c->blocks[i] = c->blocks.back();
      c->blocks.pop_back();
      return;
    }
}

extern "C" void gtb200_release_cached_memory(void) {
  std::lock_guard<std::mutex> lock(g_pool_mu);
  int cur = 0;
  cudaGetDevice(&cur);
  for (int d = 0; d < 64; ++d)
    if (!g_pool[d].empty()) {
      cudaSetDevice(d);
      pool_release_device(d);
    }
  for (void *h : g_pinned_pool) cudaFreeHost(h);
  g_pinned_pool.clear();
  for (auto &r : g_ring_pool)
    for (int i = 0; i < NR_RING; ++i) cudaFreeHost(r.buf[i]);
  g_ring_pool.clear();
  for (int d = 0; d < 64; ++d) {
    for (auto &ss : g_stream_pool[d]) {
      cudaStreamDestroy(ss.main);
      if (ss.copy) { for (int k = 0; k < 8; ++k) cudaEventDestroy(ss.ev[k]); cudaStreamDestroy(ss.copy); }
      if (ss.side) { for (int k = 0; k < 2; ++k) cudaEventDestroy(ss.sev[k]); cudaStreamDestroy(ss.side); }
    }
    g_stream_pool[d].clear();
  }
  cudaSetDevice(cur);
}

static inline unsigned grid_for(uint64_t n, unsigned per_block) {
  uint64_t g = (n + per_block - 1) / per_block;
  if (g == 0) g = 1;
  return (unsigned)g;
}

static int ensure_scratch(gtb200_ctx *c, size_t bytes) {
  if (bytes <= c->cap_scratch) return 0;
  dfree(c, c->d_scratch);
  c->d_scratch = nullptr;
  c->cap_scratch = 0;
  CK(dmalloc(c, &c->d_scratch, bytes));
  c->cap_scratch = bytes;
  return 0;
}

static int ensure_tilesums(gtb200_ctx *c, uint64_t ntiles) {
  if (ntiles <= c->cap_tilesums) return 0;
  dfree(c, c->d_tilesums);
  c->d_tilesums = nullptr;
  c->cap_tilesums = 0;
  CK(dmalloc(c, &c->d_tilesums, (ntiles + 1) * sizeof(uint32_t)));
  c->cap_tilesums = ntiles;
  return 0;
}

// exclusive scan on the context's stream (one launch); total (64-bit) lands in *d_total if non-null,
// max(in[]) is folded into *d_max (atomicMax, caller zeroes it) if non-null
static int device_scan(gtb200_ctx *c, const uint32_t *in, uint32_t *out, uint64_t n, int append_total,
                       unsigned long long *d_total, int cls = KC_SCAN, unsigned int *d_max = nullptr) {
  if (n == 0) {
    if (d_total) CK(cudaMemsetAsync(d_total, 0, sizeof *d_total, c->stream));
    return 0;
  }
  const int R = n <= (1u << 21) ? 2 : 16;
  const uint64_t tile = (uint64_t)SC_THREADS * R;
  const uint64_t ntiles = (n + tile - 1) / tile;
  if (ntiles > c->cap_scanstate || c->scan_epoch >= (1ull << 30) - 1) {
    if (ntiles > c->cap_scanstate) {
      dfree(c, c->d_scanstate);
      c->d_scanstate = nullptr;
      c->cap_scanstate = 0;
      const uint64_t cap = ntiles < 1024 ? 1024 : ntiles;
      CK(dmalloc(c, &c->d_scanstate, (cap + 1) * 8));
      c->cap_scanstate = cap;
    }
    CK(cudaMemsetAsync(c->d_scanstate, 0, (c->cap_scanstate + 1) * 8, c->stream));   // epoch 0 = never ready
    c->scan_epoch = 0;
  }
  const uint64_t epoch = ++c->scan_epoch;
  unsigned int *ticket = (unsigned int *)(c->d_scanstate + c->cap_scanstate);
  ProfScope ps(c, cls, 1);
  if (R == 2)
    k_scan_onepass<2><<<(unsigned)ntiles, SC_THREADS, 0, c->stream>>>(in, out, n, (uint32_t)ntiles, c->d_scanstate, ticket, epoch,
                                                                    append_total, d_total, d_max);
  else
    k_scan_onepass<16><<<(unsigned)ntiles, SC_THREADS, 0, c->stream>>>(in, out, n, (uint32_t)ntiles, c->d_scanstate, ticket, epoch,
                                                                     append_total, d_total, d_max);
  CK(cudaGetLastError());
  return 0;
}

// Stable LSD radix sort of n pairs on key bits [bit_lo, bit_hi).  keys_in/vals_in are left intact;
// (kA,vA) and (kB,vB) are ping-pong buffers, hist holds 256*nblocks+1 words.  Result pointers out.
static int device_radix_sort(gtb200_ctx *c, const uint64_t *keys_in, const uint32_t *vals_in, uint64_t *kA,
                             uint32_t *vA, uint64_t *kB, uint32_t *vB, uint32_t *hist, uint32_t n, int bit_lo,
                             int bit_hi, const uint64_t **keys_out, const uint32_t **vals_out) {
  const uint64_t *ksrc = keys_in;
  const uint32_t *vsrc = vals_in;
  uint64_t *kdst = kA;
  uint32_t *vdst = vA;
  const uint32_t nblocks = grid_for(n, RS_TILE);
  for (int shift = bit_lo; shift < bit_hi; shift += 8) {
    {
      ProfScope ps(c, KC_RADIX, 1);
      k_radix_hist<<<nblocks, RS_THREADS, 0, c->stream>>>(ksrc, n, shift, hist, nblocks);
    }
    if (device_scan(c, hist, hist, 256ull * nblocks, 0, nullptr, KC_RADIX)) return -1;
    {
      ProfScope ps(c, KC_RADIX, 1);
      k_radix_scatter<<<nblocks, RS_THREADS, 0, c->stream>>>(ksrc, vsrc, kdst, vdst, n, shift, hist, nblocks);
    }
    ksrc = kdst;
    vsrc = vdst;
    if (kdst == kA) { kdst = kB; vdst = vB; } else { kdst = kA; vdst = vA; }
  }
  CK(cudaGetLastError());
  *keys_out = ksrc;
  *vals_out = vsrc;
  return 0;
}

static inline int bits_for(uint64_t maxvalue) {
  int b = 1;
  while (b < 64 && (maxvalue >> b) != 0) ++b;
  return b;
}

// ---------------------------------------------------------------------------------------------
// open(): bitmap and packing kernels
// ---------------------------------------------------------------------------------------------

// mask words that lie entirely or partly at/after position n are forced to ones
__global__ void k_mask_tail(uint64_t *mask, uint64_t n, uint64_t nmaskwords) {
  const uint64_t i = (n >> 6) + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nmaskwords) return;
  if (i == (n >> 6)) mask[i] |= ~0ull << (n & 63);
  else mask[i] = ~0ull;
}

// one warp per special range: set bits [start,end)
__global__ void k_mask_from_ranges(uint64_t *mask, const gtb200_range *__restrict__ ranges, uint64_t nranges) {
  const uint64_t gw = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint64_t nw = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  const int lane = threadIdx.x & 31;
  for (uint64_t r = gw; r < nranges; r += nw) {
    const uint64_t s = ranges[r].start, e = ranges[r].end;   // e > s
    const uint64_t w0 = s >> 6, w1 = (e - 1) >> 6;
    for (uint64_t wi = w0 + lane; wi <= w1; wi += 32) {
      uint64_t bits = ~0ull;
      if (wi == w0) bits &= ~0ull << (s & 63);
      if (wi == w1) bits &= ~0ull >> (63 - ((e - 1) & 63));
      if (wi == w0 || wi == w1) atomicOr((unsigned long long *)&mask[wi], (unsigned long long)bits);
      else mask[wi] = ~0ull;   // interior words belong to this range alone
    }
  }
}

// bytes (one symbol per byte, 254/255 special) -> packed text words + mask; one thread per 64 positions
template <int BITS>
__global__ void k_pack_bytes(const uint8_t *__restrict__ enc, uint64_t n, uint64_t *__restrict__ text,
                             uint64_t *__restrict__ mask, uint64_t nchunks) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nchunks) return;
  constexpr int SPW = 64 / BITS;
  const uint64_t p0 = t << 6;
  uint64_t mw = 0;
  for (int wq = 0; wq < 64 / SPW; ++wq) {
    uint64_t word = 0;
    for (int k = 0; k < SPW; ++k) {
      const uint64_t p = p0 + (uint64_t)wq * SPW + k;
      uint8_t ch = p < n ? enc[p] : (uint8_t)255;
      const bool sp = ch >= 254;
      if (sp) { mw |= 1ull << (wq * SPW + k); ch = BITS == 8 ? 0xff : 0; }   // byte text keeps specials visible (common.cuh)
      word |= (uint64_t)ch << (64 - BITS * (k + 1));
    }
    text[t * (64 / SPW) + wq] = word;
  }
  mask[t] = mw;
}

// gather table (common.cuh): record i = {word i, word i+1}; BITS=2: bit 0 of the second word = any special among
// the record's first 59 positions
template <int BITS>
__global__ void k_build_gather(const uint64_t *__restrict__ text, const uint64_t *__restrict__ mask, uint64_t nrec,
                               ulonglong2 *__restrict__ g) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nrec) return;
  ulonglong2 r;
  r.x = text[i];
  r.y = text[i + 1];
  if (BITS == 2) {
    const uint64_t m0 = i >> 1;
    const uint64_t mw = (i & 1) ? (mask[m0] >> 32) | (mask[m0 + 1] << 32) : mask[m0];
    r.y = (r.y & ~1ull) | ((mw & ((1ull << 59) - 1ull)) ? 1ull : 0ull);
  }
  g[i] = r;
}

// Readmodes (core/readmode_api.h:24-27): the index of `-dir rev|cpl|rcl` is the index of the reversed and/or
// complemented sequence, so the device copy of the text is transformed once and everything else is unchanged.
// One thread per 64 output positions.
template <int BITS>
__global__ void k_readmode_transform(const uint64_t *__restrict__ text, const uint64_t *__restrict__ mask, uint64_t n,
                                     int reverse, int complement, uint64_t *__restrict__ otext,
                                     uint64_t *__restrict__ omask, uint64_t nchunks) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nchunks) return;
  constexpr int SPW = 64 / BITS;
  const uint64_t p0 = t << 6;
  uint64_t mw = 0;
  for (int wq = 0; wq < 64 / SPW; ++wq) {
    uint64_t word = 0;
    for (int k = 0; k < SPW; ++k) {
      const uint64_t p = p0 + (uint64_t)wq * SPW + k;
      if (p >= n) {
        if (BITS == 8) word |= 0xffull << (64 - BITS * (k + 1));
        continue;
      }
      const uint64_t src = reverse ? n - 1 - p : p;
      const bool sp = (mask[src >> 6] >> (src & 63)) & 1ull;
      uint64_t sym = (text[src / SPW] >> (64 - BITS * (src % SPW + 1))) & ((1ull << BITS) - 1);
      if (sp) { mw |= 1ull << (wq * SPW + k); sym = BITS == 8 ? 0xff : 0; }
      else if (complement) sym = 3 - sym;
      word |= sym << (64 - BITS * (k + 1));
    }
    otext[t * (64 / SPW) + wq] = word;
  }
  omask[t] = mw;
}

// range derivation from the bitmap (byte input): count run starts per mask word
__global__ void k_runs_count(const uint64_t *__restrict__ mask, uint64_t n, uint64_t nwords, uint32_t *cnt,
                             uint32_t *popc) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nwords) return;
  uint64_t mw = mask[t];
  const uint64_t lo = t << 6;
  if (lo + 64 > n) mw &= (n > lo) ? (~0ull >> (64 - (n - lo))) : 0ull;   // positions >= n are not ranges
  const uint64_t prevbit = t ? (mask[t - 1] >> 63) : 0ull;
  const uint64_t starts = mw & ~((mw << 1) | prevbit);
  cnt[t] = __popcll(starts);
  popc[t] = __popcll(mw);
}

__global__ void k_runs_write(const uint64_t *__restrict__ mask, uint64_t n, uint64_t nwords,
                             const uint32_t *__restrict__ off, gtb200_range *ranges) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nwords) return;
  uint64_t mw = mask[t];
  const uint64_t lo = t << 6;
  if (lo + 64 > n) mw &= (n > lo) ? (~0ull >> (64 - (n - lo))) : 0ull;
  const uint64_t prevbit = t ? (mask[t - 1] >> 63) : 0ull;
  uint64_t starts = mw & ~((mw << 1) | prevbit);
  uint32_t o = off[t];
  while (starts) {
    const int b = __ffsll((long long)starts) - 1;
    starts &= starts - 1;
    // walk to the end of the run (runs are long only for N-blocks; bounded by n)
    uint64_t p = lo + b, wi = t;
    uint64_t cur = ~mw >> b;    // zero bits mark the end, relative to b
    uint64_t e;
    for (;;) {
      if (cur) { e = p + (__ffsll((long long)cur) - 1); break; }
      // run continues into the next word
      wi++;
      p = wi << 6;
      if (p >= n) { e = n; break; }
      uint64_t nm = mask[wi];
      if (p + 64 > n) nm &= ~0ull >> (64 - (n - p));
      cur = ~nm;
    }
    if (e > n) e = n;
    ranges[o].start = lo + b;
    ranges[o].end = e;
    o++;
  }
}

__global__ void k_range_lengths(const gtb200_range *__restrict__ ranges, uint64_t nranges, uint32_t *len) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nranges) len[t] = (uint32_t)(ranges[t].end - ranges[t].start);
}

// ---------------------------------------------------------------------------------------------
// phase 1: counting
// ---------------------------------------------------------------------------------------------

// One thread per 32 consecutive positions.  Counts suffixes whose first pl symbols are all real.
// MODE 0: histogram (hist[code]++).  MODE 1: scatter (sa[cursor[code]++ - partoffset] = pos) for
// codes in [mincode,maxcode].
#define KM_PPT 8     // positions per thread: four threads share one 32-base word (the loads hit L1), which gives
                     // small genomes enough threads to cover the latency of the returning atomics
template <int BITS, int MODE>
__global__ void __launch_bounds__(256) k_kmers(Text<BITS> T, CodeParams cp, uint64_t pbase, uint64_t n,
                                                uint32_t *__restrict__ table, uint32_t *__restrict__ sa,
                                                uint64_t mincode, uint64_t maxcode, uint32_t partoffset) {
  // positions [pbase, n): pbase is a multiple of 32; n is the end of the text or of this rank's position shard
  // (a shard end is a multiple of 64, so no thread straddles it)
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t p0 = pbase + t * KM_PPT;
  if (p0 >= n) return;
  const unsigned pl = cp.pl;
  const uint64_t mw = T.maskwin(p0);
  const uint64_t plmask = (pl >= 64) ? ~0ull : ((1ull << pl) - 1ull);
  uint64_t hi = 0, lo = 0;
  const int i0 = (int)(p0 & 31);
  if (BITS == 2) { hi = __ldg(T.w + (p0 >> 5)); lo = __ldg(T.w + (p0 >> 5) + 1); }
  uint32_t slot[KM_PPT];
  bool ok[KM_PPT];
#pragma unroll
  for (int k = 0; k < KM_PPT; ++k) {
    ok[k] = ((mw >> k) & plmask) == 0;
    slot[k] = 0;
    if (ok[k]) {
      uint64_t win;
      const int i = i0 + k;
      if (BITS == 2) win = i ? (hi << (2 * i)) | (lo >> (64 - 2 * i)) : hi;
      else win = T.window(p0 + k);
      const uint64_t code = prefix_value<BITS>(win, pl, cp);
      if (MODE == 0) {
        atomicAdd(&table[code], 1u);
      } else if (MODE == 2) {
        if (code >= mincode && code <= maxcode) atomicAdd(&table[code], 1u);
      } else {
        ok[k] = code >= mincode && code <= maxcode;
        if (ok[k]) slot[k] = atomicAdd(&table[code], 1u);      // all atomics of the thread are issued before
      }                                                        // the first dependent store below
    }
  }
  if (MODE == 1) {
#pragma unroll
    for (int k = 0; k < KM_PPT; ++k)
      if (ok[k]) sa[slot[k] - partoffset] = (uint32_t)(p0 + k);
  }
}

// Counting for large DNA texts: k_kmers spends ~37 instructions per position on 64-bit variable shifts, and a table
// that is filled in code slices pays them once per slice (3.1 Gbp: 4 passes, issue-active 72 %: instruction-, not
// atomic-bound).  Here a thread owns 16 positions that start on a 32-bit boundary of the packed text, so with the
// loop unrolled every code is one funnel shift by a constant plus a shift; ~6 instructions per position.
// sliced: only codes in [lo, hi] are counted in this pass.
__global__ void __launch_bounds__(256) k_count_dna(Text<2> T, unsigned pl, uint64_t pbase, uint64_t pend,
                                                   uint32_t *__restrict__ table, uint32_t lo, uint32_t hi) {
  const uint64_t p0 = pbase + ((uint64_t)blockIdx.x * 256 + threadIdx.x) * 16;
  if (p0 >= pend) return;
  const uint64_t t = p0 >> 5;
  const uint64_t w0 = __ldg(T.w + t), w1 = __ldg(T.w + t + 1);
  const uint32_t u0 = (p0 & 16) ? (uint32_t)w0 : (uint32_t)(w0 >> 32);
  const uint32_t u1 = (p0 & 16) ? (uint32_t)(w1 >> 32) : (uint32_t)w0;
  const uint64_t mw = T.maskwin(p0);
  const unsigned cs = 32u - 2u * pl;                  // pl <= 15
  const uint32_t span = hi - lo;
  const uint32_t plmask = (1u << pl) - 1u;
  const uint32_t m32 = (uint32_t)mw;                  // positions p0 .. p0+31: covers i + pl - 1 <= 29
  if ((m32 & ((1u << (15 + pl)) - 1u)) == 0u) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const uint32_t code = (i ? __funnelshift_l(u1, u0, 2 * i) : u0) >> cs;
      if (code - lo <= span) atomicAdd(&table[code], 1u);
    }
  } else {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const uint32_t code = (i ? __funnelshift_l(u1, u0, 2 * i) : u0) >> cs;
      if (((m32 >> i) & plmask) == 0u && code - lo <= span) atomicAdd(&table[code], 1u);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Two-level insertion for parts whose suffix array does not fit L2.  A direct scatter writes 4
// bytes to a random place in a multi-GB array: every store evicts a partially filled sector.
// Instead k_partition first groups (code,pos) records by the leading bits of the code into <= 1024
// coarse bins (block-local histogram in shared memory, one global reservation per block and bin,
// so each bin is written as a stream), and k_scatter_records then walks the records in order:
// neighbouring records now hit a 64 KB slice of the cursor table and a few MB of the suffix array,
// both L2-resident, where the 4-byte stores are combined into full sectors.
// ---------------------------------------------------------------------------------------------
#define PT_MAXBINS 2048

template <int BITS>
__global__ void __launch_bounds__(256) k_partition(Text<BITS> T, CodeParams cp, uint64_t n, uint64_t mincode,
                                                    uint64_t maxcode, int shift, uint32_t nbins,
                                                    uint32_t *__restrict__ gcursor, uint64_t *__restrict__ rec) {
  __shared__ uint32_t h[PT_MAXBINS];
  for (uint32_t i = threadIdx.x; i < nbins; i += 256) h[i] = 0;
  __syncthreads();
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t p0 = t << 5;
  const unsigned pl = cp.pl;
  const uint64_t plmask = (1ull << pl) - 1ull;
  uint64_t mw = ~0ull, hi = 0, lo = 0;
  if (p0 < n) {
    mw = T.maskwin(p0);
    if (BITS == 2) { hi = __ldg(T.w + t); lo = __ldg(T.w + t + 1); }
  }
  uint32_t rk[16];   // rank of each of the 32 positions inside its bin (block-local, < 8192): two per word
#pragma unroll
  for (int i = 0; i < 16; ++i) rk[i] = 0;
  // pass 0: block-local rank = old value of the shared-memory bin counter
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    if (p0 < n && ((mw >> i) & plmask) == 0) {
      uint64_t win;
      if (BITS == 2) win = i ? (hi << (2 * i)) | (lo >> (64 - 2 * i)) : hi;
      else win = T.window(p0 + i);
      const uint64_t code = prefix_value<BITS>(win, pl, cp);
      if (code >= mincode && code <= maxcode) {
        const uint32_t r = atomicAdd(&h[(uint32_t)((code - mincode) >> shift)], 1u);
        rk[i >> 1] |= r << ((i & 1) * 16);
      }
    }
  }
  __syncthreads();
  // one global reservation per (block, bin); h[] becomes the bin's base offset for this block
  for (uint32_t i = threadIdx.x; i < nbins; i += 256) {
    const uint32_t cnt = h[i];
    h[i] = cnt ? atomicAdd(&gcursor[i], cnt) : 0u;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    if (p0 < n && ((mw >> i) & plmask) == 0) {
      uint64_t win;
      if (BITS == 2) win = i ? (hi << (2 * i)) | (lo >> (64 - 2 * i)) : hi;
      else win = T.window(p0 + i);
      const uint64_t code = prefix_value<BITS>(win, pl, cp);
      if (code >= mincode && code <= maxcode) {
        const uint32_t r = (rk[i >> 1] >> ((i & 1) * 16)) & 0xffffu;
        rec[h[(uint32_t)((code - mincode) >> shift)] + r] = (code << 32) | (uint64_t)(uint32_t)(p0 + i);
      }
    }
  }
}

// record offsets of the coarse bins = suffix-array offsets of their first bucket
__global__ void k_partition_init(const uint32_t *__restrict__ lb, uint64_t mincode, int shift, uint32_t nbins,
                                 uint32_t partoffset, uint32_t *gcursor) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < nbins) gcursor[j] = lb[mincode + ((uint64_t)j << shift)] - partoffset;
}

// slots a bin did not fill (its special-padded members are inserted separately) are marked empty
__global__ void k_partition_gaps(const uint32_t *__restrict__ gcursor, const uint32_t *__restrict__ lb, uint64_t mincode,
                                 uint64_t maxcode, int shift, uint32_t nbins, uint32_t partoffset, uint64_t *rec) {
  const uint32_t j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (j >= nbins) return;
  uint64_t next = mincode + ((uint64_t)(j + 1) << shift);
  if (next > maxcode + 1) next = maxcode + 1;
  const uint32_t end = lb[next] - partoffset;
  for (uint32_t i = gcursor[j] + lane; i < end; i += 32) rec[i] = ~0ull;
}

// second level with the bin's slice of the cursor table in shared memory: one block per coarse bin
__global__ void __launch_bounds__(1024) k_scatter_bins(const uint64_t *__restrict__ rec, const uint32_t *__restrict__ gcursor,
                                                        const uint32_t *__restrict__ lb, uint32_t *__restrict__ cursor,
                                                        uint32_t *__restrict__ sa, uint64_t mincode, uint64_t maxcode,
                                                        int shift, uint32_t partoffset) {
  extern __shared__ uint32_t scur[];
  const uint32_t j = blockIdx.x;
  const uint64_t c0 = mincode + ((uint64_t)j << shift);
  uint64_t c1 = c0 + (1ull << shift);
  if (c1 > maxcode + 1) c1 = maxcode + 1;
  const uint32_t ncf = (uint32_t)(c1 - c0);
  for (uint32_t i = threadIdx.x; i < ncf; i += blockDim.x) scur[i] = lb[c0 + i];
  __syncthreads();
  const uint32_t rbeg = lb[c0] - partoffset, rend = gcursor[j];
  for (uint32_t i = rbeg + threadIdx.x; i < rend; i += blockDim.x) {
    const uint64_t r = rec[i];
    const uint32_t slot = atomicAdd(&scur[(uint32_t)((r >> 32) - c0)], 1u);
    sa[slot - partoffset] = (uint32_t)r;
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < ncf; i += blockDim.x) cursor[c0 + i] = scur[i];
}

__global__ void __launch_bounds__(256) k_scatter_records(const uint64_t *__restrict__ rec, uint64_t width,
                                                          uint32_t *__restrict__ cursor, uint32_t *__restrict__ sa,
                                                          uint32_t partoffset) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= width) return;
  const uint64_t r = rec[i];
  if (r == ~0ull) return;
  const uint32_t slot = atomicAdd(&cursor[r >> 32], 1u);
  sa[slot - partoffset] = (uint32_t)r;
}

// Special-padded suffixes: for every special range (and the virtual end at n) the up to pl-1
// positions directly left of it.  Thread k handles range k (k == nranges: the end of the text).
// PASS 0: number of such suffixes per range.  PASS 1: counters + list entries at off[k].
template <int BITS, int PASS>
__global__ void k_padded(Text<BITS> T, CodeParams cp, uint64_t n, const gtb200_range *__restrict__ ranges,
                         uint64_t nranges, uint32_t *__restrict__ cntoff, uint32_t *__restrict__ hist,
                         uint32_t *__restrict__ csc, uint32_t *__restrict__ dist, uint64_t *__restrict__ padkey,
                         uint32_t *__restrict__ padpos) {
  const uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k > nranges) return;
  const uint64_t s = k < nranges ? ranges[k].start : n;
  const uint64_t prev_end = k ? ranges[k - 1].end : 0;
  const unsigned pl = cp.pl;
  uint64_t avail = s - prev_end;
  if (k == nranges && nranges && ranges[nranges - 1].end >= n) avail = 0;
  const unsigned cnt = (unsigned)(avail < pl - 1 ? avail : pl - 1);
  if (PASS == 0) { cntoff[k] = cnt; return; }
  const uint32_t off = cntoff[k];
  for (unsigned j = 1; j <= cnt; ++j) {         // suffix s-j has exactly j real symbols
    const uint64_t p = s - j;
    const uint64_t pv = prefix_value<BITS>(T.window(p), j, cp);
    const uint64_t fill = cp.pw[pl - j] - 1;    // pl-j digits of value sigma-1
    const uint64_t code = pv * (fill + 1) + fill;
    if (hist) {
      atomicAdd(&hist[code], 1u);
      atomicAdd(&csc[(code - (cp.sigma - 1)) / cp.sigma], 1u);
      if (j + 1 < pl) {
        uint64_t doff = 0;
        for (unsigned i = 1; i < j; ++i) doff += cp.pw[i];
        atomicAdd(&dist[doff + pv], 1u);
      }
    }
    padkey[off + (cnt - j)] = (code << 6) | (uint64_t)(pl - j);
    padpos[off + (cnt - j)] = (uint32_t)p;
  }
}

// Few special ranges (a genome with a handful of N-runs, or none): enumerate, count and sort the
// special-padded suffixes in ONE single-block launch instead of two passes + scan + radix passes.
#define PADSMALL_CAP 2048
template <int BITS>
__global__ void __launch_bounds__(256) k_padded_small(Text<BITS> T, CodeParams cp, uint64_t n,
                                                      const gtb200_range *__restrict__ ranges, uint64_t nranges,
                                                      uint32_t *__restrict__ hist, uint32_t *__restrict__ csc,
                                                      uint32_t *__restrict__ dist, uint64_t *__restrict__ padkey,
                                                      uint32_t *__restrict__ padpos, unsigned long long *npadded) {
  __shared__ uint64_t skey[PADSMALL_CAP];
  __shared__ uint32_t spos[PADSMALL_CAP];
  __shared__ uint32_t sm[8];
  const unsigned pl = cp.pl;
  uint32_t carry = 0;
  for (uint64_t base = 0; base <= nranges; base += 256) {
    const uint64_t k = base + threadIdx.x;
    unsigned cnt = 0;
    uint64_t s0 = 0;
    if (k <= nranges) {
      s0 = k < nranges ? ranges[k].start : n;
      const uint64_t prev_end = k ? ranges[k - 1].end : 0;
      const uint64_t avail = s0 - prev_end;
      cnt = (unsigned)(avail < pl - 1 ? avail : pl - 1);
    }
    uint32_t tot;
    const uint32_t off = carry + block_excl_scan256(cnt, &tot, sm);
    for (unsigned j = 1; j <= cnt; ++j) {
      const uint64_t p = s0 - j;
      const uint64_t pv = prefix_value<BITS>(T.window(p), j, cp);
      const uint64_t fill = cp.pw[pl - j] - 1;
      const uint64_t code = pv * (fill + 1) + fill;
      if (hist) {                              // null on ranks that only need the list (sharded counting)
        atomicAdd(&hist[code], 1u);
        atomicAdd(&csc[(code - (cp.sigma - 1)) / cp.sigma], 1u);
        if (j + 1 < pl) {
          uint64_t doff = 0;
          for (unsigned i = 1; i < j; ++i) doff += cp.pw[i];
          atomicAdd(&dist[doff + pv], 1u);
        }
      }
      skey[off + (cnt - j)] = (code << 6) | (uint64_t)(pl - j);
      spos[off + (cnt - j)] = (uint32_t)p;
    }
    carry += tot;
  }
  const uint32_t total = carry;
  uint32_t P = 1;
  while (P < total) P <<= 1;
  __syncthreads();
  for (uint32_t i = total + threadIdx.x; i < P; i += 256) { skey[i] = ~0ull; spos[i] = 0xffffffffu; }
  __syncthreads();
  for (uint32_t size = 2; size <= P; size <<= 1) {
    for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
      for (uint32_t i = threadIdx.x; i < P / 2; i += 256) {
        const uint32_t lo = 2 * i - (i & (stride - 1)), hi = lo + stride;
        const bool asc = ((lo & size) == 0);
        const uint64_t ka = skey[lo], kb = skey[hi];
        const uint32_t pa = spos[lo], pb = spos[hi];
        const bool gt = ka > kb || (ka == kb && pa > pb);
        if (gt == asc) { skey[lo] = kb; skey[hi] = ka; spos[lo] = pb; spos[hi] = pa; }
      }
      __syncthreads();
    }
  }
  for (uint32_t i = threadIdx.x; i < total; i += 256) { padkey[i] = skey[i]; padpos[i] = spos[i]; }
  if (threadIdx.x == 0) *npadded = total;
}


__global__ void k_sum_u32(const uint32_t *__restrict__ a, uint64_t n, unsigned long long *out) {
  unsigned long long s = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    s += a[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(GTB_FULL, s, o);
  if ((threadIdx.x & 31) == 0 && s) atomicAdd(out, s);
}

// ---------------------------------------------------------------------------------------------
// LCP emission
// ---------------------------------------------------------------------------------------------
struct PartOut {
  uint32_t *sa;
  uint64_t *sa64;            // non-null: finishers also write the 64-bit image (.suf format), k_widen is skipped
  uint8_t *lcp;
  gtb200_largelcp *llv;      // unordered append buffer (refinement only)
  PartStats *stats;
  uint32_t partoffset;
};

// final suffix-array entry: 32-bit working array and, on the direct path, the 64-bit image in the same sweep
__device__ __forceinline__ void put_sa(const PartOut &o, uint32_t r, uint32_t p) {
  __stcs(o.sa + r, p);
  if (o.sa64) {
    __stcs(o.sa64 + r, (uint64_t)p);
    if (p == 0) o.stats->longest = (uint64_t)o.partoffset + r;
  }
}

__device__ __forceinline__ void emit_large(const PartOut &o, uint32_t r, uint32_t value) {
  const unsigned long long slot = atomicAdd(&o.stats->llv_count, 1ull);
  o.llv[slot].position = (uint64_t)o.partoffset + r;
  o.llv[slot].value = value;
}

// ---------------------------------------------------------------------------------------------
// phase 2: padded placement, bucket sort
// ---------------------------------------------------------------------------------------------

// padded list is sorted by (code, pl-L) with ascending positions inside equal keys: entry i of
// code c goes to the right end of bucket c (reference order: src/match/sfx-suffixer.c:476-516)
__global__ void k_place_padded(const uint64_t *__restrict__ padkey, const uint32_t *__restrict__ padpos,
                               const unsigned long long *__restrict__ npadded_dev, const uint32_t *__restrict__ lb,
                               const uint32_t *__restrict__ csc,
                               unsigned sigma, unsigned pl, uint64_t mincode, uint64_t maxcode, PartOut out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= *npadded_dev) return;
  const uint64_t key = padkey[i];
  const uint64_t code = key >> 6;
  if (code < mincode || code > maxcode) return;
  // first list index with this code
  uint64_t lo = 0, hi = i;
  const uint64_t target = code << 6;
  while (lo < hi) {
    const uint64_t mid = (lo + hi) >> 1;
    if (padkey[mid] < target) lo = mid + 1; else hi = mid;
  }
  const uint32_t npad = csc[(code - (sigma - 1)) / sigma];
  const uint32_t r = lb[code + 1] - npad + (uint32_t)(i - lo) - out.partoffset;
  const uint32_t L = pl - (uint32_t)(key & 63);
  put_sa(out, r, padpos[i]);
  out.lcp[r] = (uint8_t)L;               // own number of real symbols (src/match/sfx-lcpvalues.c:125-215);
  if (r + out.partoffset > lb[code])     // a bucket-first entry is overwritten by k_boundary
    atomicMax(&out.stats->maxdepth, L);
}

template <int E>
struct WarpSorter {
  __device__ static __forceinline__ bool greater(uint64_t ka, uint32_t va, uint64_t kb, uint32_t vb) {
    return ka > kb || (ka == kb && va > vb);
  }
  // bitonic sort of 32*E (key,val) pairs held as element index = e*32 + lane
  __device__ static __forceinline__ void sort(uint64_t (&k)[E], uint32_t (&v)[E], int lane) {
#pragma unroll
    for (int size = 2; size <= 32 * E; size <<= 1) {
#pragma unroll
      for (int stride = size >> 1; stride > 0; stride >>= 1) {
        if (stride >= 32) {
          const int es = stride >> 5;
#pragma unroll
          for (int e = 0; e < E; ++e) {
            if ((e & es) == 0) {
              const bool asc = (((e * 32) & size) == 0);
              const bool gt = greater(k[e], v[e], k[e | es], v[e | es]);
              if (gt == asc) {
                const uint64_t tk = k[e]; k[e] = k[e | es]; k[e | es] = tk;
                const uint32_t tv = v[e]; v[e] = v[e | es]; v[e | es] = tv;
              }
            }
          }
        } else {
#pragma unroll
          for (int e = 0; e < E; ++e) {
            const uint64_t ok = __shfl_xor_sync(GTB_FULL, k[e], stride);
            const uint32_t ov = __shfl_xor_sync(GTB_FULL, v[e], stride);
            const bool asc = (((e * 32 + lane) & size) == 0);
            const bool lower = ((lane & stride) == 0);
            const bool mine_gt = greater(k[e], v[e], ok, ov);
            const bool take = (lower == asc) ? mine_gt : !mine_gt;
            if (take) { k[e] = ok; v[e] = ov; }
          }
        }
      }
    }
  }
};

// Sorts one bucket of cnt <= 32*E regular suffixes held by one warp and writes order + in-bucket
// LCP.  Returns false (nothing written) if two suffixes agree on all pl+W symbols.
template <int BITS, int E>
__device__ __forceinline__ bool warp_sort_bucket(const Text<BITS> &T, const PartOut &out, const uint32_t *src,
                                                 const uint32_t *pre, const uint8_t *srcL, const uint64_t *pkeys,
                                                 uint32_t start, uint32_t cnt, unsigned pl, int lane,
                                                 unsigned long long &sum, uint32_t &mx) {
  uint64_t k[E];
  uint32_t v[E];
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const uint32_t i = e * 32 + lane;
    if (i < cnt) {
      v[e] = pre ? pre[e] : src[i];
      k[e] = pkeys ? pkeys[i]
                   : srcL ? make_key_known<BITS>(T, (uint64_t)v[e] + pl, (unsigned)srcL[i] - pl)
                          : make_key<BITS>(T, (uint64_t)v[e] + pl);
    } else {
      v[e] = 0xffffffffu;
      k[e] = ~0ull;
    }
  }
  WarpSorter<E>::sort(k, v, lane);
  uint32_t l[E];
  bool tp[E];                      // tied with the predecessor on all pl+W symbols: order and lcp still open
  bool tie = false;
#pragma unroll
  for (int e = 0; e < E; ++e) {
    uint64_t pk = __shfl_up_sync(GTB_FULL, k[e], 1);
    const uint64_t last = (e > 0) ? __shfl_sync(GTB_FULL, k[e > 0 ? e - 1 : 0], 31) : 0ull;
    if (lane == 0) pk = last;
    const uint32_t i = e * 32 + lane;
    l[e] = 0;
    tp[e] = false;
    if (i >= 1 && i < cnt) {
      tp[e] = pk == k[e] && (k[e] & 0xff) == 0;
      tie |= tp[e];
      l[e] = pl + key_common<BITS>(pk, k[e]);
    }
  }
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const uint32_t i = e * 32 + lane;
    if (i < cnt) {
      put_sa(out, start + i, v[e]);
      if (i >= 1) {
        __stcs(out.lcp + start + i, (uint8_t)l[e]);      // provisional for tied pairs, rewritten by refinement
        if (!tp[e]) {
          sum += l[e];
          mx = max(mx, l[e]);
        }
      }
    }
  }
  return !__any_sync(GTB_FULL, tie);   // false: sorted on pl+W symbols, tie groups go to the refinement path
}

// 32-bit variant: keys are (12 leading bases of the window << 8) | slot; min/max network.
// Ascending-only form of the bitonic sorter (each merge starts with a flip, partner = index ^ (size-1), then
// disperses with partner = index ^ stride; the lower index always keeps the minimum).  Padding keys
// (0xffffffff at the top indices) therefore never move, and a bucket of 65..96 suffixes runs the 128-wide
// network on E = 3 registers per lane, skipping every exchange with the absent fourth.
template <int E>
struct WarpSorter32 {
  __device__ static __forceinline__ void sort(uint32_t (&k)[E], int lane) {
    constexpr int N = E <= 1 ? 32 : E <= 2 ? 64 : 128;
#pragma unroll
    for (int size = 2; size <= N; size <<= 1) {
      if (size <= 32) {
        const bool keepmin = (lane & (size >> 1)) == 0;
#pragma unroll
        for (int e = 0; e < E; ++e) {
          const uint32_t o = __shfl_xor_sync(GTB_FULL, k[e], size - 1);
          k[e] = keepmin ? min(k[e], o) : max(k[e], o);
        }
      } else {
        const int me = (size >> 5) - 1;                 // register partner: e ^ 1 (size 64), e ^ 3 (size 128)
        uint32_t o[E];
#pragma unroll
        for (int e = 0; e < E; ++e) {
          const int pe = e ^ me;
          o[e] = pe < E ? __shfl_xor_sync(GTB_FULL, k[pe < E ? pe : 0], 31) : 0xffffffffu;
        }
#pragma unroll
        for (int e = 0; e < E; ++e) {
          const int pe = e ^ me;
          if (pe < E) k[e] = e < pe ? min(k[e], o[e]) : max(k[e], o[e]);
        }
      }
#pragma unroll
      for (int stride = size >> 2; stride > 0; stride >>= 1) {
        if (stride >= 32) {
          const int es = stride >> 5;
#pragma unroll
          for (int e = 0; e < E; ++e) {
            if ((e & es) == 0 && (e | es) < E) {
              const uint32_t lo = min(k[e], k[e | es]), hi = max(k[e], k[e | es]);
              k[e] = lo;
              k[e | es] = hi;
            }
          }
        } else {
          const bool keepmin = (lane & stride) == 0;
#pragma unroll
          for (int e = 0; e < E; ++e) {
            const uint32_t o = __shfl_xor_sync(GTB_FULL, k[e], stride);
            k[e] = keepmin ? min(k[e], o) : max(k[e], o);
          }
        }
      }
    }
  }
  // the last log2(32 E) steps of a merge (strides 16 E .. 1), for a chunk whose wider steps ran elsewhere
  __device__ static __forceinline__ void tail(uint32_t (&k)[E], int lane) {
#pragma unroll
    for (int stride = 16 * E; stride > 0; stride >>= 1) {
      if (stride >= 32) {
        const int es = stride >> 5;
#pragma unroll
        for (int e = 0; e < E; ++e) {
          if ((e & es) == 0 && (e | es) < E) {
            const uint32_t lo = min(k[e], k[e | es]), hi = max(k[e], k[e | es]);
            k[e] = lo;
            k[e | es] = hi;
          }
        }
      } else {
        const bool keepmin = (lane & stride) == 0;
#pragma unroll
        for (int e = 0; e < E; ++e) {
          const uint32_t o = __shfl_xor_sync(GTB_FULL, k[e], stride);
          k[e] = keepmin ? min(k[e], o) : max(k[e], o);
        }
      }
    }
  }
};

// 20 leading key bits: 10 bases, or 4 symbols squeezed to 5 bits each (values >= 31, padding included, collapse
// to 31: a monotone map, so the order of the prefixes never contradicts the order of the keys)
template <int BITS>
__device__ __forceinline__ uint32_t key_prefix20(uint64_t key) {
  if (BITS == 2) return (uint32_t)(key >> 44);
  uint32_t r = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint32_t b = (uint32_t)(key >> (56 - 8 * j)) & 0xffu;
    r = (r << 5) | (b < 31u ? b : 31u);
  }
  return r;
}

#define GTB_MAXSLICES 8
#define SORT_MAXBUCKET 128   // warp path: 4 elements per lane keep the kernel at 64 registers (32 warps/SM)
#define SORT_WARPS 8

// Fast path: order the bucket by the 12 bases that follow the prefix (one 32-bit compare per
// network step), full 64-bit keys parked in shared memory and fetched by slot afterwards.  The
// order is then verified with the full keys; any pair it does not strictly separate (equal 12-mers
// whose rest is out of order, ties, cut-off suffixes) sends the bucket to the exact 64-bit sorter.
// Returns 0 = done, 1 = redo with warp_sort_bucket.
template <int BITS, int E>
__device__ __forceinline__ int warp_sort_bucket32(const Text<BITS> &T, const PartOut &out, const uint32_t *src,
                                                  const uint32_t *pre, const uint8_t *srcL, const uint64_t *pkeys,
                                                  uint32_t start, uint32_t cnt, unsigned pl, int lane, uint64_t *skey_w,
                                                  uint32_t *spos_w, unsigned long long &sum, uint32_t &mx, int keyskip) {
  // pkeys: the bucket's full keys are already in shared memory next to its positions (src); nothing is staged
  // byte text: the network orders by four symbols squeezed to 5 bits each, after skipping keyskip leading key bits
  // (a group whose first key symbol is known to be equal skips it)
  const uint64_t *skey = pkeys ? pkeys : skey_w;
  const uint32_t *spos = pkeys ? src : spos_w;
  uint32_t k[E];
  if (!pkeys) __syncwarp();
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const uint32_t i = e * 32 + lane;
    k[e] = 0xffffffffu;
    if (i < cnt) {
      uint64_t fk;
      if (pkeys) fk = pkeys[i];
      else {
        const uint32_t p = pre ? pre[e] : src[i];
        fk = srcL ? make_key_known<BITS>(T, (uint64_t)p + pl, (unsigned)srcL[i] - pl)
                  : make_key<BITS>(T, (uint64_t)p + pl);
        skey_w[i] = fk;
        spos_w[i] = p;
      }
      k[e] = BITS == 2 ? (((uint32_t)(fk >> 40) << 8) | i) : ((key_prefix20<BITS>(fk << keyskip) << 8) | i);
    }
  }
  if (!pkeys) __syncwarp();
  WarpSorter32<E>::sort(k, lane);
  uint64_t fk[E];
  uint32_t l[E];
  bool bad = false;
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const uint32_t i = e * 32 + lane;
    fk[e] = (i < cnt) ? skey[k[e] & 0xff] : ~0ull;
  }
#pragma unroll
  for (int e = 0; e < E; ++e) {
    uint64_t pk = __shfl_up_sync(GTB_FULL, fk[e], 1);
    const uint64_t last = (e > 0) ? __shfl_sync(GTB_FULL, fk[e > 0 ? e - 1 : 0], 31) : 0ull;
    if (lane == 0) pk = last;
    const uint32_t i = e * 32 + lane;
    l[e] = 0;
    if (i >= 1 && i < cnt) {
      if (pk >= fk[e]) bad = true;
      l[e] = pl + key_common<BITS>(pk, fk[e]);
    }
  }
  if (__any_sync(GTB_FULL, bad)) return 1;
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const uint32_t i = e * 32 + lane;
    if (i < cnt) {
      put_sa(out, start + i, spos[k[e] & 0xff]);
      if (i >= 1) {
        __stcs(out.lcp + start + i, (uint8_t)l[e]);
        sum += l[e];
        mx = max(mx, l[e]);
      }
    }
  }
  return 0;
}

template <int BITS, int E>
__device__ __forceinline__ bool warp_sort_any(const Text<BITS> &T, const PartOut &out, const uint32_t *src,
                                              const uint32_t *pre, const uint8_t *srcL, const uint64_t *pkeys,
                                              uint32_t start, uint32_t cnt, unsigned pl, int lane, uint64_t *skey,
                                              uint32_t *spos, unsigned long long &sum, uint32_t &mx, int keyskip = -1) {
  // pre: the bucket's positions already in registers (element e*32+lane in pre[e]), loaded one bucket ahead
  // pkeys: the bucket's keys already in shared memory (element i in pkeys[i], its position in src[i])
  // keyskip >= 0 (byte text with the keys in shared memory): 32-bit network first, like DNA
  if (BITS == 2 || (keyskip >= 0 && pkeys != nullptr)) {
    if (warp_sort_bucket32<BITS, E>(T, out, src, pre, srcL, pkeys, start, cnt, pl, lane, skey, spos, sum, mx,
                                    keyskip < 0 ? 0 : keyskip) == 0) return true;
  }
  return warp_sort_bucket<BITS, (E == 3 ? 4 : E)>(T, out, src, pre, srcL, pkeys, start, cnt, pl, lane, sum, mx);
}

template <int BITS>
__global__ void __launch_bounds__(256) k_sort_buckets(Text<BITS> T, PartOut out, const uint32_t *__restrict__ lb,
                                                      const uint32_t *__restrict__ cursor, uint64_t mincode,
                                                      uint64_t maxcode, unsigned pl, uint32_t *__restrict__ fbbits,
                                                      uint32_t maxbucket, uint32_t *__restrict__ medlist,
                                                      uint32_t medmax, int only_leftovers, uint64_t tieoff) {
  if (only_leftovers && out.stats->tile_leftover == 0) return;   // the tile sorter finished every bucket
  __shared__ uint64_t s_key[SORT_WARPS][SORT_MAXBUCKET];
  __shared__ uint32_t s_pos[SORT_WARPS][SORT_MAXBUCKET];
  const int lane = threadIdx.x & 31;
  uint64_t *skey = s_key[threadIdx.x >> 5];
  uint32_t *spos = s_pos[threadIdx.x >> 5];
  const uint64_t gw = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint64_t nw = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  unsigned long long sum = 0;
  uint32_t mx = 0;
  // Software pipeline over this warp's buckets: while bucket c is sorted, the positions of bucket c+nw and the
  // bounds of bucket c+2nw are already in flight, so only the text gather of the current bucket is exposed.
  uint32_t nb0 = 0, ncur = 0, n2b0 = 0, n2cur = 0;
  uint32_t npos[4] = {0, 0, 0, 0};
  {
    const uint64_t c0 = mincode + gw;
    if (c0 <= maxcode) {
      nb0 = lb[c0]; ncur = cursor[c0];
      const uint32_t ncnt = ncur - nb0;
      if (ncnt >= 1 && ncnt <= SORT_MAXBUCKET) {
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (e * 32 + lane < ncnt) npos[e] = out.sa[nb0 - out.partoffset + e * 32 + lane];
      }
    }
    if (c0 + nw <= maxcode) { n2b0 = lb[c0 + nw]; n2cur = cursor[c0 + nw]; }
  }
  for (uint64_t c = mincode + gw; c <= maxcode; c += nw) {
    const uint32_t b0 = nb0;
    const uint32_t cnt = ncur - b0;
    uint32_t pos[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) pos[e] = npos[e];
    nb0 = n2b0; ncur = n2cur;                      // bucket c+nw: bounds arrived, request its positions
    if (c + nw <= maxcode) {
      const uint32_t ncnt = ncur - nb0;
      if (ncnt >= 1 && ncnt <= SORT_MAXBUCKET) {
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (e * 32 + lane < ncnt) npos[e] = out.sa[nb0 - out.partoffset + e * 32 + lane];
      }
    }
    if (c + 2 * nw <= maxcode) { n2b0 = lb[c + 2 * nw]; n2cur = cursor[c + 2 * nw]; }
    const uint32_t start = b0 - out.partoffset;
    if (cnt <= 1) {                          // a single suffix is in place; only the 64-bit image is missing
      if (cnt == 1 && out.sa64 && lane == 0) {
        out.sa64[start] = pos[0];
        if (pos[0] == 0) out.stats->longest = (uint64_t)out.partoffset + start;
      }
      continue;
    }
    bool ok = false;
    if (cnt > maxbucket) {
      if (cnt <= medmax) {                 // too large for one warp: queue for k_sort_medium
        if (lane == 0) {
          medlist[atomicAdd(&out.stats->med_count, 1u)] = (uint32_t)c;
          atomicAdd(&out.stats->med_elems, (unsigned long long)cnt);
        }
        continue;
      }
      ok = false;
    } else if (cnt <= 32) ok = warp_sort_any<BITS, 1>(T, out, out.sa + start, pos, nullptr, nullptr, start, cnt, pl, lane, skey, spos, sum, mx);
    else if (cnt <= 64) ok = warp_sort_any<BITS, 2>(T, out, out.sa + start, pos, nullptr, nullptr, start, cnt, pl, lane, skey, spos, sum, mx);
    else if (cnt <= 96) ok = warp_sort_any<BITS, 3>(T, out, out.sa + start, pos, nullptr, nullptr, start, cnt, pl, lane, skey, spos, sum, mx);
    else if (cnt <= 128) ok = warp_sort_any<BITS, 4>(T, out, out.sa + start, pos, nullptr, nullptr, start, cnt, pl, lane, skey, spos, sum, mx);
    if (!ok && lane == 0) {
      const uint64_t rel = c - mincode;
      // cnt <= maxbucket: sorted, only its tie groups are refined; otherwise the whole bucket is
      uint32_t *bits = cnt <= maxbucket ? fbbits + tieoff : fbbits;
      atomicOr(&bits[rel >> 5], 1u << (rel & 31));
      if (cnt > maxbucket) atomicAdd(&out.stats->fb_unsorted, 1u);
      atomicAdd(&out.stats->fb_elems, (unsigned long long)cnt);
      atomicAdd(&out.stats->fb_buckets, 1u);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sum += __shfl_xor_sync(GTB_FULL, sum, o);
    mx = max(mx, __shfl_xor_sync(GTB_FULL, mx, o));
  }
  if (lane == 0) {
    if (sum) atomicAdd(&out.stats->lcpsum, sum);
    if (mx) atomicMax(&out.stats->maxdepth, mx);
  }
}

#include "tiles.cuh"
#include "keyed.cuh"

// Buckets of 129..MED_MAX suffixes: one block per bucket.  Full 64-bit keys and positions are parked in shared
// memory by slot; what is sorted is one 32-bit word per suffix, (20 leading key bits << 12) | slot:
//   phase A  every warp sorts chunks of 32 E words in registers (WarpSorter32, no block barrier),
//   phase B  log2(P / 32 E) merge levels: the wide exchanges (flip, strides >= chunk) run in shared memory
//            with a barrier each, the narrow ones again in registers per chunk,
//   fix-up   runs of equal 20-bit prefixes (short: the prefix covers 10 bases / 4 residues) are put in
//            (key, position) order by the thread that finds the run start; a run longer than 32 sends the
//            bucket to the exact block-wide 64-bit network below.
// LCP and tie detection as in the warp sorter.
#define MED_MAX 1024          // 16 KB of shared memory per block: eight blocks per SM hide the gathers and barriers
#define MED_SMEM ((size_t)MED_MAX * 16)

template <int E>
__device__ __forceinline__ void med_chunks(uint32_t *s32, uint32_t P, int wid, int lane, bool full_sort) {
  for (uint32_t cb = wid * 32 * E; cb < P; cb += 8 * 32 * E) {
    uint32_t k[E];
#pragma unroll
    for (int e = 0; e < E; ++e) k[e] = s32[cb + e * 32 + lane];
    if (full_sort) WarpSorter32<E>::sort(k, lane);
    else WarpSorter32<E>::tail(k, lane);
#pragma unroll
    for (int e = 0; e < E; ++e) s32[cb + e * 32 + lane] = k[e];
  }
}

// exact fallback: block-wide bitonic network on the (key, position) pairs themselves
__device__ __forceinline__ void med_bitonic64(uint64_t *skey, uint32_t *spos, uint32_t P) {
  for (uint32_t size = 2; size <= P; size <<= 1) {
    for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
      for (uint32_t i = threadIdx.x; i < P / 2; i += 256) {
        const uint32_t lo = 2 * i - (i & (stride - 1)), hi = lo + stride;
        const bool asc = ((lo & size) == 0);
        const uint64_t ka = skey[lo], kb = skey[hi];
        const uint32_t pa = spos[lo], pb = spos[hi];
        const bool gt = ka > kb || (ka == kb && pa > pb);
        if (gt == asc) { skey[lo] = kb; skey[hi] = ka; spos[lo] = pb; spos[hi] = pa; }
      }
      __syncthreads();
    }
  }
}

template <int BITS>
__global__ void __launch_bounds__(256) k_sort_medium(Text<BITS> T, PartOut out, const uint32_t *__restrict__ lb,
                                                     const uint32_t *__restrict__ cursor,
                                                     const uint32_t *__restrict__ medlist, uint64_t mincode,
                                                     unsigned pl, uint32_t *__restrict__ fbbits, uint64_t tieoff,
                                                     const uint64_t *__restrict__ keys_in) {
  extern __shared__ __align__(16) unsigned char sm_raw[];
  uint64_t *skey = (uint64_t *)sm_raw;                 // full keys by slot
  uint32_t *spos = (uint32_t *)(skey + MED_MAX);       // positions by slot
  uint32_t *s32 = spos + MED_MAX;                      // (prefix20 << 12) | slot, sorted in place
  __shared__ int s_tie, s_bigrun;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const uint32_t nmed = out.stats->med_count;
  unsigned long long sum = 0;
  uint32_t mx = 0;
  for (uint32_t b = blockIdx.x; b < nmed; b += gridDim.x) {
    const uint32_t c = medlist[b];
    const uint32_t b0 = lb[c], cnt = cursor[c] - b0, start = b0 - out.partoffset;
    uint32_t P = 256;
    while (P < cnt) P <<= 1;
    for (uint32_t i = threadIdx.x; i < P; i += 256) {
      if (i < cnt) {
        const uint32_t p = out.sa[start + i];
        // keys_in: the keyed tile path left the key beside the position (no text access)
        const uint64_t key = keys_in ? __ldcs(keys_in + start + i) : make_key<BITS>(T, (uint64_t)p + pl);
        skey[i] = key;
        spos[i] = p;
        s32[i] = (key_prefix20<BITS>(key) << 12) | i;
      } else {
        s32[i] = 0xffffffffu;
      }
    }
    if (threadIdx.x == 0) { s_tie = 0; s_bigrun = 0; }
    __syncthreads();
    // ---- sort the 32-bit words ----
    const int Ew = (int)(P >> 8);                      // registers per lane of a warp chunk: 1, 2 or 4
    const uint32_t C = 32u * Ew;
    if (Ew == 1) med_chunks<1>(s32, P, wid, lane, true);
    else if (Ew == 2) med_chunks<2>(s32, P, wid, lane, true);
    else med_chunks<4>(s32, P, wid, lane, true);
    __syncthreads();
    for (uint32_t size = 2 * C; size <= P; size <<= 1) {
      const uint32_t half = size >> 1;
      for (uint32_t i = threadIdx.x; i < P / 2; i += 256) {        // flip: index ^ (size-1)
        const uint32_t lo = (i / half) * size + (i & (half - 1)), hi = lo ^ (size - 1);
        const uint32_t a = s32[lo], bb = s32[hi];
        s32[lo] = min(a, bb);
        s32[hi] = max(a, bb);
      }
      __syncthreads();
      for (uint32_t stride = size >> 2; stride >= C; stride >>= 1) {
        for (uint32_t i = threadIdx.x; i < P / 2; i += 256) {
          const uint32_t lo = 2 * i - (i & (stride - 1)), hi = lo + stride;
          const uint32_t a = s32[lo], bb = s32[hi];
          s32[lo] = min(a, bb);
          s32[hi] = max(a, bb);
        }
        __syncthreads();
      }
      if (Ew == 1) med_chunks<1>(s32, P, wid, lane, false);
      else if (Ew == 2) med_chunks<2>(s32, P, wid, lane, false);
      else med_chunks<4>(s32, P, wid, lane, false);
      __syncthreads();
    }
    // ---- fix-up: order inside runs of equal 20-bit prefixes by (key, position) ----
    for (uint32_t r = threadIdx.x; r + 1 < cnt; r += 256) {
      const uint32_t me = s32[r] >> 12;
      if ((r == 0 || (s32[r - 1] >> 12) != me) && (s32[r + 1] >> 12) == me) {
        uint32_t m = 2;
        while (r + m < cnt && (s32[r + m] >> 12) == me && m <= 32) ++m;
        if (m > 32) {
          s_bigrun = 1;
        } else {
          for (uint32_t a = 1; a < m; ++a) {                         // insertion sort, m is 2 or 3 almost always
            const uint32_t w = s32[r + a];
            const uint64_t kw = skey[w & 0xfff];
            const uint32_t pw = spos[w & 0xfff];
            uint32_t j = a;
            while (j > 0) {
              const uint32_t v = s32[r + j - 1];
              const uint64_t kv = skey[v & 0xfff];
              if (kv < kw || (kv == kw && spos[v & 0xfff] < pw)) break;
              s32[r + j] = v;
              --j;
            }
            s32[r + j] = w;
          }
        }
      }
    }
    __syncthreads();
    if (s_bigrun) {                                     // block-uniform: exact network on the pairs, identity slots
      for (uint32_t i = threadIdx.x + cnt; i < P; i += 256) { skey[i] = ~0ull; spos[i] = 0xffffffffu; }
      __syncthreads();
      med_bitonic64(skey, spos, P);
      for (uint32_t i = threadIdx.x; i < cnt; i += 256) s32[i] = i;
      __syncthreads();
    }
    for (uint32_t i = threadIdx.x + 1; i < cnt; i += 256) {
      const uint64_t ka = skey[s32[i - 1] & 0xfff], kb = skey[s32[i] & 0xfff];
      if (ka == kb && (kb & 0xff) == 0) s_tie = 1;
    }
    __syncthreads();
    if (s_tie && threadIdx.x == 0) {                    // sorted on pl+W symbols; tie groups are refined
      const uint64_t rel = c - mincode;
      atomicOr(&fbbits[tieoff + (rel >> 5)], 1u << (rel & 31));
      atomicAdd(&out.stats->fb_elems, (unsigned long long)cnt);
      atomicAdd(&out.stats->fb_buckets, 1u);
    }
    for (uint32_t i = threadIdx.x; i < cnt; i += 256) {
      const uint32_t slot = s32[i] & 0xfff;
      put_sa(out, start + i, spos[slot]);
      if (i >= 1) {
        const uint64_t ka = skey[s32[i - 1] & 0xfff], kb = skey[slot];
        const uint32_t l = pl + key_common<BITS>(ka, kb);
        out.lcp[start + i] = (uint8_t)l;
        if (!(ka == kb && (kb & 0xff) == 0)) {
          sum += l;
          mx = max(mx, l);
        }
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sum += __shfl_xor_sync(GTB_FULL, sum, o);
    mx = max(mx, __shfl_xor_sync(GTB_FULL, mx, o));
  }
  if ((threadIdx.x & 31) == 0) {
    if (sum) atomicAdd(&out.stats->lcpsum, sum);
    if (mx) atomicMax(&out.stats->maxdepth, mx);
  }
}

// ---------------------------------------------------------------------------------------------
// refinement path (buckets > MED_MAX or still tied after pl+W symbols)
// ---------------------------------------------------------------------------------------------
__global__ void k_fb_popc(const uint32_t *__restrict__ fbbits, const uint32_t *__restrict__ tiebits, uint64_t nwords,
                          uint32_t *cnt) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nwords) cnt[t] = __popc(fbbits[t] | tiebits[t]);
}

// flagged buckets in code order; kind 0 = never sorted (whole bucket, depth pl), 1 = sorted, has tie groups
__global__ void k_fb_codes(const uint32_t *__restrict__ fbbits, const uint32_t *__restrict__ tiebits, uint64_t nwords,
                           const uint32_t *__restrict__ off, uint64_t mincode, uint32_t *codes, uint32_t *kind) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nwords) return;
  const uint32_t fb = fbbits[t];
  uint32_t bits = fb | tiebits[t];
  uint32_t o = off[t];
  while (bits) {
    const int b = __ffs(bits) - 1;
    bits &= bits - 1;
    codes[o] = (uint32_t)(mincode + (t << 5) + b);
    kind[o] = (fb >> b) & 1u ? 0u : 1u;
    o++;
  }
}

__device__ __forceinline__ uint32_t regular_count(const uint32_t *__restrict__ lb, const uint32_t *__restrict__ csc,
                                                  uint32_t c, unsigned sigma) {
  const uint32_t npad = (c % sigma == sigma - 1) ? csc[(c - (sigma - 1)) / sigma] : 0u;
  return lb[c + 1] - lb[c] - npad;
}

// One warp per flagged bucket.  FILL = false: sizes[b] = number of list entries the bucket contributes
// (all regular suffixes of a never-sorted bucket; the members of tie groups of a sorted one).
// FILL = true: writes the entries (ascending suffix-array index) at eoff[b].
template <int BITS, bool FILL>
__global__ void k_fb_list(Text<BITS> T, const uint32_t *__restrict__ sa, const uint32_t *__restrict__ codes,
                          const uint32_t *__restrict__ kind, uint32_t nb, const uint32_t *__restrict__ lb,
                          const uint32_t *__restrict__ csc, unsigned sigma, uint32_t partoffset, unsigned pl,
                          uint32_t *sizes, const uint32_t *__restrict__ eoff, uint32_t *idx, uint32_t *gs, uint32_t *dep) {
  constexpr int W = Text<BITS>::W;
  const uint64_t gw = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint64_t nw = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  const int lane = threadIdx.x & 31;
  for (uint64_t b = gw; b < nb; b += nw) {
    const uint32_t c = codes[b];
    const uint32_t start = lb[c] - partoffset, cnt = regular_count(lb, csc, c, sigma);
    if (kind[b] == 0) {
      if (!FILL) {
        if (lane == 0) sizes[b] = cnt;
      } else {
        const uint32_t o = eoff[b];
        for (uint32_t i = lane; i < cnt; i += 32) { idx[o + i] = start + i; gs[o + i] = start; dep[o + i] = pl; }
      }
      continue;
    }
    uint32_t total = 0, carry_head = 0;        // carry_head: run start inherited from the previous 32-chunk
    for (uint32_t base = 0; base < cnt; base += 32) {
      const uint32_t i = base + lane;
      uint64_t kp = 1, kc = 0, kn = 1;
      if (i < cnt) {
        kc = make_key<BITS>(T, (uint64_t)sa[start + i] + pl);
        if (i >= 1) kp = make_key<BITS>(T, (uint64_t)sa[start + i - 1] + pl);
        if (i + 1 < cnt) kn = make_key<BITS>(T, (uint64_t)sa[start + i + 1] + pl);
      }
      const bool open_key = i < cnt && (kc & 0xff) == 0;
      const bool tp = open_key && i >= 1 && kp == kc;          // tied with predecessor
      const bool tn = open_key && i + 1 < cnt && kn == kc;     // tied with successor
      const bool member = tp || tn;
      const unsigned mm = __ballot_sync(GTB_FULL, member);
      if (FILL) {
        // run start = index of the latest member that is not tied with its predecessor
        const bool hd = member && !tp;
        const unsigned hm = __ballot_sync(GTB_FULL, hd);
        const unsigned below = hm & ((2u << lane) - 1u);       // heads at lanes <= mine
        const uint32_t rs = below ? base + (31 - __clz(below)) : carry_head;
        if (member) {
          const uint32_t o = eoff[b] + total + __popc(mm & ((1u << lane) - 1u));
          idx[o] = start + i;
          gs[o] = start + rs;
          dep[o] = pl + W;
        }
        if (hm) carry_head = base + (31 - __clz(hm));
      }
      total += __popc(mm);
    }
    if (!FILL && lane == 0) sizes[b] = total;
  }
}

template <int BITS>
__global__ void k_fb_keys(Text<BITS> T, const uint32_t *__restrict__ sa, const uint32_t *__restrict__ idx,
                          const uint32_t *__restrict__ dep, uint32_t A, uint64_t *keys, uint32_t *posbuf,
                          uint32_t *vals) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= A) return;
  const uint32_t p = sa[idx[j]];
  const uint32_t d = dep[j];
  posbuf[j] = p;
  keys[j] = (d == GTB_POSMODE) ? (uint64_t)p : make_key<BITS>(T, (uint64_t)p + d);
  vals[j] = j;
}

__global__ void k_fb_gather_gs(const uint32_t *__restrict__ vals, const uint32_t *__restrict__ gs, uint32_t A,
                               uint64_t *keys2) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < A) keys2[j] = gs[vals[j]];
}

// after the segmented sort: write the new order, emit LCPs of newly separated neighbours, flag heads
template <int BITS>
__global__ void k_fb_regroup(PartOut out, const uint32_t *__restrict__ vals, const uint32_t *__restrict__ idx,
                             const uint32_t *__restrict__ gs, const uint32_t *__restrict__ dep,
                             const uint64_t *__restrict__ keys, const uint32_t *__restrict__ posbuf, uint32_t A,
                             uint32_t *head, uint32_t *newdep) {
  constexpr int W = Text<BITS>::W;
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long sum = 0;
  uint32_t mx = 0;
  if (j < A) {
    const uint32_t e = vals[j];
    const uint32_t g = gs[e], d = dep[e];
    const uint64_t kc = keys[e];
    const uint32_t r = idx[j];
    out.sa[r] = posbuf[e];
    bool hd = true;
    if (j > 0) {
      const uint32_t e0 = vals[j - 1];
      const uint64_t k0 = keys[e0];
      if (gs[e0] == g) {
        if (d == GTB_POSMODE) {
          hd = true;                        // positions are unique; lcp was emitted when the tie was found
        } else if (k0 != kc) {
          const uint32_t l = d + key_common<BITS>(k0, kc);
          out.lcp[r] = (uint8_t)(l < 255 ? l : 255);
          if (l >= 255) emit_large(out, r, l);
          sum = l; mx = l;
        } else if ((kc & 0xff) != 0) {      // same real symbols, both cut by a special at the same depth
          const uint32_t l = d + (W - (uint32_t)(kc & 0xff));
          out.lcp[r] = (uint8_t)(l < 255 ? l : 255);
          if (l >= 255) emit_large(out, r, l);
          sum = l; mx = l;
          hd = false;                       // stays one group, to be ordered by position
        } else {
          hd = false;                       // still tied on real symbols
        }
      }
    }
    head[j] = hd ? 1u : 0u;
    newdep[j] = (d == GTB_POSMODE) ? GTB_POSMODE : ((kc & 0xff) ? GTB_POSMODE : d + W);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sum += __shfl_xor_sync(GTB_FULL, sum, o);
    mx = max(mx, __shfl_xor_sync(GTB_FULL, mx, o));
  }
  if ((threadIdx.x & 31) == 0) {
    if (sum) atomicAdd(&out.stats->lcpsum, sum);
    if (mx) atomicMax(&out.stats->maxdepth, mx);
  }
}

// groupno = exclusive scan of head; head index table
__global__ void k_fb_headpos(const uint32_t *__restrict__ head, const uint32_t *__restrict__ hscan, uint32_t A,
                             uint32_t *headpos, const unsigned long long *__restrict__ ngroups) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < A && head[j]) headpos[hscan[j]] = j;
  if (j == 0) headpos[*ngroups] = A;
}

__global__ void k_fb_finalize(const uint32_t *__restrict__ head, const uint32_t *__restrict__ hscan,
                              const uint32_t *__restrict__ headpos, const uint32_t *__restrict__ idx, uint32_t A,
                              uint32_t *keep, uint32_t *newgs) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= A) return;
  const uint32_t g = hscan[j] + head[j] - 1;
  const uint32_t hs = headpos[g], he = headpos[g + 1];
  keep[j] = (he - hs) > 1 ? 1u : 0u;
  newgs[j] = idx[hs];
}

__global__ void k_fb_compact(const uint32_t *__restrict__ keep, const uint32_t *__restrict__ kscan,
                             const uint32_t *__restrict__ idx, const uint32_t *__restrict__ newgs,
                             const uint32_t *__restrict__ newdep, uint32_t A, uint32_t *idx2, uint32_t *gs2,
                             uint32_t *dep2, PartStats *stats) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= A || !keep[j]) return;
  const uint32_t o = kscan[j];
  idx2[o] = idx[j];
  gs2[o] = newgs[j];
  dep2[o] = newdep[j];
  if (stats && newdep[j] == GTB_POSMODE) atomicAdd(&stats->posmode_kept, 1u);
}

__global__ void k_reset_posmode(PartStats *stats) { stats->posmode_kept = 0; }

// Skip-ahead for text-window rounds (parts that do not cover the whole suffix array cannot use rank doubling, the
// ranks of other parts' suffixes are unknown): a group tied on d symbols whose members go on agreeing for another
// e windows of W real symbols is moved to depth d + e*W at once, so the next round meets the window in which the
// group splits (or a special cuts it).  A repeat of length R then costs R/(maxwin*W) rounds instead of R/W.
// The members of a group occupy consecutive suffix-array slots, so sa[gs[j]] is the group's first member: every
// member counts the windows it shares with that one, the minimum over the group goes to the slot of the first
// member's list entry (list index j - (idx[j] - gs[j])).  Equal keys with tag 0 = W equal real symbols.
template <int BITS>
__global__ void k_fb_extend_count(Text<BITS> T, const uint32_t *__restrict__ sa, const uint32_t *__restrict__ idx,
                                  const uint32_t *__restrict__ gs, const uint32_t *__restrict__ dep, uint32_t A,
                                  uint32_t maxwin, uint32_t *extmin) {
  constexpr int W = Text<BITS>::W;
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= A) return;
  const uint32_t d = dep[j], r = idx[j], g = gs[j];
  if (d == GTB_POSMODE || r == g) return;
  const uint64_t p = (uint64_t)sa[r] + d, q = (uint64_t)sa[g] + d;
  uint32_t e = 0;
  while (e < maxwin) {
    const uint64_t ka = make_key<BITS>(T, p + (uint64_t)e * W);
    if ((ka & 0xff) != 0 || ka != make_key<BITS>(T, q + (uint64_t)e * W)) break;
    e++;
  }
  atomicMin(&extmin[j - (r - g)], e);
}

__global__ void k_fb_extend_apply(const uint32_t *__restrict__ idx, const uint32_t *__restrict__ gs, uint32_t A,
                                  unsigned W, const uint32_t *__restrict__ extmin, uint32_t *dep) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= A || dep[j] == GTB_POSMODE) return;
  dep[j] += extmin[j - (idx[j] - gs[j])] * W;
}

// Pivot rounds (GTB200_EXTEND=2): the skip-ahead above moves a group by what ALL members share, which is nothing for
// a tandem array or a homopolymer run (some member always ends in the next window).  Here every group picks one
// member as pivot (pseudo-random: largest hash of the position, k_fb_pivot_pick) and every member j walks along the
// pivot until the first window e_j in which it differs from it, is cut by a special, or the cap is reached; k_j is
// its key in that window, kp_j the pivot's.  Order inside the group: members below the pivot (k_j < kp_j) by
// (e_j, k_j) ascending - one that leaves the pivot earlier towards smaller keys is smaller than one that still
// agrees with it -, then the members equal to the pivot in its last window (still tied, or cut at the same place),
// then the members above it by e_j DESCENDING, k_j ascending.  The class number c_j encodes exactly that order, so
// one stable sort on (group, c, k) places every member; neighbours of different classes have
// lcp = d + e_s*W + common(k_s, kp_s), s = the one that left the pivot first.  A group of m members whose ends
// are spread out halves in expectation every round.
#define GTB_PIVOT_MAXWIN 4096u      // c < 2^14 must hold: c <= 2 * GTB_PIVOT_MAXWIN

__global__ void k_fb_pivot_pick(const uint32_t *__restrict__ sa, const uint32_t *__restrict__ idx,
                                const uint32_t *__restrict__ gs, const uint32_t *__restrict__ dep, uint32_t A, uint32_t salt,
                                unsigned long long *pvt) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= A || dep[j] == GTB_POSMODE) return;
  const uint32_t r = idx[j], g = gs[j];
  uint32_t h = (sa[r] + salt) * 0x9E3779B1u;
  h ^= h >> 15; h *= 0x85EBCA77u; h ^= h >> 13;
  atomicMax(&pvt[j - (r - g)], ((unsigned long long)h << 32) | (unsigned long long)(r - g));
}

template <int BITS>
__global__ void k_fb_keys_pivot(Text<BITS> T, const uint32_t *__restrict__ sa, const uint32_t *__restrict__ idx,
                                const uint32_t *__restrict__ gs, const uint32_t *__restrict__ dep, uint32_t A,
                                const unsigned long long *__restrict__ pvt, uint64_t *keys, uint64_t *kp, uint32_t *ebuf,
                                uint32_t *cbuf, uint32_t *posbuf, uint32_t *vals) {
  constexpr int W = Text<BITS>::W;
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= A) return;
  const uint32_t r = idx[j], g = gs[j], d = dep[j];
  const uint32_t p = sa[r];
  posbuf[j] = p;
  vals[j] = j;
  if (d == GTB_POSMODE) { keys[j] = p; kp[j] = 0; ebuf[j] = 0; cbuf[j] = 0; return; }
  const uint32_t q = sa[g + (uint32_t)pvt[j - (r - g)]];
  uint32_t e = 0;
  uint64_t ka, kb;
  for (;;) {
    ka = make_key<BITS>(T, (uint64_t)p + d + (uint64_t)e * W);
    kb = make_key<BITS>(T, (uint64_t)q + d + (uint64_t)e * W);
    if (ka != kb || (ka & 0xff) != 0 || e == GTB_PIVOT_MAXWIN - 1) break;
    e++;
  }
  keys[j] = ka;
  kp[j] = kb;
  ebuf[j] = e;
  cbuf[j] = ka < kb ? e : (ka == kb ? GTB_PIVOT_MAXWIN : 2 * GTB_PIVOT_MAXWIN - e);
}

__global__ void k_fb_gather_gsc(const uint32_t *__restrict__ vals, const uint32_t *__restrict__ gs,
                                const uint32_t *__restrict__ cbuf, uint32_t A, uint64_t *keys2) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < A) keys2[j] = ((uint64_t)gs[vals[j]] << 16) | cbuf[vals[j]];
}

template <int BITS>
__global__ void k_fb_regroup_pivot(PartOut out, const uint32_t *__restrict__ vals, const uint32_t *__restrict__ idx,
                                   const uint32_t *__restrict__ gs, const uint32_t *__restrict__ dep,
                                   const uint64_t *__restrict__ keys, const uint64_t *__restrict__ kp,
                                   const uint32_t *__restrict__ ebuf, const uint32_t *__restrict__ cbuf,
                                   const uint32_t *__restrict__ posbuf, uint32_t A, uint32_t *head, uint32_t *newdep) {
  constexpr int W = Text<BITS>::W;
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long sum = 0;
  uint32_t mx = 0;
  if (j < A) {
    const uint32_t e = vals[j];
    const uint32_t g = gs[e], d = dep[e];
    const uint64_t kc = keys[e];
    const uint32_t r = idx[j];
    out.sa[r] = posbuf[e];
    bool hd = true;
    if (j > 0 && d != GTB_POSMODE) {
      const uint32_t e0 = vals[j - 1];
      if (gs[e0] == g) {
        const uint32_t c0 = cbuf[e0], c1 = cbuf[e];
        uint32_t l = 0;
        bool emit = true;
        if (c0 == c1) {                       // same class: both left the pivot in the same window (or never did)
          const uint32_t D = d + ebuf[e] * W;
          const uint64_t k0 = keys[e0];
          if (k0 != kc) l = D + key_common<BITS>(k0, kc);
          else if ((kc & 0xff) != 0) { l = D + (W - (uint32_t)(kc & 0xff)); hd = false; }   // cut at the same place: by position
          else { hd = false; emit = false; }  // still tied on real symbols
        } else {
          // the one that left the pivot first (the other still agreed with the pivot in that window); with equal
          // windows one of the two is a member of the pivot's own class and the other one left
          const uint32_t w0 = ebuf[e0], w1 = ebuf[e];
          const uint32_t s = w0 < w1 ? e0 : (w1 < w0 ? e : (c0 == GTB_PIVOT_MAXWIN ? e : e0));
          l = d + ebuf[s] * W + key_common<BITS>(keys[s], kp[s]);
        }
        if (emit) {
          out.lcp[r] = (uint8_t)(l < 255 ? l : 255);
          if (l >= 255) emit_large(out, r, l);
          sum = l; mx = l;
        }
      }
    }
    head[j] = hd ? 1u : 0u;
    newdep[j] = (d == GTB_POSMODE) ? GTB_POSMODE : ((kc & 0xff) ? GTB_POSMODE : d + (ebuf[e] + 1) * W);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sum += __shfl_xor_sync(GTB_FULL, sum, o);
    mx = max(mx, __shfl_xor_sync(GTB_FULL, mx, o));
  }
  if ((threadIdx.x & 31) == 0) {
    if (sum) atomicAdd(&out.stats->lcpsum, sum);
    if (mx) atomicMax(&out.stats->maxdepth, mx);
  }
}

// ---------------------------------------------------------------------------------------------
// Prefix doubling (Larsson-Sadakane style) for parts that cover the whole suffix array: the rank
// array ISA (group start index of every suffix; final index for resolved ones, tail suffixes
// included) lets a group tied on h symbols be split by the rank of the suffix h positions further
// on, so a repeat of length L costs log2(L/h) rounds instead of L/28.
// ---------------------------------------------------------------------------------------------
__global__ void k_isa_from_sa(const uint32_t *__restrict__ sa, uint64_t width, uint32_t *__restrict__ isa) {
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < width) isa[sa[r]] = (uint32_t)r;
}

// suffixes that start with a special symbol (and the empty suffix n): rank = their place in the tail
__global__ void k_isa_tail(const gtb200_range *__restrict__ ranges, const uint32_t *__restrict__ rangeoff,
                           uint64_t nranges, uint64_t specialchars, uint64_t n, uint64_t nsort, uint32_t *isa) {
  const uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e > specialchars) return;
  if (e == specialchars) { isa[n] = (uint32_t)(nsort + e); return; }
  uint64_t lo = 0, hi = nranges;
  while (hi - lo > 1) {
    const uint64_t mid = (lo + hi) >> 1;
    if (rangeoff[mid] <= e) lo = mid; else hi = mid;
  }
  isa[ranges[lo].start + (e - rangeoff[lo])] = (uint32_t)(nsort + e);
}

__global__ void k_isa_groups(const uint32_t *__restrict__ idx, const uint32_t *__restrict__ gs,
                             const uint32_t *__restrict__ sa, uint32_t A, uint32_t *isa) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < A) isa[sa[idx[j]]] = gs[j];
}

__global__ void k_dbl_keys(const uint32_t *__restrict__ sa, const uint32_t *__restrict__ idx,
                           const uint32_t *__restrict__ gs, uint32_t A, uint64_t h, const uint32_t *__restrict__ isa,
                           uint64_t *keys, uint32_t *posbuf, uint32_t *vals) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= A) return;
  const uint32_t p = sa[idx[j]];
  posbuf[j] = p;
  keys[j] = ((uint64_t)gs[j] << 32) | (uint64_t)isa[(uint64_t)p + h];
  vals[j] = j;
}

// sorted by (group, rank of the suffix h further on): new order, heads of the new groups
__global__ void k_dbl_regroup(uint32_t *__restrict__ sa, const uint64_t *__restrict__ skeys,
                              const uint32_t *__restrict__ svals, const uint32_t *__restrict__ idx,
                              const uint32_t *__restrict__ posbuf, uint32_t A, uint32_t *head) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= A) return;
  sa[idx[j]] = posbuf[svals[j]];
  head[j] = (j == 0 || skeys[j] != skeys[j - 1]) ? 1u : 0u;
}

__global__ void k_dbl_update_isa(const uint32_t *__restrict__ sa, const uint32_t *__restrict__ idx,
                                 const uint32_t *__restrict__ newgs, uint32_t A, uint32_t *isa) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < A) isa[sa[idx[j]]] = newgs[j];
}

// entries whose lcp the fast path could not determine: everything in the initial list but group heads
__global__ void k_phi_flags(const uint32_t *__restrict__ idx0, const uint32_t *__restrict__ gs0, uint32_t A,
                            uint32_t *flag) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < A) flag[j] = idx0[j] != gs0[j] ? 1u : 0u;
}

__global__ void k_phi_build(const uint32_t *__restrict__ idx0, const uint32_t *__restrict__ flag,
                            const uint32_t *__restrict__ fscan, const uint32_t *__restrict__ sa, uint32_t A,
                            uint64_t *keys, uint32_t *vals) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= A || !flag[j]) return;
  const uint32_t o = fscan[j], r = idx0[j];
  keys[o] = sa[r];          // text position: the Kasai order
  vals[o] = r;
}

// Phi / Kasai over the open entries in text order: lcp(i, Phi(i)) >= lcp(i-1, Phi(i-1)) - 1, so a thread
// that walks consecutive text positions re-compares only what is new.  One thread per PHI_CHUNK entries;
// a chunk start (or a gap in the positions) restarts from the prefix length shared inside a bucket.
#define PHI_CHUNK 32
template <int BITS>
__global__ void k_phi_lcp(Text<BITS> T, PartOut out, const uint64_t *__restrict__ skeys,
                          const uint32_t *__restrict__ svals, uint32_t U, unsigned pl) {
  constexpr int SPW = Text<BITS>::SPW;
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long sum = 0;
  uint32_t mx = 0;
  const uint64_t e0 = (uint64_t)t * PHI_CHUNK;
  uint64_t l = 0, previ = ~0ull - 1;
  for (uint64_t e = e0; e < e0 + PHI_CHUNK && e < U; ++e) {
    const uint64_t i = skeys[e];
    const uint32_t r = svals[e];
    const uint64_t phi = out.sa[r - 1];
    l = (i == previ + 1 && l > 0) ? l - 1 : 0;
    if (l < pl) l = pl;                       // neighbours inside a bucket share the prefix
    for (;;) {
      const uint64_t x = T.window(i + l) ^ T.window(phi + l);
      unsigned step = x ? (unsigned)__clzll((long long)x) / BITS : (unsigned)SPW;
      const unsigned ra = T.reallen(i + l), rb = T.reallen(phi + l);
      step = step < ra ? step : ra;
      step = step < rb ? step : rb;
      l += step;
      if (step < (unsigned)SPW) break;
    }
    out.lcp[r] = (uint8_t)(l < 255 ? l : 255);
    if (l >= 255) emit_large(out, r, (uint32_t)l);
    sum += l;
    mx = max(mx, (uint32_t)l);
    previ = i;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sum += __shfl_xor_sync(GTB_FULL, sum, o);
    mx = max(mx, __shfl_xor_sync(GTB_FULL, mx, o));
  }
  if ((threadIdx.x & 31) == 0) {
    if (sum) atomicAdd(&out.stats->lcpsum, sum);
    if (mx) atomicMax(&out.stats->maxdepth, mx);
  }
}

__global__ void k_llv_keys(const gtb200_largelcp *__restrict__ llv, uint32_t nl, uint64_t *keys, uint32_t *vals) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < nl) { keys[j] = llv[j].position; vals[j] = j; }
}

__global__ void k_llv_gather(const gtb200_largelcp *__restrict__ src, const uint32_t *__restrict__ vals, uint32_t nl,
                             gtb200_largelcp *dst) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < nl) dst[j] = src[vals[j]];
}

// ---------------------------------------------------------------------------------------------
// bucket boundaries, widening, tail
// ---------------------------------------------------------------------------------------------

// lcp of the first entry of every non-empty bucket with its left neighbour in the suffix array
// (src/match/sfx-lcpvalues.c:91-111,591-693 compute the same value from codes and prefix indices)
template <int BITS>
__global__ void k_boundary(Text<BITS> T, CodeParams cp, PartOut out, const uint32_t *__restrict__ lb,
                           const uint32_t *__restrict__ csc, uint64_t mincode, uint64_t maxcode) {
  const uint64_t c = mincode + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned pl = cp.pl;
  unsigned long long sum = 0;
  uint32_t mx = 0;
  if (c <= maxcode) {
    const uint32_t b0 = lb[c], b1 = lb[c + 1];
    const uint32_t r = b0 - out.partoffset;
    if (b1 != b0 && r != 0) {                 // non-empty and not the first entry of the part (k_part_first)
      uint64_t cprev = c - 1;                 // previous non-empty bucket; exists inside the part because r > 0
      while (lb[cprev] == b0) --cprev;
      const uint32_t padc = (c % cp.sigma == cp.sigma - 1) ? csc[(c - (cp.sigma - 1)) / cp.sigma] : 0u;
      const uint32_t padp = (cprev % cp.sigma == cp.sigma - 1) ? csc[(cprev - (cp.sigma - 1)) / cp.sigma] : 0u;
      unsigned l;
      bool regular;
      if (padp == 0 && b1 - b0 > padc) {
        // both neighbours have >= pl real symbols: the value is the common prefix of the two codes
        regular = true;
        if (cp.pow2) {
          const uint64_t x = cprev ^ c;
          l = pl - 1 - (unsigned)(63 - __clzll((long long)x)) / BITS;
        } else {
          l = 0;
          for (unsigned i = 0; i < pl; ++i) {
            if ((cprev / cp.pw[pl - 1 - i]) % cp.sigma != (c / cp.pw[pl - 1 - i]) % cp.sigma) break;
            ++l;
          }
        }
      } else {
        // a special-padded suffix is involved: compare the two suffixes themselves
        const uint32_t a = out.sa[r - 1], b = out.sa[r];
        const unsigned la = T.reallen(a), lbn = T.reallen(b);
        const uint64_t x = T.window(a) ^ T.window(b);
        const unsigned common = x ? (unsigned)__clzll((long long)x) / BITS : 64u / BITS;
        l = common < la ? common : la;
        l = l < lbn ? l : lbn;
        if (l > pl) l = pl;
        regular = lbn >= pl;
      }
      out.lcp[r] = (uint8_t)l;
      if (regular) sum = l;                   // special-padded entries do not enter lcptabsum
      mx = l;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sum += __shfl_xor_sync(GTB_FULL, sum, o);
    mx = max(mx, __shfl_xor_sync(GTB_FULL, mx, o));
  }
  if ((threadIdx.x & 31) == 0) {
    if (sum) atomicAdd(&out.stats->bsum, sum);
    if (mx) atomicMax(&out.stats->bmax, mx);
  }
}

// first entry of a part that is not the first part: its left neighbour lives on another device /
// in an earlier part, so the value is derived from codes and the padded list alone
template <int BITS>
__global__ void k_part_first(Text<BITS> T, CodeParams cp, PartOut out, const uint32_t *__restrict__ lb,
                             uint64_t numcodes, const uint32_t *__restrict__ csc,
                             const uint64_t *__restrict__ padkey, const unsigned long long *__restrict__ npadded_dev) {
  const uint64_t npadded = *npadded_dev;
  if (blockIdx.x || threadIdx.x) return;
  const uint32_t po = out.partoffset;
  // cprev = max{x : lb[x] < po}; ccur = max{x : lb[x] <= po, x < numcodes}
  uint64_t lo = 0, hi = numcodes;           // invariant lb[lo] < po
  while (hi - lo > 1) {
    const uint64_t mid = (lo + hi) >> 1;
    if (lb[mid] < po) lo = mid; else hi = mid;
  }
  const uint64_t cprev = lo;
  lo = 0; hi = numcodes;                    // invariant lb[lo] <= po
  while (hi - lo > 1) {
    const uint64_t mid = (lo + hi) >> 1;
    if (lb[mid] <= po) lo = mid; else hi = mid;
  }
  const uint64_t ccur = lo;
  const unsigned pl = cp.pl;
  unsigned la = pl;
  if (cprev % cp.sigma == cp.sigma - 1 && csc[(cprev - (cp.sigma - 1)) / cp.sigma] > 0) {
    // last padded entry of cprev = smallest number of real symbols
    uint64_t l2 = 0, h2 = npadded;          // first index with code > cprev
    const uint64_t target = (cprev + 1) << 6;
    while (l2 < h2) {
      const uint64_t mid = (l2 + h2) >> 1;
      if (padkey[mid] < target) l2 = mid + 1; else h2 = mid;
    }
    la = pl - (unsigned)(padkey[l2 - 1] & 63);
  }
  const uint32_t b = out.sa[0];
  unsigned lbn = T.reallen(b);
  if (lbn > pl) lbn = pl;
  // common leading digits of the two codes
  unsigned common = 0;
  for (unsigned i = 0; i < pl; ++i) {
    const uint64_t da = (cprev / cp.pw[pl - 1 - i]) % cp.sigma, db = (ccur / cp.pw[pl - 1 - i]) % cp.sigma;
    if (da != db) break;
    common++;
  }
  unsigned l = common < la ? common : la;
  l = l < lbn ? l : lbn;
  out.lcp[0] = (uint8_t)l;
  if (lbn >= pl && l) atomicAdd(&out.stats->bsum, (unsigned long long)l);
  if (l) atomicMax(&out.stats->bmax, l);
}

// per-part initialisation in one launch: scatter cursors, refinement flags, statistics
__global__ void k_part_init(const uint32_t *__restrict__ lb, uint32_t *__restrict__ cursor, uint64_t mincode,
                            uint64_t ncodes, uint32_t *__restrict__ fbbits, uint64_t fbwords, PartStats *stats,
                            uint8_t *lcp) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < ncodes) cursor[mincode + i] = lb[mincode + i];
  if (i < fbwords) fbbits[i] = 0;
  if (i == 0) {
    PartStats z;
    memset(&z, 0, sizeof z);
    z.longest = ~0ull;
    *stats = z;
    lcp[0] = 0;
  }
}

__global__ void k_reset_medcount(PartStats *stats) { stats->med_count = 0; }

__global__ void k_reset_boundary_stats(PartStats *stats) {
  stats->bsum = 0;
  stats->bmax = 0;
}

__global__ void k_widen(const uint32_t *__restrict__ sa, uint64_t width, uint64_t *__restrict__ out64,
                        uint64_t partoffset, PartStats *stats) {
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= width) return;
  const uint32_t p = sa[r];
  out64[r] = p;
  if (p == 0) stats->longest = partoffset + r;
}

// After refinement only the flagged buckets (never sorted, or sorted with tie groups) have a new order: when the
// sorters wrote the 64-bit image themselves, just those buckets are re-imaged.  One block per 32 codes.
__global__ void __launch_bounds__(256) k_widen_flagged(const uint32_t *__restrict__ sa, uint64_t *__restrict__ out64,
                                                       const uint32_t *__restrict__ lb, const uint32_t *__restrict__ fbbits,
                                                       const uint32_t *__restrict__ tiebits, uint64_t mincode,
                                                       uint64_t maxcode, uint32_t partoffset, PartStats *stats) {
  uint32_t bits = fbbits[blockIdx.x] | tiebits[blockIdx.x];
  while (bits) {
    const int b = __ffs(bits) - 1;
    bits &= bits - 1;
    const uint64_t code = mincode + ((uint64_t)blockIdx.x << 5) + b;
    if (code > maxcode) break;
    const uint32_t r0 = lb[code] - partoffset, r1 = lb[code + 1] - partoffset;
    for (uint32_t r = r0 + threadIdx.x; r < r1; r += 256) {
      const uint32_t p = sa[r];
      out64[r] = p;
      if (p == 0) stats->longest = (uint64_t)partoffset + r;
    }
  }
}

template <typename OUT>
__global__ void k_tail(const gtb200_range *__restrict__ ranges, const uint32_t *__restrict__ rangeoff,
                       uint64_t nranges, uint64_t specialchars, uint64_t n, uint64_t first, uint64_t count, OUT *out) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= count) return;
  const uint64_t e = first + t;
  if (e >= specialchars) { out[t] = (OUT)n; return; }
  uint64_t lo = 0, hi = nranges;            // rangeoff[lo] <= e < rangeoff[hi]
  while (hi - lo > 1) {
    const uint64_t mid = (lo + hi) >> 1;
    if (rangeoff[mid] <= e) lo = mid; else hi = mid;
  }
  out[t] = (OUT)(ranges[lo].start + (e - rangeoff[lo]));
}

__global__ void k_narrow(const uint64_t *__restrict__ in, uint64_t n, uint32_t *__restrict__ out) {
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < n) out[r] = (uint32_t)in[r];
}

// ---------------------------------------------------------------------------------------------
// host API
// ---------------------------------------------------------------------------------------------
static void free_ctx(gtb200_ctx *c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  if (c->l2_window) cudaCtxResetPersistingL2Cache();
  if (c->copy_stream) cudaStreamSynchronize(c->copy_stream);
  if (c->side_stream) cudaStreamSynchronize(c->side_stream);
  {
    std::lock_guard<std::mutex> lock(g_pool_mu);
    for (auto &b : c->blocks) g_pool[c->device].push_back(b);
    c->blocks.clear();
    if (c->h_stats) g_pinned_pool.push_back(c->h_stats);
  }
  for (auto &p : c->pending) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
  for (auto e : c->evpool) cudaEventDestroy(e);
  if (c->stream) {
    if (c->l2_window) {
      cudaStreamDestroy(c->stream);               // carries an access-policy attribute: not reusable
      if (c->copy_stream) { for (int k = 0; k < 8; ++k) cudaEventDestroy(c->slice_ev[k]); cudaStreamDestroy(c->copy_stream); }
      if (c->side_stream) { for (int k = 0; k < 2; ++k) cudaEventDestroy(c->side_ev[k]); cudaStreamDestroy(c->side_stream); }
    } else {
      StreamSet ss;
      ss.main = c->stream;
      ss.copy = c->copy_stream;
      ss.side = c->side_stream;
      for (int k = 0; k < 8; ++k) ss.ev[k] = c->slice_ev[k];
      for (int k = 0; k < 2; ++k) ss.sev[k] = c->side_ev[k];
      std::lock_guard<std::mutex> lock(g_pool_mu);
      g_stream_pool[c->device].push_back(ss);
    }
  }
  delete c;
}

extern "C" void gtb200_close(gtb200_ctx *c) { free_ctx(c); }

extern "C" void *gtb200_stream(gtb200_ctx *c) { return (void *)c->stream; }

static int sync_stats(gtb200_ctx *c) {
  CK(cudaMemcpyAsync(c->h_stats, c->d_stats, sizeof(PartStats), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

template <int BITS>
static Text<BITS> text_of(gtb200_ctx *c) {
  Text<BITS> T;
  T.w = c->d_text;
  T.m = c->d_mask;
  T.g = c->d_gather;
  return T;
}

static int open_impl(gtb200_ctx *c, const gtb200_seqdesc *seq) {
  const uint64_t n = c->n;
  const int spw = 64 / c->bits;
  const cudaMemcpyKind h2d = seq->memory == GTB200_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
  // sizes: whole 64-position chunks, plus two spare words
  const uint64_t nchunks = (n >> 6) + 1;
  c->nmaskwords = nchunks + 2;
  c->ntextwords = nchunks * (64 / spw) + 2;
  CK(dmalloc(c, &c->d_text, (c->ntextwords + c->nmaskwords) * 8));   // one block: text, then the bitmap
  c->d_mask = c->d_text + c->ntextwords;
  {
    // Keep the packed sequence + bitmap in L2 while suffix-array sized streams pass through: the
    // sort kernels gather 16+16 bytes per suffix from it at random.  Persisting-L2 set-aside and an
    // access-policy window on this context's stream; scaled down if the sequence is larger.
    int maxpersist = 0, maxwin = 0;
    cudaDeviceGetAttribute(&maxpersist, cudaDevAttrMaxPersistingL2CacheSize, c->device);
    cudaDeviceGetAttribute(&maxwin, cudaDevAttrMaxAccessPolicyWindowSize, c->device);
    const size_t bytes = (c->ntextwords + c->nmaskwords) * 8;
    const char *off = getenv("GTB200_L2_PERSIST");
    // Opt-in (GTB200_L2_PERSIST=1): measured on B200 the set-aside speeds the tile sorter up by ~20 % at
    // 250 Mbp but starves every other kernel of L2, a net loss; the default keeps L2 unpartitioned.
    if (maxpersist > 0 && maxwin > 0 && bytes >= (8u << 20) && off && atoi(off) == 1) {
      const size_t setaside = bytes < (size_t)maxpersist ? bytes : (size_t)maxpersist;
      if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, setaside) == cudaSuccess) {
        cudaStreamAttrValue av;
        memset(&av, 0, sizeof av);
        av.accessPolicyWindow.base_ptr = c->d_text;
        av.accessPolicyWindow.num_bytes = bytes < (size_t)maxwin ? bytes : (size_t)maxwin;
        av.accessPolicyWindow.hitRatio = bytes <= setaside ? 1.0f : (float)((double)setaside / (double)bytes);
        av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        if (cudaStreamSetAttribute(c->stream, cudaStreamAttributeAccessPolicyWindow, &av) == cudaSuccess) c->l2_window = true;
      }
      (void)cudaGetLastError();
    }
  }
  CK(dmalloc(c, &c->d_stats, sizeof(PartStats)));
  {
    std::lock_guard<std::mutex> lock(g_pool_mu);
    if (!g_pinned_pool.empty()) { c->h_stats = (PartStats *)g_pinned_pool.back(); g_pinned_pool.pop_back(); }
  }
  if (!c->h_stats) CK(cudaMallocHost(&c->h_stats, 256));
  CK(cudaMemsetAsync(c->d_stats, 0, sizeof(PartStats), c->stream));
  // byte text: every position that is not a real symbol reads 0xff, padding words included (make_key<8>)
  CK(cudaMemsetAsync(c->d_text, c->bits == 8 ? 0xff : 0, c->ntextwords * 8, c->stream));
  CK(cudaMemsetAsync(c->d_text + c->ntextwords, 0, c->nmaskwords * 8, c->stream));
  if (seq->twobitencoding) {
    uint64_t nw = (n + 31) / 32;
    if (seq->numofwords < nw) return set_error("twobitencoding has %llu words, %llu needed",
                                               (unsigned long long)seq->numofwords, (unsigned long long)nw);
    {
      ProfScope ps(c, KC_PREP, 0);
      if (nw) CK(cudaMemcpyAsync(c->d_text, seq->twobitencoding, nw * 8, h2d, c->stream));
    }
    c->nranges = seq->numofspecialranges;
    CK(dmalloc(c, &c->d_ranges, (c->nranges + 1) * sizeof(gtb200_range)));
    if (c->nranges) {
      CK(cudaMemcpyAsync(c->d_ranges, seq->specialranges, c->nranges * sizeof(gtb200_range), h2d, c->stream));
      ProfScope ps(c, KC_PREP);
      k_mask_from_ranges<<<grid_for(c->nranges * 32, 256), 256, 0, c->stream>>>(c->d_mask, c->d_ranges, c->nranges);
    }
  } else {
    const uint8_t *d_enc = seq->encoded;
    uint8_t *tmp = nullptr;
    if (seq->memory != GTB200_MEM_DEVICE) {
      CK(dmalloc(c, &tmp, n ? n : 1));
      if (n) CK(cudaMemcpyAsync(tmp, seq->encoded, n, cudaMemcpyHostToDevice, c->stream));
      d_enc = tmp;
    }
    {
      ProfScope ps(c, KC_PREP);
      if (c->bits == 2) k_pack_bytes<2><<<grid_for(nchunks, 256), 256, 0, c->stream>>>(d_enc, n, c->d_text, c->d_mask, nchunks);
      else k_pack_bytes<8><<<grid_for(nchunks, 256), 256, 0, c->stream>>>(d_enc, n, c->d_text, c->d_mask, nchunks);
    }
    CK(cudaStreamSynchronize(c->stream));
    dfree(c, tmp);
  }
  if (seq->readmode != 0) {
    const int rev = seq->readmode == 1 || seq->readmode == 3, cpl = seq->readmode == 2 || seq->readmode == 3;
    uint64_t *tmp2;
    const size_t words = c->ntextwords + c->nmaskwords;
    CK(dmalloc(c, &tmp2, words * 8));
    CK(cudaMemsetAsync(tmp2, c->bits == 8 ? 0xff : 0, c->ntextwords * 8, c->stream));
    CK(cudaMemsetAsync(tmp2 + c->ntextwords, 0, c->nmaskwords * 8, c->stream));
    {
      ProfScope ps(c, KC_PREP);
      if (c->bits == 2)
        k_readmode_transform<2><<<grid_for(nchunks, 256), 256, 0, c->stream>>>(c->d_text, c->d_mask, n, rev, cpl, tmp2,
                                                                             tmp2 + c->ntextwords, nchunks);
      else
        k_readmode_transform<8><<<grid_for(nchunks, 256), 256, 0, c->stream>>>(c->d_text, c->d_mask, n, rev, cpl, tmp2,
                                                                             tmp2 + c->ntextwords, nchunks);
    }
    CK(cudaMemcpyAsync(c->d_text, tmp2, words * 8, cudaMemcpyDeviceToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    dfree(c, tmp2);
    dfree(c, c->d_ranges);          // the host's ranges describe the forward sequence
    c->d_ranges = nullptr;
  }
  if (!c->d_ranges) {
    // derive the special ranges from the bitmap
    uint32_t *cnt, *popc;
    CK(dmalloc(c, &cnt, (nchunks + 1) * 4));
    CK(dmalloc(c, &popc, (nchunks + 1) * 4));
    {
      ProfScope ps(c, KC_PREP);
      k_runs_count<<<grid_for(nchunks, 256), 256, 0, c->stream>>>(c->d_mask, n, nchunks, cnt, popc);
    }
    if (device_scan(c, cnt, cnt, nchunks, 0, &c->d_stats->scan_total, KC_PREP)) return -1;
    if (sync_stats(c)) return -1;
    c->nranges = c->h_stats->scan_total;
    CK(dmalloc(c, &c->d_ranges, (c->nranges + 1) * sizeof(gtb200_range)));
    if (c->nranges) {
      ProfScope ps(c, KC_PREP);
      k_runs_write<<<grid_for(nchunks, 256), 256, 0, c->stream>>>(c->d_mask, n, nchunks, cnt, c->d_ranges);
    }
    CK(cudaStreamSynchronize(c->stream));
    dfree(c, cnt);
    dfree(c, popc);
  }
  {
    ProfScope ps(c, KC_PREP);
    const uint64_t tailwords = c->nmaskwords - (n >> 6);
    k_mask_tail<<<grid_for(tailwords, 64), 64, 0, c->stream>>>(c->d_mask, n, c->nmaskwords);
  }
  // The sorters' gather table, built after the bitmap is final (ranges, readmode transform, tail bits) - only while
  // it is small enough (<= 64 MB) to stay in the 126 MB L2 next to the text and the streams passing through.
  // Measured: 4.6 Mbp sort 92 -> 71 us with it; at 250 Mbp (125 MB table) it evicts the text and costs 9 %;
  // at 3.1 Gbp neither copy fits, DRAM fetches 64 bytes per random access either way: no difference.
  // GTB200_GATHER = 0 never, 2 always (testing knob).
  int gather_mode = 1;
  if (const char *eg = getenv("GTB200_GATHER")) gather_mode = atoi(eg);
  const uint64_t gather_bytes = (c->ntextwords - 1) * sizeof(ulonglong2);
  c->d_gather = nullptr;
  if (gather_mode == 2 || (gather_mode == 1 && gather_bytes <= (64ull << 20))) {
    const uint64_t nrec = c->ntextwords - 1;
    CK(dmalloc(c, &c->d_gather, nrec * sizeof(ulonglong2)));
    ProfScope ps(c, KC_PREP);
    if (c->bits == 2) k_build_gather<2><<<grid_for(nrec, 256), 256, 0, c->stream>>>(c->d_text, c->d_mask, nrec, c->d_gather);
    else k_build_gather<8><<<grid_for(nrec, 256), 256, 0, c->stream>>>(c->d_text, c->d_mask, nrec, c->d_gather);
  }
  // tail paging table: exclusive scan of range lengths; total = specialcharacters
  CK(dmalloc(c, &c->d_rangeoff, (c->nranges + 2) * 4));
  if (c->nranges) {
    {
      ProfScope ps(c, KC_PREP);
      k_range_lengths<<<grid_for(c->nranges, 256), 256, 0, c->stream>>>(c->d_ranges, c->nranges, c->d_rangeoff);
    }
    if (device_scan(c, c->d_rangeoff, c->d_rangeoff, c->nranges, 1, &c->d_stats->scan_total, KC_PREP)) return -1;
  } else {
    CK(cudaMemsetAsync(c->d_rangeoff, 0, 8, c->stream));
    CK(cudaMemsetAsync(&c->d_stats->scan_total, 0, 8, c->stream));
  }
  if (sync_stats(c)) return -1;
  c->specialchars = c->h_stats->scan_total;
  c->nsort = n - c->specialchars;
  CK(cudaGetLastError());
  return 0;
}

// A context without a sequence: alphabet and bucket-table geometry, stream (reused from closed contexts), knobs.
static gtb200_ctx *ctx_create(int device, unsigned numofchars, uint64_t totallength, unsigned prefixlength) {
  if (cudaSetDevice(device) != cudaSuccess) { set_error("cudaSetDevice(%d) failed", device); return nullptr; }
  gtb200_ctx *c = new gtb200_ctx();
  c->device = device;
  c->n = totallength;
  c->bits = numofchars <= 4 ? 2 : 8;
  c->cp.sigma = numofchars;
  c->cp.pl = prefixlength;
  c->cp.pow2 = (numofchars == (1u << c->bits)) ? 1 : 0;
  const unsigned maxpl = c->bits == 2 ? 16 : 8;
  if (prefixlength > maxpl) { delete c; set_error("gtb200_open: prefixlength %u > %u unsupported", prefixlength, maxpl); return nullptr; }
  uint64_t pw = 1;
  for (unsigned i = 0; i < 34; ++i) {
    c->cp.pw[i] = pw;
    if (pw < (1ull << 40)) pw *= numofchars;
  }
  c->numcodes = c->cp.pw[prefixlength];
  if (c->numcodes > (1ull << 31)) {
    const unsigned long long nc = c->numcodes;
    delete c;
    set_error("gtb200_open: %llu buckets are too many", nc);
    return nullptr;
  }
  c->numspecialcodes = c->cp.pw[prefixlength - 1];
  c->numdist = 0;
  for (unsigned i = 1; i + 1 < prefixlength; ++i) c->numdist += c->cp.pw[i];
  {
    std::lock_guard<std::mutex> lock(g_pool_mu);
    if (!g_stream_pool[device].empty()) {
      const StreamSet ss = g_stream_pool[device].back();
      g_stream_pool[device].pop_back();
      c->stream = ss.main;
      c->copy_stream = ss.copy;
      c->side_stream = ss.side;
      for (int k = 0; k < 8; ++k) c->slice_ev[k] = ss.ev[k];
      for (int k = 0; k < 2; ++k) c->side_ev[k] = ss.sev[k];
    }
  }
  if (!c->stream && cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
    delete c;
    set_error("gtb200_open: cudaStreamCreate failed");
    return nullptr;
  }
  c->sort_maxbucket = SORT_MAXBUCKET;
  if (const char *e2 = getenv("GTB200_MAXBUCKET")) {   // testing knob: force buckets into the refinement path
    const long v = atol(e2);
    if (v >= 0 && v < SORT_MAXBUCKET) c->sort_maxbucket = (uint32_t)v;
  }
  c->pipeline = 1;
  if (const char *e7 = getenv("GTB200_PIPELINE")) c->pipeline = atoi(e7);
  c->narrow = 1;
  if (const char *e12 = getenv("GTB200_NARROW")) c->narrow = atoi(e12);
  {
    unsigned hw = std::thread::hardware_concurrency();
    c->widen_threads = hw == 0 ? 4 : hw > 16 ? 16 : (int)hw;
    if (const char *e13 = getenv("GTB200_WIDEN_THREADS")) c->widen_threads = atoi(e13) > 0 ? atoi(e13) : 1;
  }
  c->overlap = 1;
  if (const char *e9 = getenv("GTB200_OVERLAP")) c->overlap = atoi(e9);
  c->fused_widen = 1;
  if (const char *e8 = getenv("GTB200_FUSED_WIDEN")) c->fused_widen = atoi(e8);
  c->doubling = 1;
  if (const char *e6 = getenv("GTB200_DOUBLING")) c->doubling = atoi(e6);
  c->extend = 2;
  if (const char *e7 = getenv("GTB200_EXTEND")) c->extend = atoi(e7);
  c->tile_mode = 1;
  if (const char *e4 = getenv("GTB200_TILE")) c->tile_mode = atoi(e4);
  c->keyed_mode = 1;
  c->subsort = 1;
  if (const char *e14 = getenv("GTB200_SUBSORT")) c->subsort = atoi(e14);
  if (const char *e10 = getenv("GTB200_KEYED")) c->keyed_mode = atoi(e10);
  if (const char *e5 = getenv("GTB200_TILE_LEVELS")) c->tile_levels = atoi(e5);
  c->twolevel_bytes = 48ull << 20;
  if (const char *e3 = getenv("GTB200_TWOLEVEL_BYTES")) c->twolevel_bytes = strtoull(e3, nullptr, 10);   // testing knob
  return c;
}

extern "C" int gtb200_open(gtb200_ctx **out, int device, const gtb200_seqdesc *seq, unsigned prefixlength) {
  if (!out || !seq) return set_error("gtb200_open: null argument");
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return set_error("gtb200_open: no CUDA device available (%s); this library has no CPU fallback",
                     e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
  if (device < 0 || device >= ndev) return set_error("gtb200_open: device %d out of range (0..%d)", device, ndev - 1);
  if (seq->totallength >= 0xfffffffeull)
    return set_error("gtb200_open: totallength %llu does not fit the 32-bit bucket table (reference: bcktab.c:258-261)",
                     (unsigned long long)seq->totallength);
  if (seq->numofchars < 2 || seq->numofchars > 253) return set_error("gtb200_open: numofchars %u unsupported", seq->numofchars);
  if (!seq->twobitencoding && !seq->encoded && seq->totallength)
    return set_error("gtb200_open: neither twobitencoding nor encoded given");
  if (seq->twobitencoding && seq->numofchars != 4) return set_error("gtb200_open: twobitencoding requires numofchars == 4");
  if (seq->twobitencoding && seq->numofspecialranges && !seq->specialranges)
    return set_error("gtb200_open: specialranges missing");
  if (prefixlength < 1) return set_error("gtb200_open: prefixlength must be >= 1");
  if (seq->readmode < 0 || seq->readmode > 3) return set_error("gtb200_open: readmode %d unknown", seq->readmode);
  if ((seq->readmode == 2 || seq->readmode == 3) && seq->numofchars != 4)
    return set_error("gtb200_open: complement readmodes need the DNA alphabet");
  gtb200_ctx *c = ctx_create(device, seq->numofchars, seq->totallength, prefixlength);
  if (!c) return -1;
  if (open_impl(c, seq)) { free_ctx(c); return -1; }
  *out = c;
  return 0;
}

// Side stream: fork = it waits for everything issued on the main stream so far, join = the main stream waits for it.
// Off while kernel classes are being timed (their events sit on the main stream).
static bool side_usable(gtb200_ctx *c) {
  if (!c->overlap || c->prof_on) return false;
  if (c->side_stream) return true;
  if (cudaStreamCreateWithFlags(&c->side_stream, cudaStreamNonBlocking) != cudaSuccess) { c->side_stream = nullptr; return false; }
  for (int k = 0; k < 2; ++k)
    if (cudaEventCreateWithFlags(&c->side_ev[k], cudaEventDisableTiming) != cudaSuccess) {
      c->overlap = 0;                       // stay on one stream for the rest of this context's life
      return false;
    }
  return true;
}
static int side_fork(gtb200_ctx *c) {
  CK(cudaEventRecord(c->side_ev[0], c->stream));
  CK(cudaStreamWaitEvent(c->side_stream, c->side_ev[0], 0));
  return 0;
}
static int side_join(gtb200_ctx *c) {
  CK(cudaEventRecord(c->side_ev[1], c->side_stream));
  CK(cudaStreamWaitEvent(c->stream, c->side_ev[1], 0));
  return 0;
}

template <int BITS>
static int count_impl(gtb200_ctx *c, bool need_host_values, uint64_t pos_lo, uint64_t pos_hi, bool padded_counts,
                      int phase /* 0 = everything, 1 = counting only, 2 = finish only */) {
  const Text<BITS> T = text_of<BITS>(c);
  const uint64_t n = c->n;
  const uint64_t nr1 = c->nranges + 1;
  const uint64_t tablewords = (c->numcodes + 2) + (c->numspecialcodes + 1) + (c->numdist + 1);
  if (!c->d_lb) {                            // first call: allocate the tables once, contiguously
    c->padcap = nr1 * (c->cp.pl - 1) + 1;
    // the tables and the 64-byte block of persistent counters share one allocation: one memset clears both
    CK(dmalloc(c, &c->d_lb, (tablewords + (tablewords & 1)) * 4 + 64));
    c->d_persist = (unsigned long long *)(c->d_lb + tablewords + (tablewords & 1));
    c->d_csc = c->d_lb + (c->numcodes + 2);
    c->d_dist = c->d_csc + (c->numspecialcodes + 1);
    CK(dmalloc(c, &c->d_cursor, (c->numcodes + 2) * 4));
    CK(dmalloc(c, &c->d_padcnt, (nr1 + 1) * 4));
    CK(dmalloc(c, &c->d_padkey, c->padcap * 8));
    CK(dmalloc(c, &c->d_padpos, c->padcap * 4));
  }
  const bool small = nr1 <= PADSMALL_CAP && c->padcap <= PADSMALL_CAP;
  uint32_t *cntoff = c->d_padcnt;
  uint32_t *hist_or_null = padded_counts ? c->d_lb : nullptr;
  if (phase != 2) {
  CK(cudaMemsetAsync(c->d_lb, 0, (tablewords + (tablewords & 1)) * 4 + 64, c->stream));   // tables + persistent counters
  // the short list of special-padded suffixes is built beside the k-mer count (both only add into the zeroed tables)
  const bool beside = small && side_usable(c);
  if (beside && side_fork(c)) return -1;
  if (pos_hi > pos_lo) {
    // A table larger than L2 is filled slice by slice (64 MB of counters per pass over the text), so
    // that the RED.ADDs always hit L2-resident lines; the text scan itself is cheap (n/4 bytes).
    uint64_t slice = 1ull << 24;
    if (const char *es = getenv("GTB200_COUNT_SLICE_LOG2")) slice = 1ull << atoi(es);     // testing knob
    // large DNA texts: the funnel-shift scanner (the per-position cost is what bounds the sliced count)
    int fastmode = 1;                                  // GTB200_FASTCOUNT: 0 never, 2 always (testing knob)
    if (const char *ef = getenv("GTB200_FASTCOUNT")) fastmode = atoi(ef);
    const bool fast = BITS == 2 && c->cp.pow2 && c->cp.pl <= 15 && fastmode != 0 &&
                      (fastmode == 2 || (pos_hi - pos_lo) >= (32ull << 20));
    const Text<2> T2 = text_of<2>(c);
    if (c->numcodes * 4 <= (96ull << 20)) {
      ProfScope ps(c, KC_COUNT);
      if (fast)
        k_count_dna<<<grid_for((pos_hi - pos_lo + 15) / 16, 256), 256, 0, c->stream>>>(T2, c->cp.pl, pos_lo, pos_hi, c->d_lb, 0u,
                                                                                     (uint32_t)(c->numcodes - 1));
      else
        k_kmers<BITS, 0><<<grid_for((pos_hi - pos_lo + KM_PPT - 1) / KM_PPT, 256), 256, 0, c->stream>>>(T, c->cp, pos_lo, pos_hi, c->d_lb, nullptr, 0, 0, 0);
    } else {
      for (uint64_t lo = 0; lo < c->numcodes; lo += slice) {
        const uint64_t hi = lo + slice - 1 < c->numcodes - 1 ? lo + slice - 1 : c->numcodes - 1;
        ProfScope ps(c, KC_COUNT);
        if (fast)
          k_count_dna<<<grid_for((pos_hi - pos_lo + 15) / 16, 256), 256, 0, c->stream>>>(T2, c->cp.pl, pos_lo, pos_hi, c->d_lb,
                                                                                       (uint32_t)lo, (uint32_t)hi);
        else
          k_kmers<BITS, 2><<<grid_for((pos_hi - pos_lo + KM_PPT - 1) / KM_PPT, 256), 256, 0, c->stream>>>(T, c->cp, pos_lo, pos_hi, c->d_lb, nullptr, lo, hi, 0);
      }
    }
  }
  // special-padded suffixes
  if (small) {
    ProfScope ps(c, KC_PADDED);
    k_padded_small<BITS><<<1, 256, 0, beside ? c->side_stream : c->stream>>>(T, c->cp, n, c->d_ranges, c->nranges, hist_or_null,
                                                                             c->d_csc, c->d_dist, c->d_padkey, c->d_padpos,
                                                                             c->d_persist);
    if (beside && side_join(c)) return -1;
  } else {
    {
      ProfScope ps(c, KC_PADDED);
      k_padded<BITS, 0><<<grid_for(nr1, 128), 128, 0, c->stream>>>(T, c->cp, n, c->d_ranges, c->nranges, cntoff, nullptr,
                                                                   nullptr, nullptr, nullptr, nullptr);
    }
    if (device_scan(c, cntoff, cntoff, nr1, 0, c->d_persist, KC_PADDED)) return -1;
    {
      ProfScope ps(c, KC_PADDED);
      k_padded<BITS, 1><<<grid_for(nr1, 128), 128, 0, c->stream>>>(T, c->cp, n, c->d_ranges, c->nranges, cntoff,
                                                                   hist_or_null, c->d_csc, c->d_dist, c->d_padkey, c->d_padpos);
    }
  }
  }   // phase != 2
  if (phase == 1) {
    CK(cudaGetLastError());
    return 0;
  }
  // prefix sum of the bucket counts; the largest bucket is reduced in the same launch
  if (device_scan(c, c->d_lb, c->d_lb, c->numcodes, 1, c->d_persist + 2, KC_SCAN, (unsigned int *)(c->d_persist + 1))) return -1;
  c->counted = true;
  if (small && !need_host_values) {           // nothing the host needs: stay asynchronous
    CK(cudaGetLastError());
    return 0;
  }
  unsigned long long *hp = (unsigned long long *)((char *)c->h_stats + 128);
  CK(cudaMemcpyAsync(hp, c->d_persist, 24, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  c->npadded = hp[0];
  c->largestbucket = (uint32_t)hp[1];
  if (hp[2] != c->nsort)
    return set_error("internal: bucket counts sum to %llu, expected %llu", hp[2], (unsigned long long)c->nsort);
  // large lists: sort by (code, pl-L); stable, so positions stay ascending
  if (!small && c->npadded > 1) {
    const uint32_t np = (uint32_t)c->npadded;
    const uint32_t nblocks = grid_for(np, RS_TILE);
    const size_t need = (size_t)np * (8 + 8 + 4 + 4) + (256ull * nblocks + 1) * 4 + 64;
    if (ensure_scratch(c, need)) return -1;
    char *s = (char *)c->d_scratch;
    uint64_t *kA = (uint64_t *)s; s += (size_t)np * 8;
    uint64_t *kB = (uint64_t *)s; s += (size_t)np * 8;
    uint32_t *vA = (uint32_t *)s; s += (size_t)np * 4;
    uint32_t *vB = (uint32_t *)s; s += (size_t)np * 4;
    uint32_t *hist = (uint32_t *)s;
    const uint64_t *ko;
    const uint32_t *vo;
    const int hib = 6 + bits_for(c->numcodes - 1);
    if (device_radix_sort(c, c->d_padkey, c->d_padpos, kA, vA, kB, vB, hist, np, 0, hib, &ko, &vo)) return -1;
    CK(cudaMemcpyAsync(c->d_padkey, ko, (size_t)np * 8, cudaMemcpyDeviceToDevice, c->stream));
    CK(cudaMemcpyAsync(c->d_padpos, vo, (size_t)np * 4, cudaMemcpyDeviceToDevice, c->stream));
  }
  CK(cudaGetLastError());
  c->counted = true;
  return 0;
}

extern "C" int gtb200_count(gtb200_ctx *c, uint32_t *leftborder, uint32_t *countspecialcodes, uint32_t *distpfxidx,
                            gtb200_countinfo *info) {
  if (!c) return set_error("gtb200_count: null context");
  CK(cudaSetDevice(c->device));
  {                                           // recomputed on every call (the tables are reused)
    const bool need = leftborder || countspecialcodes || distpfxidx || info;
    const int rc = c->bits == 2 ? count_impl<2>(c, need, 0, c->n, true, 0) : count_impl<8>(c, need, 0, c->n, true, 0);
    if (rc) return rc;
  }
  if (leftborder) CK(cudaMemcpyAsync(leftborder, c->d_lb, (c->numcodes + 1) * 4, cudaMemcpyDeviceToHost, c->stream));
  if (countspecialcodes)
    CK(cudaMemcpyAsync(countspecialcodes, c->d_csc, c->numspecialcodes * 4, cudaMemcpyDeviceToHost, c->stream));
  if (distpfxidx && c->numdist)
    CK(cudaMemcpyAsync(distpfxidx, c->d_dist, c->numdist * 4, cudaMemcpyDeviceToHost, c->stream));
  if (leftborder || countspecialcodes || distpfxidx || info) CK(cudaStreamSynchronize(c->stream));
  if (info) {
    info->numofallcodes = c->numcodes;
    info->numofspecialcodes = c->numspecialcodes;
    info->numofdistpfxidx = c->numdist;
    info->numofsuffixestosort = c->nsort;
    info->specialcharacters = c->specialchars;
    info->largestbucket = c->largestbucket;
  }
  return 0;
}

extern "C" int gtb200_count_shard(gtb200_ctx *c, uint64_t pos_lo, uint64_t pos_hi, int with_padded_counts,
                                  void **tables_dev, uint64_t *ntablewords) {
  if (!c) return set_error("gtb200_count_shard: null context");
  if (pos_hi > c->n) pos_hi = c->n;
  if (pos_lo > pos_hi || (pos_lo & 63) || (pos_hi != c->n && (pos_hi & 63)))
    return set_error("gtb200_count_shard: [%llu,%llu) must be 64-aligned inside the sequence",
                     (unsigned long long)pos_lo, (unsigned long long)pos_hi);
  CK(cudaSetDevice(c->device));
  const int rc = c->bits == 2 ? count_impl<2>(c, false, pos_lo, pos_hi, with_padded_counts != 0, 1)
                              : count_impl<8>(c, false, pos_lo, pos_hi, with_padded_counts != 0, 1);
  if (rc) return rc;
  if (tables_dev) *tables_dev = c->d_lb;
  if (ntablewords) *ntablewords = (c->numcodes + 2) + (c->numspecialcodes + 1) + (c->numdist + 1);
  return 0;
}

extern "C" int gtb200_count_finish(gtb200_ctx *c, uint32_t *leftborder, uint32_t *countspecialcodes,
                                   uint32_t *distpfxidx, gtb200_countinfo *info) {
  if (!c) return set_error("gtb200_count_finish: null context");
  if (!c->d_lb) return set_error("gtb200_count_finish: call gtb200_count_shard first");
  CK(cudaSetDevice(c->device));
  const bool need = leftborder || countspecialcodes || distpfxidx || info;
  const int rc = c->bits == 2 ? count_impl<2>(c, need, 0, 0, true, 2) : count_impl<8>(c, need, 0, 0, true, 2);
  if (rc) return rc;
  if (leftborder) CK(cudaMemcpyAsync(leftborder, c->d_lb, (c->numcodes + 1) * 4, cudaMemcpyDeviceToHost, c->stream));
  if (countspecialcodes)
    CK(cudaMemcpyAsync(countspecialcodes, c->d_csc, c->numspecialcodes * 4, cudaMemcpyDeviceToHost, c->stream));
  if (distpfxidx && c->numdist)
    CK(cudaMemcpyAsync(distpfxidx, c->d_dist, c->numdist * 4, cudaMemcpyDeviceToHost, c->stream));
  if (need) CK(cudaStreamSynchronize(c->stream));
  if (info) {
    info->numofallcodes = c->numcodes;
    info->numofspecialcodes = c->numspecialcodes;
    info->numofdistpfxidx = c->numdist;
    info->numofsuffixestosort = c->nsort;
    info->specialcharacters = c->specialchars;
    info->largestbucket = c->largestbucket;
  }
  return 0;
}

static int ensure_part_buffers(gtb200_ctx *c, uint64_t width, uint64_t ncodes_in_part, bool keyed) {
  // keyed tile path: the level-1 records (16 bytes per suffix) pass through the image buffer, which is dead until
  // the tile sorter writes the 8-byte .suf entries into its first half
  const uint64_t suf64_bytes = (width + 1) * (keyed ? 16 : 8);
  if (width > c->cap_width || suf64_bytes > c->cap_suf64_bytes) {
    dfree(c, c->d_sa);
    dfree(c, c->d_lcp);
    dfree(c, c->d_suf64);
    dfree(c, c->d_medlist);
    c->d_sa = nullptr; c->d_lcp = nullptr; c->d_suf64 = nullptr; c->d_medlist = nullptr;
    c->cap_width = 0; c->cap_suf64_bytes = 0;
    CK(dmalloc(c, &c->d_sa, (width + 1) * 4));
    CK(dmalloc(c, &c->d_lcp, width + 1));
    CK(dmalloc(c, &c->d_suf64, suf64_bytes));
    CK(dmalloc(c, &c->d_medlist, (width / (SORT_MAXBUCKET + 1) + 2) * 4));
    c->cap_width = width;
    c->cap_suf64_bytes = suf64_bytes;
  }
  const uint64_t fbwords = 2 * ((ncodes_in_part + 31) / 32) + 2;   // unsorted-bucket bitmap, then tied-bucket bitmap
  if (fbwords > c->cap_fbwords) {
    dfree(c, c->d_fbbits);
    c->d_fbbits = nullptr; c->cap_fbwords = 0;
    CK(dmalloc(c, &c->d_fbbits, fbwords * 4));
    c->cap_fbwords = fbwords;
  }
  return 0;
}

// refinement of flagged buckets; A elements, nb buckets (both from the sort kernel's counters)
template <int BITS>
static int refine_impl(gtb200_ctx *c, const PartOut &out_in, uint64_t mincode, uint64_t maxcode, uint64_t width,
                       gtb200_partinfo *info) {
  const Text<BITS> T = text_of<BITS>(c);
  const uint32_t A0 = (uint32_t)c->h_stats->fb_elems;
  const uint32_t nb = c->h_stats->fb_buckets;
  const uint64_t fbwords = (maxcode - mincode + 1 + 31) / 32;
  info->fallback_suffixes = A0;
  // llv append buffer: at most one value per refined suffix
  if (A0 > c->cap_llv) {
    dfree(c, c->d_llv);
    c->d_llv = nullptr; c->cap_llv = 0;
    CK(dmalloc(c, &c->d_llv, ((size_t)A0 + 1) * sizeof(gtb200_largelcp)));
    c->cap_llv = A0;
  }
  PartOut out = out_in;
  // scratch carve-up
  const uint32_t nblocks = grid_for(A0, RS_TILE);
  size_t need = 0;
  auto take = [&](size_t bytes) { size_t o = need; need += (bytes + 255) & ~(size_t)255; return o; };
  const size_t o_wcnt = take((fbwords + 1) * 4), o_codes = take(((size_t)nb + 1) * 4), o_sizes = take(((size_t)nb + 1) * 4);
  const size_t o_kind = take(((size_t)nb + 1) * 4);
  const size_t a4 = ((size_t)A0 + 1) * 4, a8 = ((size_t)A0 + 1) * 8;
  const size_t o_idx0 = take(a4), o_idx1 = take(a4), o_gs0 = take(a4), o_gs1 = take(a4), o_dep0 = take(a4), o_dep1 = take(a4);
  const size_t o_keys = take(a8), o_kA = take(a8), o_kB = take(a8), o_vals = take(a4), o_vA = take(a4), o_vB = take(a4);
  const size_t o_pos = take(a4), o_head = take(a4), o_hscan = take(a4), o_headpos = take(a4 + 4), o_keep = take(a4);
  const size_t o_kscan = take(a4), o_newgs = take(a4), o_newdep = take(a4), o_hist = take((256ull * nblocks + 1) * 4);
  const size_t o_llvtmp = take(((size_t)A0 + 1) * sizeof(gtb200_largelcp));
  const size_t o_idx00 = take(a4), o_gs00 = take(a4), o_ext = take(a4);
  const size_t o_pvt = take(a8), o_kpv = take(a8), o_ebuf = take(a4), o_cbuf = take(a4);
  if (ensure_scratch(c, need)) return -1;
  char *S = (char *)c->d_scratch;
  uint32_t *wcnt = (uint32_t *)(S + o_wcnt), *codes = (uint32_t *)(S + o_codes), *sizes = (uint32_t *)(S + o_sizes);
  uint32_t *kind = (uint32_t *)(S + o_kind);
  const uint32_t *tiebits = c->d_fbbits + fbwords;
  uint32_t *idx[2] = {(uint32_t *)(S + o_idx0), (uint32_t *)(S + o_idx1)};
  uint32_t *gs[2] = {(uint32_t *)(S + o_gs0), (uint32_t *)(S + o_gs1)};
  uint32_t *dep[2] = {(uint32_t *)(S + o_dep0), (uint32_t *)(S + o_dep1)};
  uint64_t *keys = (uint64_t *)(S + o_keys), *kA = (uint64_t *)(S + o_kA), *kB = (uint64_t *)(S + o_kB);
  uint32_t *vals = (uint32_t *)(S + o_vals), *vA = (uint32_t *)(S + o_vA), *vB = (uint32_t *)(S + o_vB);
  uint32_t *posbuf = (uint32_t *)(S + o_pos), *head = (uint32_t *)(S + o_head), *hscan = (uint32_t *)(S + o_hscan);
  uint32_t *headpos = (uint32_t *)(S + o_headpos), *keep = (uint32_t *)(S + o_keep), *kscan = (uint32_t *)(S + o_kscan);
  uint32_t *newgs = (uint32_t *)(S + o_newgs), *newdep = (uint32_t *)(S + o_newdep), *hist = (uint32_t *)(S + o_hist);
  gtb200_largelcp *llvtmp = (gtb200_largelcp *)(S + o_llvtmp);
  uint32_t *idx00 = (uint32_t *)(S + o_idx00), *gs00 = (uint32_t *)(S + o_gs00), *extmin = (uint32_t *)(S + o_ext);
  unsigned long long *pvt = (unsigned long long *)(S + o_pvt);
  uint64_t *kpv = (uint64_t *)(S + o_kpv);
  uint32_t *ebuf = (uint32_t *)(S + o_ebuf), *cbuf = (uint32_t *)(S + o_cbuf);
  out.llv = llvtmp;

  // expand the two flagged-bucket bitmaps into the ordered element list
  {
    ProfScope ps(c, KC_REFINE, 1);
    k_fb_popc<<<grid_for(fbwords, 256), 256, 0, c->stream>>>(c->d_fbbits, tiebits, fbwords, wcnt);
  }
  if (device_scan(c, wcnt, wcnt, fbwords, 0, nullptr, KC_REFINE)) return -1;
  {
    ProfScope ps(c, KC_REFINE, 2);
    k_fb_codes<<<grid_for(fbwords, 256), 256, 0, c->stream>>>(c->d_fbbits, tiebits, fbwords, wcnt, mincode, codes, kind);
    k_fb_list<BITS, false><<<grid_for((uint64_t)nb * 32, 256), 256, 0, c->stream>>>(
        T, out.sa, codes, kind, nb, c->d_lb, c->d_csc, c->cp.sigma, out.partoffset, c->cp.pl, sizes, nullptr, nullptr,
        nullptr, nullptr);
  }
  if (device_scan(c, sizes, sizes, nb, 0, &c->d_stats->scan_total2, KC_REFINE)) return -1;
  {
    ProfScope ps(c, KC_REFINE, 1);
    k_fb_list<BITS, true><<<grid_for((uint64_t)nb * 32, 256), 256, 0, c->stream>>>(
        T, out.sa, codes, kind, nb, c->d_lb, c->d_csc, c->cp.sigma, out.partoffset, c->cp.pl, nullptr, sizes, idx[0], gs[0],
        dep[0]);
  }
  CK(cudaGetLastError());
  if (sync_stats(c)) return -1;
  const uint32_t A1 = (uint32_t)c->h_stats->scan_total2;     // suffixes that really need refinement
  info->fallback_suffixes = A1;
  uint32_t A = A1;
  int cur = 0;
  uint64_t rounds = 0;
  const int gsbits = ((bits_for(width) + 7) / 8) * 8;
  const bool whole = mincode == 0 && maxcode == c->numcodes - 1 && out.partoffset == 0;
  // Text-window rounds: the only strategy for sharded parts; for whole-array parts they run first when few
  // suffixes are open (building the rank array costs a pass over the whole suffix array), and as long as a
  // group cut by a special still has to be ordered by position.
  const bool can_double = whole && c->doubling;
  const uint64_t text_limit = !can_double ? ~0ull : ((uint64_t)A1 * 256 < c->n ? 6 : 0);
  bool posmode_open = false;
  while (A > 0 && (rounds < text_limit || posmode_open)) {
    rounds++;
    const unsigned g = grid_for(A, 256);
    // what survived the first round is mostly repeats (c->extend: 1 = groups jump over the windows all members share,
    // 2 = pivot rounds, see k_fb_keys_pivot)
    const bool pivot = !can_double && c->extend == 2 && rounds >= 2;
    if (pivot) {
      CK(cudaMemsetAsync(pvt, 0, (size_t)A * 8, c->stream));
      ProfScope ps(c, KC_REFINE, 2);
      k_fb_pivot_pick<<<g, 256, 0, c->stream>>>(out.sa, idx[cur], gs[cur], dep[cur], A, (uint32_t)rounds * 0x632BE5ABu, pvt);
      k_fb_keys_pivot<BITS><<<g, 256, 0, c->stream>>>(T, out.sa, idx[cur], gs[cur], dep[cur], A, pvt, keys, kpv, ebuf, cbuf,
                                                       posbuf, vals);
    } else {
      if (!can_double && c->extend == 1 && rounds >= 2) {
        CK(cudaMemsetAsync(extmin, 0xff, (size_t)A * 4, c->stream));
        ProfScope ps(c, KC_REFINE, 2);
        k_fb_extend_count<BITS><<<g, 256, 0, c->stream>>>(T, out.sa, idx[cur], gs[cur], dep[cur], A, 4096u, extmin);
        k_fb_extend_apply<<<g, 256, 0, c->stream>>>(idx[cur], gs[cur], A, (unsigned)Text<BITS>::W, extmin, dep[cur]);
      }
      ProfScope ps(c, KC_REFINE, 1);
      k_fb_keys<BITS><<<g, 256, 0, c->stream>>>(T, out.sa, idx[cur], dep[cur], A, keys, posbuf, vals);
    }
    const uint64_t *ko;
    const uint32_t *vo;
    if (device_radix_sort(c, keys, vals, kA, vA, kB, vB, hist, A, 0, 64, &ko, &vo)) return -1;
    // second key: group start; vo must not be one of the output buffers of the next sort's first pass
    uint64_t *k2 = (ko == kA) ? kB : kA;   // free key buffer
    {
      ProfScope ps(c, KC_REFINE, 1);
      if (pivot) k_fb_gather_gsc<<<g, 256, 0, c->stream>>>(vo, gs[cur], cbuf, A, k2);
      else k_fb_gather_gs<<<g, 256, 0, c->stream>>>(vo, gs[cur], A, k2);
    }
    // copy vo into vals so that both ping-pong value buffers are free again
    CK(cudaMemcpyAsync(vals, vo, (size_t)A * 4, cudaMemcpyDeviceToDevice, c->stream));
    uint64_t *k3a = (k2 == kA) ? kB : kA;
    // keys buffer `keys` must stay intact (regroup reads it), so ping-pong between k3a and a third buffer:
    // reuse posbuf? no - use kscan/newgs area is too small; sort (k2 -> k3a -> k2 ...) is safe because
    // device_radix_sort never writes to its input on the first pass and alternates afterwards.
    const uint64_t *ko2;
    const uint32_t *vo2;
    if (device_radix_sort(c, k2, vals, k3a, vA, k2, vB, hist, A, 0, pivot ? 16 + gsbits : gsbits, &ko2, &vo2)) return -1;
    {
      ProfScope ps(c, KC_REFINE, 1);
      if (pivot)
        k_fb_regroup_pivot<BITS><<<g, 256, 0, c->stream>>>(out, vo2, idx[cur], gs[cur], dep[cur], keys, kpv, ebuf, cbuf, posbuf, A,
                                                            head, newdep);
      else
        k_fb_regroup<BITS><<<g, 256, 0, c->stream>>>(out, vo2, idx[cur], gs[cur], dep[cur], keys, posbuf, A, head, newdep);
    }
    if (device_scan(c, head, hscan, A, 0, &c->d_stats->scan_total, KC_REFINE)) return -1;
    {
      ProfScope ps(c, KC_REFINE, 2);
      k_fb_headpos<<<g, 256, 0, c->stream>>>(head, hscan, A, headpos, &c->d_stats->scan_total);
      k_fb_finalize<<<g, 256, 0, c->stream>>>(head, hscan, headpos, idx[cur], A, keep, newgs);
    }
    if (device_scan(c, keep, kscan, A, 0, &c->d_stats->scan_total2, KC_REFINE)) return -1;
    {
      // newdep is indexed by sorted position j, like idx; gather of dep happened in regroup
      ProfScope ps(c, KC_REFINE, 1);
      k_reset_posmode<<<1, 1, 0, c->stream>>>(c->d_stats);
      k_fb_compact<<<g, 256, 0, c->stream>>>(keep, kscan, idx[cur], newgs, newdep, A, idx[cur ^ 1], gs[cur ^ 1], dep[cur ^ 1],
                                             c->d_stats);
    }
    CK(cudaGetLastError());
    if (sync_stats(c)) return -1;
    A = (uint32_t)c->h_stats->scan_total2;
    posmode_open = c->h_stats->posmode_kept > 0;
    cur ^= 1;
    if (rounds > (c->n / Text<BITS>::W) + 8) return set_error("internal: refinement does not converge");
  }
  if (A > 0) {                                 // what text rounds left open (or everything): prefix doubling
    const uint32_t AD = A;                       // list size when doubling starts
    // ---- prefix doubling on ranks, then LCP of the refined entries by Phi/Kasai in text order ----
    if (c->n + 2 > c->cap_isa) {
      dfree(c, c->d_isa);
      c->d_isa = nullptr;
      c->cap_isa = 0;
      CK(dmalloc(c, &c->d_isa, (c->n + 2) * 4));
      c->cap_isa = c->n + 2;
    }
    uint32_t *isa = c->d_isa;
    CK(cudaMemcpyAsync(idx00, idx[cur], (size_t)A * 4, cudaMemcpyDeviceToDevice, c->stream));
    CK(cudaMemcpyAsync(gs00, gs[cur], (size_t)A * 4, cudaMemcpyDeviceToDevice, c->stream));
    {
      ProfScope ps(c, KC_REFINE, 3);
      k_isa_from_sa<<<grid_for(width, 256), 256, 0, c->stream>>>(out.sa, width, isa);
      k_isa_tail<<<grid_for(c->specialchars + 1, 256), 256, 0, c->stream>>>(c->d_ranges, c->d_rangeoff, c->nranges,
                                                                            c->specialchars, c->n, c->nsort, isa);
      k_isa_groups<<<grid_for(A, 256), 256, 0, c->stream>>>(idx[cur], gs[cur], out.sa, A, isa);
    }
    uint64_t h = c->h_stats->fb_unsorted ? c->cp.pl : c->cp.pl + Text<BITS>::W;   // every open group is tied >= h
    const int rbits = ((bits_for(c->n + 1) + 7) / 8) * 8;
    while (A > 0) {
      rounds++;
      const unsigned g = grid_for(A, 256);
      {
        ProfScope ps(c, KC_REFINE, 1);
        k_dbl_keys<<<g, 256, 0, c->stream>>>(out.sa, idx[cur], gs[cur], A, h, isa, keys, posbuf, vals);
      }
      const uint64_t *ko, *ko2;
      const uint32_t *vo, *vo2;
      if (device_radix_sort(c, keys, vals, kA, vA, kB, vB, hist, A, 0, rbits, &ko, &vo)) return -1;
      if (ko == kA) { if (device_radix_sort(c, ko, vo, kB, vB, kA, vA, hist, A, 32, 32 + gsbits, &ko2, &vo2)) return -1; }
      else          { if (device_radix_sort(c, ko, vo, kA, vA, kB, vB, hist, A, 32, 32 + gsbits, &ko2, &vo2)) return -1; }
      {
        ProfScope ps(c, KC_REFINE, 1);
        k_dbl_regroup<<<g, 256, 0, c->stream>>>(out.sa, ko2, vo2, idx[cur], posbuf, A, head);
      }
      if (device_scan(c, head, hscan, A, 0, &c->d_stats->scan_total, KC_REFINE)) return -1;
      {
        ProfScope ps(c, KC_REFINE, 3);
        k_fb_headpos<<<g, 256, 0, c->stream>>>(head, hscan, A, headpos, &c->d_stats->scan_total);
        k_fb_finalize<<<g, 256, 0, c->stream>>>(head, hscan, headpos, idx[cur], A, keep, newgs);
        k_dbl_update_isa<<<g, 256, 0, c->stream>>>(out.sa, idx[cur], newgs, A, isa);
      }
      if (device_scan(c, keep, kscan, A, 0, &c->d_stats->scan_total2, KC_REFINE)) return -1;
      {
        ProfScope ps(c, KC_REFINE, 1);
        k_fb_compact<<<g, 256, 0, c->stream>>>(keep, kscan, idx[cur], newgs, dep[cur], A, idx[cur ^ 1], gs[cur ^ 1],
                                               dep[cur ^ 1], nullptr);
      }
      CK(cudaGetLastError());
      if (sync_stats(c)) return -1;
      A = (uint32_t)c->h_stats->scan_total2;
      cur ^= 1;
      h = h * 2 < c->n ? h * 2 : c->n;
      if (rounds > 70) return set_error("internal: prefix doubling does not converge");
    }
    // LCP of every entry that was not a group head when refinement started
    {
      ProfScope ps(c, KC_REFINE, 1);
      k_phi_flags<<<grid_for(AD, 256), 256, 0, c->stream>>>(idx00, gs00, AD, head);
    }
    if (device_scan(c, head, hscan, AD, 0, &c->d_stats->scan_total, KC_REFINE)) return -1;
    if (sync_stats(c)) return -1;
    const uint32_t U = (uint32_t)c->h_stats->scan_total;
    if (U) {
      {
        ProfScope ps(c, KC_REFINE, 1);
        k_phi_build<<<grid_for(AD, 256), 256, 0, c->stream>>>(idx00, head, hscan, out.sa, AD, keys, vals);
      }
      const uint64_t *ko;
      const uint32_t *vo;
      if (device_radix_sort(c, keys, vals, kA, vA, kB, vB, hist, U, 0, rbits, &ko, &vo)) return -1;
      {
        ProfScope ps(c, KC_REFINE, 1);
        k_phi_lcp<BITS><<<grid_for(((uint64_t)U + PHI_CHUNK - 1) / PHI_CHUNK, 128), 128, 0, c->stream>>>(T, out, ko, vo, U,
                                                                                                    c->cp.pl);
      }
      CK(cudaGetLastError());
      if (sync_stats(c)) return -1;
    }
  }
  info->fallback_rounds = rounds;
  // order the large-lcp records by suffix-array index
  const uint32_t nl = (uint32_t)c->h_stats->llv_count;
  if (nl) {
    const unsigned g = grid_for(nl, 256);
    {
      ProfScope ps(c, KC_REFINE, 1);
      k_llv_keys<<<g, 256, 0, c->stream>>>(llvtmp, nl, keys, vals);
    }
    const uint64_t *ko;
    const uint32_t *vo;
    const int pbits = ((bits_for(out.partoffset + width) + 7) / 8) * 8;
    if (device_radix_sort(c, keys, vals, kA, vA, kB, vB, hist, nl, 0, pbits, &ko, &vo)) return -1;
    {
      ProfScope ps(c, KC_REFINE, 1);
      k_llv_gather<<<g, 256, 0, c->stream>>>(llvtmp, vo, nl, c->d_llv);
    }
    CK(cudaGetLastError());
  }
  return 0;
}

// Everything that can be enqueued for the code range [lo,hi] of the current part without a host round trip:
// insertion, special-padded suffixes, the sorters, bucket borders and the 64-bit image of the slice
// [roff, roff+rwidth) of the part's buffers.  (lo - part mincode) is a multiple of 32.
template <int BITS>
static int enqueue_range(gtb200_ctx *c, const Text<BITS> &T, const PartOut &out, uint64_t pmin, uint64_t lo, uint64_t hi,
                         uint64_t roff, uint64_t rwidth, uint64_t fbwords, bool first_of_part, bool keyed_ok) {
  const uint64_t rcodes = hi - lo + 1;
  uint32_t *fbb = c->d_fbbits + ((lo - pmin) >> 5);
  if (rwidth == 0) return 0;
  if (!first_of_part) {                       // k_part_init zeroed the statistics for the first range
    ProfScope ps(c, KC_INIT);
    k_reset_medcount<<<1, 1, 0, c->stream>>>(c->d_stats);
  }
  // ---- tile path (DNA): staged partition into shared-memory sized bins, then sort inside the tile ----
  bool tile = c->tile_mode != 0 && rcodes >= 2 && (c->tile_mode == 2 || (rwidth >= (1u << 24) && rcodes >= 1024));
  int shiftF = 0, shift1 = 0;
  uint32_t nbinsF = 0, nb1 = 0;
  if (tile) {
    if (rwidth / rcodes > 4096) tile = false;                      // buckets alone exceed a tile
    while (tile && shiftF < 10 && (((uint64_t)rwidth << (shiftF + 1)) / rcodes) <= 4608 && (rcodes >> (shiftF + 1)) >= 1) ++shiftF;
    const uint64_t nf = ((rcodes - 1) >> shiftF) + 1;
    bool two = nf > SP_MAXBINS;
    if (c->tile_levels == 2 && nf >= 4) two = true;
    if (c->tile_levels == 1 && nf <= SP_MAXBINS) two = false;
    if (nf > (1ull << 20)) tile = false;
    nbinsF = (uint32_t)nf;
    if (tile && two) {
      int lg = 0;
      while ((1ull << lg) < nf) ++lg;
      shift1 = shiftF + (lg + 1) / 2;
      while ((((rcodes - 1) >> shift1) + 1) > SP_MAXBINS) ++shift1;
      if (shift1 - shiftF > 10) tile = false;
      nb1 = (uint32_t)(((rcodes - 1) >> shift1) + 1);
    }
  }
  // records carry min(real length, 63) in their top 6 bits when the code leaves room for it
  const int recL = (c->numcodes <= (1ull << 26) && c->cp.pl + Text<BITS>::W <= 63) ? 1 : 0;
  const bool keyed = tile && keyed_ok && recL;
  uint64_t *medkeys = nullptr;               // byte text, keyed: keys of the buckets the tile sorter only inserts
  if (tile) {
    c->diag[0] = 1; c->diag[1] = nb1 ? 2 : 1; c->diag[2] = keyed ? 1 : 0;
    const uint64_t binwords = (uint64_t)nbinsF + 3 * SP_MAXBINS + 64;
    if (binwords > c->cap_bins) {
      dfree(c, c->d_bins);
      c->d_bins = nullptr;
      c->cap_bins = 0;
      CK(dmalloc(c, &c->d_bins, binwords * 4));
      c->cap_bins = binwords;
    }
    // The records the tile sorter reads always end up in recB, so that the 64-bit buffer is free for the .suf image
    // the sorters write themselves (with two levels the first one passes through the image buffer, dead by then).
    const uint64_t recB_bytes = (c->part_width + 1) * (keyed ? 16 : 8);
    if (recB_bytes > c->cap_recB) {
      dfree(c, c->d_recB);
      c->d_recB = nullptr;
      c->cap_recB = 0;
      CK(dmalloc(c, &c->d_recB, recB_bytes));
      c->cap_recB = recB_bytes;
    }
    uint32_t *gcurF = c->d_bins, *gcur1 = gcurF + nbinsF, *start1 = gcur1 + SP_MAXBINS, *blkfirst = start1 + SP_MAXBINS;
    // blocks of the text-level partition scan reps x chunk positions when the range covers only 1/reps of the suffixes
    int reps = 1;
    if (c->nsort > 0 && rwidth > 0) {
      const uint64_t ratio = c->nsort / rwidth;
      reps = ratio >= 8 ? 8 : ratio >= 2 ? (int)ratio : 1;
    }
    if (keyed) {
      // ---- keyed records (keyed.cuh): 16 bytes per suffix, the key travels with the record ----
      CK(cudaFuncSetAttribute(k_keyed_partition_dense, cudaFuncAttributeMaxDynamicSharedMemorySize, KP_SMEM));
      CK(cudaFuncSetAttribute(k_keyed_partition_sparse, cudaFuncAttributeMaxDynamicSharedMemorySize, KP_SMEM));
      CK(cudaFuncSetAttribute(k_keyed_partition_bytes, cudaFuncAttributeMaxDynamicSharedMemorySize, KP_SMEM));
      CK(cudaFuncSetAttribute(k_keyed_partition_rec, cudaFuncAttributeMaxDynamicSharedMemorySize, KP_SMEM));
      const unsigned textblocks = grid_for(c->n, (uint64_t)KP_CHUNK * reps);
      const int whole = (lo == 0 && hi == c->numcodes - 1) ? 1 : 0;
      ulonglong2 *recX = (ulonglong2 *)c->d_suf64, *recY = (ulonglong2 *)c->d_recB;
      const Text<2> T2 = text_of<2>(c);
      const Text<8> T8 = text_of<8>(c);
      if (BITS == 8) medkeys = c->d_suf64 + (c->part_width + 1);      // second half of the 16-byte-per-suffix buffer
      {
        ProfScope ps(c, KC_SCATTER, nb1 ? 5 : 2);
        k_bins_init<<<grid_for(nbinsF, 256), 256, 0, c->stream>>>(c->d_lb, lo, shiftF, nbinsF, out.partoffset, gcurF,
                                                                  nullptr);
        uint32_t *gc = nb1 ? gcur1 : gcurF;
        const int sh = nb1 ? shift1 : shiftF;
        const uint32_t nb = nb1 ? nb1 : nbinsF;
        ulonglong2 *dst = nb1 ? recX : recY;
        if (nb1) k_bins_init<<<grid_for(nb1, 256), 256, 0, c->stream>>>(c->d_lb, lo, shift1, nb1, out.partoffset, gcur1, start1);
        if (BITS == 8)
          k_keyed_partition_bytes<<<textblocks, 256, KP_SMEM, c->stream>>>(T8, c->cp, c->n, lo, hi, sh, nb, reps, gc, dst);
        else if (reps > 1)
          k_keyed_partition_sparse<<<textblocks, 256, KP_SMEM, c->stream>>>(T2, c->cp.pl, c->n, lo, hi, sh, nb, reps, gc, dst);
        else
          k_keyed_partition_dense<<<textblocks, 256, KP_SMEM, c->stream>>>(T2, c->cp.pl, c->n, lo, hi, whole, sh, nb, gc, dst);
        if (nb1) {
          k_keyed_chunk_table<<<1, 256, 0, c->stream>>>(start1, gcur1, nb1, blkfirst);
          const unsigned recblocks = (unsigned)(rwidth / (KP_CHUNK * KP_SUPER) + nb1 + 1);
          k_keyed_partition_rec<<<recblocks, 512, KP_SMEM, c->stream>>>(lo, shiftF, 1u << (shift1 - shiftF), recX, start1, gcur1,
                                                                        blkfirst, nb1, shift1, gcurF, recY);
        }
      }
      {
        ProfScope ps(c, KC_SORT);
        const unsigned tb = nbinsF < 148u * 2 ? nbinsF : 148u * 2;
        CK(cudaFuncSetAttribute(k_tile_sort_keyed<BITS>, cudaFuncAttributeMaxDynamicSharedMemorySize, TS_SMEM));
        PartOut tfin = out;
        if (c->fused_widen) tfin.sa64 = c->d_suf64;          // the level-1 records in it are dead by now
        if (BITS == 8 && c->subsort && (((uint64_t)c->cp.sigma + 1) << shiftF) <= TSUB_MAX) {
          // byte text: buckets split once more by the first key symbol inside the tile, warp networks on the groups
          CK(cudaFuncSetAttribute(k_tile_subsort, cudaFuncAttributeMaxDynamicSharedMemorySize, TSUB_SMEM));
          k_tile_subsort<<<tb, TS_THREADS, TSUB_SMEM, c->stream>>>(T8, tfin, c->d_lb, c->d_cursor, recY, gcurF, lo, hi, shiftF,
                                                                   nbinsF, c->cp.pl, c->cp.sigma, fbb, fbwords, medkeys);
        } else
        k_tile_sort_keyed<BITS><<<tb, TS_THREADS, TS_SMEM, c->stream>>>(T, tfin, c->d_lb, c->d_cursor, recY, gcurF, lo, hi,
                                                                        shiftF, nbinsF, c->cp.pl, fbb, fbwords, medkeys);
      }
      CK(cudaGetLastError());
    } else {
    CK(cudaFuncSetAttribute(k_stage_partition<BITS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SP_SMEM));
    CK(cudaFuncSetAttribute(k_stage_partition<BITS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SP_SMEM));
    const unsigned textblocks = grid_for(c->n, (uint64_t)SP_CHUNK * reps);
    if (reps > 1)
      CK(cudaFuncSetAttribute(k_stage_partition_sparse<BITS>, cudaFuncAttributeMaxDynamicSharedMemorySize, SP_SMEM_SPARSE));
    {
      ProfScope ps(c, KC_SCATTER, nb1 ? 5 : 2);
      k_bins_init<<<grid_for(nbinsF, 256), 256, 0, c->stream>>>(c->d_lb, lo, shiftF, nbinsF, out.partoffset, gcurF,
                                                                nullptr);
      if (!nb1) {
        if (reps > 1)
          k_stage_partition_sparse<BITS><<<textblocks, 256, SP_SMEM_SPARSE, c->stream>>>(T, c->cp, c->n, lo, hi, shiftF, nbinsF,
                                                                                         reps, gcurF, c->d_recB, recL);
        else
          k_stage_partition<BITS, true><<<textblocks, 256, SP_SMEM, c->stream>>>(T, c->cp, c->n, lo, hi, shiftF, nbinsF,
                                                                                 nullptr, nullptr, nullptr, nullptr, 0, 0, gcurF,
                                                                                 c->d_recB, recL);
      } else {
        k_bins_init<<<grid_for(nb1, 256), 256, 0, c->stream>>>(c->d_lb, lo, shift1, nb1, out.partoffset, gcur1, start1);
        if (reps > 1)
          k_stage_partition_sparse<BITS><<<textblocks, 256, SP_SMEM_SPARSE, c->stream>>>(T, c->cp, c->n, lo, hi, shift1, nb1,
                                                                                         reps, gcur1, c->d_suf64, recL);
        else
          k_stage_partition<BITS, true><<<textblocks, 256, SP_SMEM, c->stream>>>(T, c->cp, c->n, lo, hi, shift1, nb1,
                                                                                 nullptr, nullptr, nullptr, nullptr, 0, 0, gcur1,
                                                                                 c->d_suf64, recL);
        k_chunk_table<<<1, 256, 0, c->stream>>>(start1, gcur1, nb1, blkfirst);
        const unsigned recblocks = (unsigned)(rwidth / SP_CHUNK + nb1 + 1);
        k_stage_partition<BITS, false><<<recblocks, 256, SP_SMEM, c->stream>>>(T, c->cp, c->n, lo, hi, shiftF,
                                                                               1u << (shift1 - shiftF), c->d_suf64, start1, gcur1,
                                                                               blkfirst, nb1, shift1, gcurF, c->d_recB, recL);
      }
    }
    {
      ProfScope ps(c, KC_SORT);
      const unsigned tb = nbinsF < 148u * 2 ? nbinsF : 148u * 2;     // two 512-thread blocks per SM, tiles dealt round-robin
      CK(cudaFuncSetAttribute(k_tile_sort<BITS>, cudaFuncAttributeMaxDynamicSharedMemorySize, TS_SMEM));
      PartOut tfin = out;
      if (c->fused_widen) tfin.sa64 = c->d_suf64;          // records are in recB: the image buffer is free
      k_tile_sort<BITS><<<tb, TS_THREADS, TS_SMEM, c->stream>>>(T, tfin, c->d_lb, c->d_cursor, c->d_recB, gcurF, lo, hi, shiftF,
                                                   nbinsF, c->cp.pl, recL, fbb, fbwords);
    }
    CK(cudaGetLastError());
    }
  }
  const bool twolevel = !tile && rwidth * 4 > c->twolevel_bytes && rcodes >= 4096;
  // Direct insertion and the tile path leave the 64-bit buffer free (the old two-level insertion keeps its records in
  // it): the sorters then write the .suf image themselves and k_widen is not launched.  A later refinement redoes it.
  PartOut fin = out;
  if (!twolevel && c->fused_widen) fin.sa64 = c->d_suf64;
  if (!fin.sa64) c->part_imaged = false;
  // special-padded suffixes go to the right end of their buckets: slots no other kernel of the range writes, so
  // the placement runs beside the insertion (joined before the bucket borders are computed)
  bool padded_beside = false;
  if (c->padcap > 1) {
    padded_beside = !tile && side_usable(c);
    if (padded_beside && side_fork(c)) return -1;
    ProfScope ps(c, KC_PADDED);
    k_place_padded<<<grid_for(c->padcap, 256), 256, 0, padded_beside ? c->side_stream : c->stream>>>(
        c->d_padkey, c->d_padpos, c->d_persist, c->d_lb, c->d_csc, c->cp.sigma, c->cp.pl, lo, hi, fin);
  }
  if (tile) {
    // insertion already done
  } else if (!twolevel) {
    ProfScope ps(c, KC_SCATTER, 1);
    k_kmers<BITS, 1><<<grid_for((c->n + KM_PPT - 1) / KM_PPT, 256), 256, 0, c->stream>>>(T, c->cp, 0, c->n, c->d_cursor, c->d_sa, lo,
                                                                            hi, out.partoffset);
  } else {
    int shift = 0;
    while (((rcodes - 1) >> shift) + 1 > PT_MAXBINS) ++shift;
    const uint32_t nbins = (uint32_t)(((rcodes - 1) >> shift) + 1);
    if (ensure_tilesums(c, PT_MAXBINS)) return -1;          // reused as the bins' record cursors
    uint32_t *gcursor = c->d_tilesums;
    ProfScope ps(c, KC_SCATTER, 3);
    k_partition_init<<<grid_for(nbins, 256), 256, 0, c->stream>>>(c->d_lb, lo, shift, nbins, out.partoffset, gcursor);
    k_partition<BITS><<<grid_for((c->n + 31) / 32, 256), 256, 0, c->stream>>>(T, c->cp, c->n, lo, hi, shift, nbins,
                                                                             gcursor, c->d_suf64);
    const size_t smem = ((size_t)4 << shift);
    if (smem <= 160 * 1024) {
      if (smem > 48 * 1024)
        CK(cudaFuncSetAttribute(k_scatter_bins, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      k_scatter_bins<<<nbins, 1024, smem, c->stream>>>(c->d_suf64, gcursor, c->d_lb, c->d_cursor, c->d_sa, lo, hi,
                                                       shift, out.partoffset);
    } else {
      k_partition_gaps<<<grid_for((uint64_t)nbins * 32, 256), 256, 0, c->stream>>>(gcursor, c->d_lb, lo, hi, shift,
                                                                                  nbins, out.partoffset, c->d_suf64);
      k_scatter_records<<<grid_for(rwidth, 256), 256, 0, c->stream>>>(c->d_suf64 + roff, rwidth, c->d_cursor, c->d_sa,
                                                                     out.partoffset);
    }
  }
  {
    ProfScope ps(c, KC_SORT);
    uint64_t warps = rcodes < 148ull * 64 ? rcodes : 148ull * 64;   // 8 blocks of 8 warps per SM
    const uint32_t medmax = c->sort_maxbucket == SORT_MAXBUCKET ? MED_MAX : 0;   // testing knob disables both
    k_sort_buckets<BITS><<<grid_for(warps * 32, 256), 256, 0, c->stream>>>(T, fin, c->d_lb, c->d_cursor, lo, hi,
                                                                          c->cp.pl, fbb, c->sort_maxbucket,
                                                                          c->d_medlist, medmax, tile ? 1 : 0, fbwords);
  }
  if (c->sort_maxbucket == SORT_MAXBUCKET && rwidth > SORT_MAXBUCKET) {
    ProfScope ps(c, KC_SORTMED);
    const size_t smem = MED_SMEM;
    CK(cudaFuncSetAttribute(k_sort_medium<BITS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_sort_medium<BITS><<<148 * 8, 256, smem, c->stream>>>(T, fin, c->d_lb, c->d_cursor, c->d_medlist, lo, c->cp.pl,
                                                          fbb, fbwords, medkeys);
  }
  if (padded_beside && side_join(c)) return -1;
  {
    ProfScope ps(c, KC_BOUNDARY, first_of_part && out.partoffset > 0 ? 2 : 1);
    k_boundary<BITS><<<grid_for(rcodes, 256), 256, 0, c->stream>>>(T, c->cp, out, c->d_lb, c->d_csc, lo, hi);
    if (first_of_part && out.partoffset > 0)
      k_part_first<BITS><<<1, 32, 0, c->stream>>>(T, c->cp, out, c->d_lb, c->numcodes, c->d_csc, c->d_padkey, c->d_persist);
  }
  if (!fin.sa64) {
    ProfScope ps(c, KC_WIDEN);
    k_widen<<<grid_for(rwidth, 256), 256, 0, c->stream>>>(c->d_sa + roff, rwidth, c->d_suf64 + roff, (uint64_t)out.partoffset + roff,
                                                         c->d_stats);
  }
  CK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Narrow transfer of a part's suffix array: the 32-bit working array crosses PCIe in chunks through a small ring of
// page-locked buffers, host threads widen every chunk into the caller's GtUword array while the next chunks are in
// flight (csrc/host_logic.c, streaming stores).  Half the bus traffic of copying the 64-bit image, and the caller's
// buffer may be ordinary pageable memory.  The ring is cached per process.
// ---------------------------------------------------------------------------------------------
extern "C" void gtb200_widen_u32_u64(const uint32_t *src, uint64_t *dst, size_t n);


static bool ring_acquire(NarrowRing *r) {
  {
    std::lock_guard<std::mutex> lock(g_pool_mu);
    if (!g_ring_pool.empty()) { *r = g_ring_pool.back(); g_ring_pool.pop_back(); return true; }
  }
  for (int i = 0; i < NR_RING; ++i) r->buf[i] = nullptr;
  for (int i = 0; i < NR_RING; ++i)
    if (cudaHostAlloc((void **)&r->buf[i], (size_t)NR_CHUNK * 4, cudaHostAllocPortable) != cudaSuccess) {
      (void)cudaGetLastError();
      for (int j = 0; j < i; ++j) cudaFreeHost(r->buf[j]);
      return false;
    }
  return true;
}
static void ring_release(const NarrowRing &r) {
  std::lock_guard<std::mutex> lock(g_pool_mu);
  g_ring_pool.push_back(r);
}

// Persistent host workers for the widening (spawning threads per 64 MB chunk cost a quarter of the widening time).
// One pool per process; a job is split into equal 64-entry-aligned shares, callers are serialised by the job mutex.
#include <condition_variable>
struct WidenPool {
  std::mutex job_mu;                    // one widening job at a time
  std::mutex mu;
  std::condition_variable cv_work, cv_done;
  std::vector<std::thread> workers;
  const uint32_t *src = nullptr;
  uint64_t *dst = nullptr;
  size_t n = 0, per = 0;
  int shares = 0, next_share = 0, done = 0;
  uint64_t generation = 0;
  void worker() {
    uint64_t seen = 0;
    for (;;) {
      std::unique_lock<std::mutex> lk(mu);
      cv_work.wait(lk, [&] { return generation != seen && next_share < shares; });
      while (next_share < shares) {
        const int sh = next_share++;
        const uint64_t gen = generation;
        lk.unlock();
        const size_t lo = (size_t)sh * per;
        gtb200_widen_u32_u64(src + lo, dst + lo, lo + per <= n ? per : n - lo);
        lk.lock();
        if (gen == generation && ++done == shares) cv_done.notify_all();
      }
      seen = generation;
    }
  }
  void run(const uint32_t *s, uint64_t *d, size_t count, int threads) {
    std::lock_guard<std::mutex> job(job_mu);
    {
      std::unique_lock<std::mutex> lk(mu);
      while ((int)workers.size() < threads - 1) {
        workers.emplace_back([this] { worker(); });
        workers.back().detach();
      }
      src = s; dst = d; n = count;
      per = ((count + threads - 1) / threads + 63) & ~(size_t)63;
      shares = (int)((count + per - 1) / per);
      next_share = 0; done = 0;
      ++generation;
    }
    cv_work.notify_all();
    std::unique_lock<std::mutex> lk(mu);
    while (next_share < shares) {                     // the calling thread takes shares too
      const int sh = next_share++;
      lk.unlock();
      const size_t lo = (size_t)sh * per;
      gtb200_widen_u32_u64(src + lo, dst + lo, lo + per <= n ? per : n - lo);
      lk.lock();
      if (++done == shares) cv_done.notify_all();
    }
    cv_done.wait(lk, [&] { return done == shares; });
  }
};
static WidenPool *g_widen_pool = new WidenPool();     // never destroyed: its detached workers outlive static teardown

static void widen_parallel(const uint32_t *src, uint64_t *dst, size_t n, int threads) {
  if (threads <= 1 || n < (1u << 16)) { gtb200_widen_u32_u64(src, dst, n); return; }
  g_widen_pool->run(src, dst, n, threads);
}

struct NarrowJob {
  gtb200_ctx *c;
  uint64_t *h_suf;
  uint8_t *h_lcp;
  int S;
  uint64_t boff[GTB_MAXSLICES + 1];
  bool wait_events;            // slices become ready one by one (slice_ev[k] on the compute stream)
  cudaStream_t stream;         // the stream the copies run on
  int rc;
  std::string err;
};

static void narrow_copy_run(NarrowJob *j) {
  gtb200_ctx *c = j->c;
  j->rc = -1;
  if (cudaSetDevice(c->device) != cudaSuccess) { j->err = "cudaSetDevice failed in the copy thread"; return; }
  NarrowRing ring;
  if (!ring_acquire(&ring)) { j->err = "cudaHostAlloc of the transfer ring failed"; return; }
  cudaEvent_t rev[NR_RING];
  for (int i = 0; i < NR_RING; ++i) cudaEventCreateWithFlags(&rev[i], cudaEventDisableTiming);
  struct Chunk { int slice; uint64_t off; uint32_t len; bool first, last; };
  std::vector<Chunk> chunks;
  for (int k = 0; k < j->S; ++k) {
    const uint64_t lo = j->boff[k], hi = j->boff[k + 1];
    for (uint64_t o = lo; o < hi; o += NR_CHUNK) {
      const uint32_t len = (uint32_t)(hi - o < NR_CHUNK ? hi - o : NR_CHUNK);
      chunks.push_back({k, o, len, o == lo, o + len == hi});
    }
  }
  cudaError_t e = cudaSuccess;
  auto issue = [&](size_t i) {
    const Chunk &ch = chunks[i];
    if (ch.first && j->wait_events && e == cudaSuccess) e = cudaStreamWaitEvent(j->stream, c->slice_ev[ch.slice], 0);
    if (e == cudaSuccess) e = cudaMemcpyAsync(ring.buf[i % NR_RING], c->d_sa + ch.off, (size_t)ch.len * 4, cudaMemcpyDeviceToHost, j->stream);
    if (e == cudaSuccess) e = cudaEventRecord(rev[i % NR_RING], j->stream);
    if (ch.last && j->h_lcp && e == cudaSuccess) {
      const uint64_t lo = j->boff[ch.slice], hi = j->boff[ch.slice + 1];
      e = cudaMemcpyAsync(j->h_lcp + lo, c->d_lcp + lo, hi - lo, cudaMemcpyDeviceToHost, j->stream);
    }
  };
  for (size_t i = 0; i < chunks.size() && i < NR_RING; ++i) issue(i);
  for (size_t i = 0; i < chunks.size() && e == cudaSuccess; ++i) {
    e = cudaEventSynchronize(rev[i % NR_RING]);
    if (e != cudaSuccess) break;
    widen_parallel(ring.buf[i % NR_RING], j->h_suf + chunks[i].off, chunks[i].len, c->widen_threads);
    if (i + NR_RING < chunks.size()) issue(i + NR_RING);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(j->stream);
  for (int i = 0; i < NR_RING; ++i) cudaEventDestroy(rev[i]);
  ring_release(ring);
  if (e != cudaSuccess) { j->err = std::string("narrow transfer: ") + cudaGetErrorString(e); return; }
  j->rc = 0;
}

template <int BITS>
static int sort_part_impl(gtb200_ctx *c, uint64_t mincode, uint64_t maxcode, gtb200_partinfo *info, void *h_suf,
                          int h_suf_u32, uint8_t *h_lcp) {
  const Text<BITS> T = text_of<BITS>(c);
  // part geometry from the bucket table (known without a round trip when the part is everything)
  uint64_t partoffset = 0, width = c->nsort;
  if (!(mincode == 0 && maxcode == c->numcodes - 1)) {
    uint32_t h_lb[2];
    CK(cudaMemcpyAsync(&h_lb[0], c->d_lb + mincode, 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(&h_lb[1], c->d_lb + maxcode + 1, 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    partoffset = h_lb[0];
    width = (uint64_t)h_lb[1] - h_lb[0];
  }
  const uint64_t ncodes = maxcode - mincode + 1;
  const uint64_t fbwords = (ncodes + 31) / 32;
  memset(info, 0, sizeof *info);
  info->partoffset = partoffset;
  info->width = width;
  info->longest = UINT64_MAX;
  c->part_offset = partoffset;
  c->part_width = width;
  c->part_nllv = 0;
  if (width == 0) return 0;
  // keyed records once the packed text (and its bitmap) no longer fit L2 beside the streams passing through:
  // the tile sorter's one random text access per suffix is then a 64-byte DRAM access.
  // Measured at 250 Mbp (94 MB of text + bitmap): 7.15 ms keyed against 8.17 ms with 8-byte records.
  const bool keyed_ok = ((BITS == 2 && c->cp.pow2 && c->cp.pl <= 15) || (BITS == 8 && c->cp.pl <= 8)) &&
                        c->numcodes <= (1ull << 26) && c->keyed_mode != 0 &&
                        (c->keyed_mode == 2 || (c->ntextwords + c->nmaskwords) * 8 > (64ull << 20));
  memset(c->diag, 0, sizeof c->diag);
  if (ensure_part_buffers(c, width, ncodes, keyed_ok)) return -1;
  PartOut out;
  out.sa = c->d_sa;
  out.sa64 = nullptr;
  out.lcp = c->d_lcp;
  out.llv = nullptr;
  out.stats = c->d_stats;
  out.partoffset = (uint32_t)partoffset;
  // With host buffers to fill, the part is cut into up to 8 code ranges: while range k+1 is inserted and sorted
  // on the compute stream, the finished slice of range k already crosses PCIe on the copy stream.
  const bool to_host = h_suf != nullptr || h_lcp != nullptr;
  int S = 1;
  if (to_host && c->pipeline && width >= (2u << 20) && ncodes >= 4096) {
    S = width >= (64u << 20) ? GTB_MAXSLICES : width >= (16u << 20) ? 4 : 2;
    // every slice of a tile-path part scans the whole text for its code range: a part that is itself a small share of
    // the suffix array (one of many parts, one GPU's range of a part) is cut into fewer slices or none
    const uint64_t fair = c->nsort ? width * 16 / c->nsort : 16;
    if ((uint64_t)S > fair) S = fair < 2 ? 1 : (int)fair;
  }
  uint64_t blo[GTB_MAXSLICES + 1], boff[GTB_MAXSLICES + 1];
  blo[0] = mincode; boff[0] = 0;
  blo[S] = maxcode + 1; boff[S] = width;
  if (S > 1) {
    uint32_t *hb = (uint32_t *)((char *)c->h_stats + 160);
    for (int k = 1; k < S; ++k) {
      blo[k] = mincode + ((ncodes * (uint64_t)k / S) & ~31ull);
      CK(cudaMemcpyAsync(hb + k, c->d_lb + blo[k], 4, cudaMemcpyDeviceToHost, c->stream));
    }
    CK(cudaStreamSynchronize(c->stream));
    for (int k = 1; k < S; ++k) boff[k] = (uint64_t)hb[k] - partoffset;
    if (!c->copy_stream) {
      CK(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
      for (int k = 0; k < GTB_MAXSLICES; ++k) CK(cudaEventCreateWithFlags(&c->slice_ev[k], cudaEventDisableTiming));
    }
  }
  {
    ProfScope ps(c, KC_INIT);
    k_part_init<<<grid_for(ncodes, 256), 256, 0, c->stream>>>(c->d_lb, c->d_cursor, mincode, ncodes, c->d_fbbits, 2 * fbwords,
                                                              c->d_stats, c->d_lcp);
  }
  c->part_imaged = true;
  for (int k = 0; k < S; ++k) {
    if (enqueue_range<BITS>(c, T, out, mincode, blo[k], blo[k + 1] - 1, boff[k], boff[k + 1] - boff[k], fbwords, k == 0,
                            keyed_ok))
      return -1;
    if (S > 1) CK(cudaEventRecord(c->slice_ev[k], c->stream));
  }
  // a GtUword suftab of a large part: 32-bit positions over the bus, widened on the host (see narrow_copy_run)
  // (c->narrow == 2 forces it for tests; small parts are latency-, not PCIe-bound: 4.6 Mbp 1.1 ms direct, 2.1 ms narrow)
  const bool narrow = h_suf != nullptr && !h_suf_u32 && c->narrow && (c->narrow == 2 ? width >= (4u << 20) : width >= (32u << 20));
  NarrowJob job;
  std::thread copier;
  if (S > 1 && narrow) {
    job.c = c; job.h_suf = (uint64_t *)h_suf; job.h_lcp = h_lcp; job.S = S; job.wait_events = true;
    job.stream = c->copy_stream;
    for (int k = 0; k <= S; ++k) job.boff[k] = boff[k];
    copier = std::thread(narrow_copy_run, &job);
  } else if (S > 1) {
    for (int k = 0; k < S; ++k) {
      const uint64_t o = boff[k], w = boff[k + 1] - boff[k];
      if (w == 0) continue;
      CK(cudaStreamWaitEvent(c->copy_stream, c->slice_ev[k], 0));
      if (h_suf) {
        if (h_suf_u32) CK(cudaMemcpyAsync((uint32_t *)h_suf + o, c->d_sa + o, w * 4, cudaMemcpyDeviceToHost, c->copy_stream));
        else CK(cudaMemcpyAsync((uint64_t *)h_suf + o, c->d_suf64 + o, w * 8, cudaMemcpyDeviceToHost, c->copy_stream));
      }
      if (h_lcp) CK(cudaMemcpyAsync(h_lcp + o, c->d_lcp + o, w, cudaMemcpyDeviceToHost, c->copy_stream));
    }
  }
  const int rc_stats = sync_stats(c);
  if (copier.joinable()) {
    copier.join();
    if (job.rc) return set_error("%s", job.err.c_str());
  }
  if (rc_stats) return -1;
  bool copied = S > 1;
  if (S > 1) CK(cudaStreamSynchronize(c->copy_stream));
  if (c->h_stats->fb_elems != 0) {
    // some buckets need refinement: finish them, then redo borders and the image on the final order
    if (refine_impl<BITS>(c, out, mincode, maxcode, width, info)) return -1;
    {
      ProfScope ps(c, KC_BOUNDARY, partoffset > 0 ? 3 : 2);
      k_reset_boundary_stats<<<1, 1, 0, c->stream>>>(c->d_stats);
      k_boundary<BITS><<<grid_for(ncodes, 256), 256, 0, c->stream>>>(T, c->cp, out, c->d_lb, c->d_csc, mincode, maxcode);
      if (partoffset > 0)
        k_part_first<BITS><<<1, 32, 0, c->stream>>>(T, c->cp, out, c->d_lb, c->numcodes, c->d_csc, c->d_padkey, c->d_persist);
    }
    {
      ProfScope ps(c, KC_WIDEN);
      if (c->part_imaged)
        k_widen_flagged<<<(unsigned)fbwords, 256, 0, c->stream>>>(c->d_sa, c->d_suf64, c->d_lb, c->d_fbbits, c->d_fbbits + fbwords,
                                                                 mincode, maxcode, (uint32_t)partoffset, c->d_stats);
      else
        k_widen<<<grid_for(width, 256), 256, 0, c->stream>>>(c->d_sa, width, c->d_suf64, partoffset, c->d_stats);
    }
    CK(cudaGetLastError());
    if (sync_stats(c)) return -1;
    copied = false;
  }
  if (to_host && !copied && narrow) {
    job.c = c; job.h_suf = (uint64_t *)h_suf; job.h_lcp = h_lcp; job.S = 1; job.wait_events = false;
    job.stream = c->stream;
    job.boff[0] = 0; job.boff[1] = width;
    narrow_copy_run(&job);
    if (job.rc) return set_error("%s", job.err.c_str());
  } else if (to_host && !copied) {
    if (h_suf) {
      if (h_suf_u32) CK(cudaMemcpyAsync(h_suf, c->d_sa, width * 4, cudaMemcpyDeviceToHost, c->stream));
      else CK(cudaMemcpyAsync(h_suf, c->d_suf64, width * 8, cudaMemcpyDeviceToHost, c->stream));
    }
    if (h_lcp) CK(cudaMemcpyAsync(h_lcp, c->d_lcp, width, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
  }
  c->diag[3] = c->h_stats->med_elems;
  c->diag[4] = c->padcap ? c->padcap - 1 : 0;     // upper bound: (special ranges + 1) x (pl - 1)
  info->numoflargelcpvalues = c->h_stats->llv_count;
  info->maxbranchdepth = c->h_stats->maxdepth > c->h_stats->bmax ? c->h_stats->maxdepth : c->h_stats->bmax;
  info->lcptabsum = (double)(c->h_stats->lcpsum + c->h_stats->bsum);
  info->longest = c->h_stats->longest;
  c->part_nllv = c->h_stats->llv_count;
  return 0;
}

extern "C" int gtb200_sort_part_device(gtb200_ctx *c, uint64_t mincode, uint64_t maxcode, gtb200_partinfo *info) {
  if (!c || !info) return set_error("gtb200_sort_part_device: null argument");
  if (!c->counted) return set_error("gtb200_sort_part_device: call gtb200_count first");
  if (mincode > maxcode || maxcode >= c->numcodes)
    return set_error("gtb200_sort_part_device: code range [%llu,%llu] invalid (numofallcodes %llu)",
                     (unsigned long long)mincode, (unsigned long long)maxcode, (unsigned long long)c->numcodes);
  CK(cudaSetDevice(c->device));
  return c->bits == 2 ? sort_part_impl<2>(c, mincode, maxcode, info, nullptr, 0, nullptr)
                      : sort_part_impl<8>(c, mincode, maxcode, info, nullptr, 0, nullptr);
}

extern "C" int gtb200_part_device(gtb200_ctx *c, void **suftab_dev, void **lcptab_dev, void **llv_dev) {
  if (!c) return set_error("gtb200_part_device: null context");
  if (suftab_dev) *suftab_dev = c->d_suf64;
  if (lcptab_dev) *lcptab_dev = c->d_lcp;
  if (llv_dev) *llv_dev = c->d_llv;
  return 0;
}

extern "C" int gtb200_fetch_part(gtb200_ctx *c, void *suftab, int suftab_is_uint32, uint8_t *lcptab,
                                 gtb200_largelcp *llv, uint64_t llv_capacity) {
  if (!c) return set_error("gtb200_fetch_part: null context");
  CK(cudaSetDevice(c->device));
  const uint64_t w = c->part_width;
  if (w == 0) return 0;
  if (suftab) {
    if (suftab_is_uint32) CK(cudaMemcpyAsync(suftab, c->d_sa, w * 4, cudaMemcpyDeviceToHost, c->stream));
    else CK(cudaMemcpyAsync(suftab, c->d_suf64, w * 8, cudaMemcpyDeviceToHost, c->stream));
  }
  if (lcptab) CK(cudaMemcpyAsync(lcptab, c->d_lcp, w, cudaMemcpyDeviceToHost, c->stream));
  if (llv && c->part_nllv) {
    if (c->part_nllv > llv_capacity)
      return set_error("gtb200_fetch_part: %llu large lcp values, capacity %llu", (unsigned long long)c->part_nllv,
                       (unsigned long long)llv_capacity);
    CK(cudaMemcpyAsync(llv, c->d_llv, c->part_nllv * sizeof(gtb200_largelcp), cudaMemcpyDeviceToHost, c->stream));
  }
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" int gtb200_sort_part(gtb200_ctx *c, uint64_t mincode, uint64_t maxcode, void *suftab, int suftab_is_uint32,
                                uint8_t *lcptab, gtb200_largelcp *llv, uint64_t llv_capacity, gtb200_partinfo *info) {
  if (!c || !info) return set_error("gtb200_sort_part: null argument");
  if (!c->counted) return set_error("gtb200_sort_part: call gtb200_count first");
  if (mincode > maxcode || maxcode >= c->numcodes)
    return set_error("gtb200_sort_part: code range [%llu,%llu] invalid (numofallcodes %llu)", (unsigned long long)mincode,
                     (unsigned long long)maxcode, (unsigned long long)c->numcodes);
  CK(cudaSetDevice(c->device));
  const int rc = c->bits == 2 ? sort_part_impl<2>(c, mincode, maxcode, info, suftab, suftab_is_uint32, lcptab)
                              : sort_part_impl<8>(c, mincode, maxcode, info, suftab, suftab_is_uint32, lcptab);
  if (rc) return rc;
  return gtb200_fetch_part(c, nullptr, 0, nullptr, llv, llv_capacity);   // suftab and lcptab are already on the host
}

extern "C" int gtb200_tail(gtb200_ctx *c, uint64_t first, uint64_t count, void *suftab, int suftab_is_uint32) {
  if (!c) return set_error("gtb200_tail: null context");
  if (first + count > c->specialchars + 1)
    return set_error("gtb200_tail: [%llu,%llu) exceeds the %llu tail entries", (unsigned long long)first,
                     (unsigned long long)(first + count), (unsigned long long)(c->specialchars + 1));
  if (count == 0) return 0;
  CK(cudaSetDevice(c->device));
  const size_t esz = suftab_is_uint32 ? 4 : 8;
  if (count * esz > c->cap_tail) {
    dfree(c, c->d_tail);
    c->d_tail = nullptr;
    c->cap_tail = 0;
    CK(dmalloc(c, &c->d_tail, count * esz));
    c->cap_tail = count * esz;
  }
  {
    ProfScope ps(c, KC_TAIL);
    if (suftab_is_uint32)
      k_tail<uint32_t><<<grid_for(count, 256), 256, 0, c->stream>>>(c->d_ranges, c->d_rangeoff, c->nranges, c->specialchars,
                                                                   c->n, first, count, (uint32_t *)c->d_tail);
    else
      k_tail<uint64_t><<<grid_for(count, 256), 256, 0, c->stream>>>(c->d_ranges, c->d_rangeoff, c->nranges, c->specialchars,
                                                                   c->n, first, count, (uint64_t *)c->d_tail);
  }
  CK(cudaGetLastError());
  if (suftab) {                               // without a host buffer the call stays stream-ordered (device consumers)
    CK(cudaMemcpyAsync(suftab, c->d_tail, count * esz, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
  }
  return 0;
}

extern "C" int gtb200_tail_device(gtb200_ctx *c, void **tail_dev) {
  if (!c || !tail_dev) return set_error("gtb200_tail_device: null argument");
  *tail_dev = c->d_tail;
  return 0;
}

extern "C" int gtb200_part_diag(gtb200_ctx *c, uint64_t *values, int capacity) {
  if (!c || !values) return set_error("gtb200_part_diag: null argument");
  for (int i = 0; i < 8 && i < capacity; ++i) values[i] = c->diag[i];
  return 8;
}

extern "C" int gtb200_profile_enable(gtb200_ctx *c, int on) {
  if (!c) return set_error("null context");
  if (!on) prof_fold(c);
  c->prof_on = on != 0;
  return 0;
}

extern "C" void gtb200_profile_reset(gtb200_ctx *c) {
  if (!c) return;
  prof_fold(c);
  for (int i = 0; i < KC_NUM; ++i) { c->prof_ms[i] = 0.0; c->prof_launches[i] = 0; }
}

extern "C" int gtb200_profile_read(gtb200_ctx *c, const char **names, double *ms, uint64_t *launches, int capacity) {
  if (!c) return set_error("null context");
  if (prof_fold(c)) return -1;
  for (int i = 0; i < KC_NUM && i < capacity; ++i) {
    if (names) names[i] = kclass_names[i];
    if (ms) ms[i] = c->prof_ms[i];
    if (launches) launches[i] = c->prof_launches[i];
  }
  return KC_NUM;
}
#include "multi.cuh"
#include "fasta.cuh"

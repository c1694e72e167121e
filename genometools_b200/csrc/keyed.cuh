// Keyed tile path (DNA whose packed text does not fit L2): the records that travel through the two partition
// levels carry the sort key of their suffix, so the tile sorter never touches the text.
//
//   record = 16 bytes {a, k}
//     a = (min(real symbols ahead, 63) << 58) | (code << 32) | position        (code < 2^26)
//     k = make_key_known(T, position + pl, real - pl): the 28 bases that follow the prefix, padded with ones
//         from the first special on, low byte = 28 - (number of real bases among them)      (common.cuh)
//
// The level-1 partition reads the text sequentially and has those 28 bases in registers anyway (three
// consecutive words cover every window of a thread's 16 positions); level 2 and the tile sorter stream the
// records.  Against the 8-byte records of tiles.cuh this adds 32 bytes of streaming per suffix and removes one
// random 64-byte DRAM access per suffix (3.1 Gbp: the text is 775 MB, L2 126 MB).
//
//   k_keyed_partition_dense / _sparse   block = 256 threads x 16 positions (x reps when the code range selects
//                            only a fraction of the suffixes); counts per bin in shared memory, one global
//                            reservation per block and bin, records staged bin by bin, written as contiguous runs.
//   k_keyed_partition_rec    block = 512 threads x 8 records of one level-1 bin, same scheme.
//   k_tile_sort_keyed        tiles.cuh's tile sorter with the gather removed.
// Included by gtsfx_b200.cu after tiles.cuh.
#pragma once

#define KP_CHUNK 4096                   // records staged per block (64 KB)
#define KP_PPT 16                       // positions per thread and repetition
#define KP_MAXREPS 8
// dynamic shared memory: stage[KP_CHUNK] 16 B | h[SP_MAXBINS] | lstart[SP_MAXBINS] | misc[16]   (72 KB: three blocks per SM)
#define KP_SMEM (KP_CHUNK * 16 + 2 * SP_MAXBINS * 4 + 64)

// exclusive scan of h[0..nbins) into lstart[] for a block of NT threads (NT = 256 or 512), nbins <= 1024;
// sm holds NT/32 words
template <int NT>
__device__ __forceinline__ uint32_t kp_scan_bins(const uint32_t *h, uint32_t *lstart, uint32_t nbins, uint32_t *sm) {
  constexpr int PER = SP_MAXBINS / NT;                 // 4 or 2 bins per thread
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  uint32_t v[PER], s = 0;
#pragma unroll
  for (int k = 0; k < PER; ++k) {
    const uint32_t i = threadIdx.x * PER + k;
    v[k] = i < nbins ? h[i] : 0;
    s += v[k];
  }
  uint32_t x = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(GTB_FULL, x, o);
    if (lane >= o) x += y;
  }
  __syncthreads();
  if (lane == 31) sm[wid] = x;
  __syncthreads();
  uint32_t woff = 0, tot = 0;
#pragma unroll
  for (int k = 0; k < NT / 32; ++k) {
    const uint32_t t = sm[k];
    if (k < wid) woff += t;
    tot += t;
  }
  uint32_t ex = woff + x - s;
#pragma unroll
  for (int k = 0; k < PER; ++k) {
    const uint32_t i = threadIdx.x * PER + k;
    if (i < nbins) lstart[i] = ex;
    ex += v[k];
  }
  return tot;
}

// the 16 positions p0..p0+15 of one thread (p0 a multiple of 16): three text words and 80 mask bits
struct KpWin {
  uint64_t w0, w1, w2, mw, mw1;
  unsigned o0;
};

__device__ __forceinline__ KpWin kp_load(const Text<2> &T, uint64_t p0) {
  KpWin k;
  const uint64_t t = p0 >> 5;
  const unsigned s = (unsigned)(p0 & 63);
  k.o0 = (unsigned)(p0 & 31);
  k.w0 = __ldg(T.w + t);
  k.w1 = __ldg(T.w + t + 1);
  k.w2 = __ldg(T.w + t + 2);
  const uint64_t m0 = __ldg(T.m + (p0 >> 6)), m1 = __ldg(T.m + (p0 >> 6) + 1);
  k.mw = s ? (m0 >> s) | (m1 << (64 - s)) : m0;                       // positions p0 .. p0+63
  k.mw1 = s ? (m1 >> s) | (__ldg(T.m + (p0 >> 6) + 2) << (64 - s)) : m1;   // positions p0+64 .. p0+127
  return k;
}

__device__ __forceinline__ uint64_t kp_code(const KpWin &k, int i, unsigned pl) {
  const unsigned o = k.o0 + i;
  const uint64_t win = o ? (k.w0 << (2 * o)) | (k.w1 >> (64 - 2 * o)) : k.w0;
  return win >> (64 - 2 * pl);
}

// bit i = position p0+i starts a suffix with >= pl real symbols whose code lies in [mincode, maxcode]
__device__ __forceinline__ uint32_t kp_select(const KpWin &k, unsigned pl, uint64_t mincode, uint64_t maxcode, bool whole) {
  const uint64_t plmask = (1ull << pl) - 1ull;
  uint32_t sel = 0;
#pragma unroll
  for (int i = 0; i < KP_PPT; ++i) {
    if (((k.mw >> i) & plmask) != 0) continue;
    if (!whole) {
      const uint64_t code = kp_code(k, i, pl);
      if (code < mincode || code > maxcode) continue;
    }
    sel |= 1u << i;
  }
  return sel;
}

// the 16-byte record of position p0+i (selected, so it has >= pl real symbols ahead)
__device__ __forceinline__ ulonglong2 kp_record(const KpWin &k, int i, uint64_t p0, unsigned pl, uint64_t code) {
  constexpr unsigned W = Text<2>::W;
  const uint64_t m = i ? (k.mw >> i) | (k.mw1 << (64 - i)) : k.mw;          // specials ahead of p0+i
  const unsigned L = m ? (unsigned)(__ffsll((long long)m) - 1) : 63u;
  const unsigned Lc = L < 63u ? L : 63u;
  const unsigned ko = k.o0 + i + pl;                                         // 1 .. 31+pl
  uint64_t kw;
  if (ko < 32) kw = (k.w0 << (2 * ko)) | (k.w1 >> (64 - 2 * ko));
  else if (ko == 32) kw = k.w1;
  else kw = (k.w1 << (2 * (ko - 32))) | (k.w2 >> (64 - 2 * (ko - 32)));
  unsigned avail = Lc - pl;
  if (avail > W) avail = W;
  uint64_t x = kw >> 8;
  if (avail < W) x |= (~0ull) >> (64 - (W - avail) * 2);
  ulonglong2 r;
  r.x = ((uint64_t)Lc << 58) | (code << 32) | (uint64_t)(uint32_t)(p0 + i);
  r.y = (x << 8) | (uint64_t)(W - avail);
  return r;
}

// Level 1 when the code range selects only a fraction of the suffixes (one of P parts, one of N GPUs, one slice of
// the pipelined fetch): block b scans positions [b*reps*4096, (b+1)*reps*4096), reps <= 8, so that it still collects
// about 4096 records; bin = (code-mincode)>>shift.  Same funnel-shift extraction as the dense kernel below for
// windows without specials; only the selected positions (about 16/reps per thread and repetition) reach the
// shared-memory atomics.  A block that collects more than the staging buffer holds writes its records straight to
// their reserved ranges (correct, just not coalesced).
__global__ void __launch_bounds__(256) k_keyed_partition_sparse(Text<2> T, unsigned pl, uint64_t n, uint64_t mincode,
                                                                 uint64_t maxcode, int shift, uint32_t nbins, int reps,
                                                                 uint32_t *__restrict__ gcursor,
                                                                 ulonglong2 *__restrict__ rec_out) {
  extern __shared__ __align__(16) unsigned char kp_raw[];
  ulonglong2 *stage = (ulonglong2 *)kp_raw;
  uint32_t *h = (uint32_t *)(stage + KP_CHUNK);
  uint32_t *lstart = h + SP_MAXBINS;                     // local bin starts, then the staging cursors
  uint32_t *misc = lstart + SP_MAXBINS;
  for (uint32_t i = threadIdx.x; i < nbins; i += 256) h[i] = 0;
  __syncthreads();
  const uint64_t pbase = (uint64_t)blockIdx.x * reps * KP_CHUNK + (uint64_t)threadIdx.x * KP_PPT;
  const unsigned cs = 32u - 2u * pl;
  const uint32_t min32 = (uint32_t)mincode, max32 = (uint32_t)maxcode;
  uint64_t sb0 = 0, sb1 = 0;                             // 16 selection bits per repetition
  // the window of repetition r+1 is requested before repetition r is processed: one memory round trip per pass
  // instead of one per repetition (the scan is latency-, not bandwidth-bound: a rank reads the text once)
  KpWin knext;
  knext.w0 = knext.w1 = knext.w2 = 0; knext.mw = knext.mw1 = ~0ull; knext.o0 = 0;
  if (pbase < n) knext = kp_load(T, pbase);
#pragma unroll 1
  for (int rep = 0; rep < reps; ++rep) {
    const uint64_t p0 = pbase + (uint64_t)rep * KP_CHUNK;
    if (p0 >= n) break;
    const KpWin k = knext;
    if (rep + 1 < reps && p0 + KP_CHUNK < n) knext = kp_load(T, p0 + KP_CHUNK);
    uint32_t sel = 0;
    if (k.mw == 0ull && (k.mw1 & 0x7fffull) == 0ull) {
      const uint32_t u0 = k.o0 == 0 ? (uint32_t)(k.w0 >> 32) : (uint32_t)k.w0;
      const uint32_t u1 = k.o0 == 0 ? (uint32_t)k.w0 : (uint32_t)(k.w1 >> 32);
      // select with the loop unrolled (four instructions per position), then touch shared memory only for the
      // selected ones: predicated bodies for all 16 positions cost 62 instructions per position (ncu), seven in
      // eight of them for nothing when the range is 1/8 of the codes
#pragma unroll
      for (int i = 0; i < KP_PPT; ++i) {
        const uint32_t code = (i ? __funnelshift_l(u1, u0, 2 * i) : u0) >> cs;
        if (code - min32 <= max32 - min32) sel |= 1u << i;
      }
      uint32_t s2 = sel;
      while (s2) {
        const int i = __ffs(s2) - 1;
        s2 &= s2 - 1;
        atomicAdd(&h[((__funnelshift_l(u1, u0, 2 * i) >> cs) - min32) >> shift], 1u);
      }
    } else {
      sel = kp_select(k, pl, mincode, maxcode, false);
      uint32_t s2 = sel;
      while (s2) {
        const int i = __ffs(s2) - 1;
        s2 &= s2 - 1;
        atomicAdd(&h[(uint32_t)((kp_code(k, i, pl) - mincode) >> shift)], 1u);
      }
    }
    if (rep < 4) sb0 |= (uint64_t)sel << (16 * rep); else sb1 |= (uint64_t)sel << (16 * (rep - 4));
  }
  __syncthreads();
  const uint32_t total = kp_scan_bins<256>(h, lstart, nbins, misc);
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < nbins; i += 256) {
    const uint32_t cnt = h[i];
    h[i] = cnt ? atomicAdd(&gcursor[i], cnt) - lstart[i] : 0u;      // global index = h[bin] + local index
  }
  __syncthreads();
  const bool staged = total <= KP_CHUNK;
  if (pbase < n) knext = kp_load(T, pbase);
#pragma unroll 1
  for (int rep = 0; rep < reps; ++rep) {
    const uint32_t sel = (uint32_t)((rep < 4 ? sb0 >> (16 * rep) : sb1 >> (16 * (rep - 4))) & 0xffffull);
    const uint64_t p0 = pbase + (uint64_t)rep * KP_CHUNK;
    if (p0 >= n) break;
    const KpWin k = knext;
    if (rep + 1 < reps && p0 + KP_CHUNK < n) knext = kp_load(T, p0 + KP_CHUNK);
    if (!sel) continue;
    if (k.mw == 0ull && (k.mw1 & 0x7fffull) == 0ull) {
      uint32_t u0, u1, u2, u3;
      if (k.o0 == 0) { u0 = (uint32_t)(k.w0 >> 32); u1 = (uint32_t)k.w0; u2 = (uint32_t)(k.w1 >> 32); u3 = (uint32_t)k.w1; }
      else           { u0 = (uint32_t)k.w0; u1 = (uint32_t)(k.w1 >> 32); u2 = (uint32_t)k.w1; u3 = (uint32_t)(k.w2 >> 32); }
      const unsigned ks = 2u * pl;
      const uint32_t v0 = __funnelshift_l(u1, u0, ks), v1 = __funnelshift_l(u2, u1, ks), v2 = __funnelshift_l(u3, u2, ks);
      uint32_t s2 = sel;
      while (s2) {                                           // a funnel shift takes its amount from a register just as well
        const int i = __ffs(s2) - 1;
        s2 &= s2 - 1;
        const uint32_t code = __funnelshift_l(u1, u0, 2 * i) >> cs;
        const uint32_t bin = (code - min32) >> shift;
        ulonglong2 rec;
        rec.x = ((uint64_t)((63u << 26) | code) << 32) | (uint64_t)(uint32_t)(p0 + i);
        rec.y = ((uint64_t)__funnelshift_l(v1, v0, 2 * i) << 32) | (uint64_t)(__funnelshift_l(v2, v1, 2 * i) & 0xffffff00u);
        const uint32_t r = atomicAdd(&lstart[bin], 1u);
        if (staged) stage[r] = rec;
        else rec_out[h[bin] + r] = rec;
      }
    } else {
      uint32_t s2 = sel;
      while (s2) {
        const int i = __ffs(s2) - 1;
        s2 &= s2 - 1;
        const uint64_t code = kp_code(k, i, pl);
        const uint32_t bin = (uint32_t)((code - mincode) >> shift);
        const ulonglong2 rec = kp_record(k, i, p0, pl, code);
        const uint32_t r = atomicAdd(&lstart[bin], 1u);
        if (staged) stage[r] = rec;
        else rec_out[h[bin] + r] = rec;
      }
    }
  }
  if (!staged) return;
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < total; i += 256) {
    const ulonglong2 r = stage[i];
    const uint32_t bin = (((uint32_t)(r.x >> 32) & (uint32_t)REC_CODE_MASK) - min32) >> shift;
    __stcs(rec_out + h[bin] + i, r);
  }
}

// Level 1 when the code range covers (nearly) all suffixes: block = 4096 consecutive positions, 16 per thread, at
// most 4096 records.  The thread's 16 positions start on a 32-bit boundary of the packed text, so with the loop
// over its positions unrolled every code and every key is a funnel shift by a constant: u[] = the text from p0 on
// as 32-bit words, v[] = the same stream advanced by pl bases.  A thread whose next 79 positions hold no special
// (all 16 suffixes have full keys) never looks at the bitmap again; the others take the generic per-position code.
// Ranks inside the bins come from the counting pass (old value of the shared-memory counter) and stay in registers.
__global__ void __launch_bounds__(256) k_keyed_partition_dense(Text<2> T, unsigned pl, uint64_t n, uint64_t mincode,
                                                                uint64_t maxcode, int whole, int shift, uint32_t nbins,
                                                                uint32_t *__restrict__ gcursor,
                                                                ulonglong2 *__restrict__ rec_out) {
  extern __shared__ __align__(16) unsigned char kp_raw[];
  ulonglong2 *stage = (ulonglong2 *)kp_raw;
  uint32_t *h = (uint32_t *)(stage + KP_CHUNK);
  uint32_t *lstart = h + SP_MAXBINS;
  uint32_t *misc = lstart + SP_MAXBINS;
  for (uint32_t i = threadIdx.x; i < nbins; i += 256) h[i] = 0;
  __syncthreads();
  const uint64_t p0 = (uint64_t)blockIdx.x * KP_CHUNK + (uint64_t)threadIdx.x * KP_PPT;
  KpWin k;
  k.w0 = k.w1 = k.w2 = 0; k.mw = k.mw1 = ~0ull; k.o0 = 0;
  if (p0 < n) k = kp_load(T, p0);
  // the text from position p0 on as 32-bit words (16 bases each)
  uint32_t u0, u1, u2, u3;
  if (k.o0 == 0) { u0 = (uint32_t)(k.w0 >> 32); u1 = (uint32_t)k.w0; u2 = (uint32_t)(k.w1 >> 32); u3 = (uint32_t)k.w1; }
  else           { u0 = (uint32_t)k.w0; u1 = (uint32_t)(k.w1 >> 32); u2 = (uint32_t)k.w1; u3 = (uint32_t)(k.w2 >> 32); }
  const unsigned cs = 32u - 2u * pl;                     // code = leading 2 pl bits of a 32-bit window (pl <= 15)
  const uint32_t min32 = (uint32_t)mincode, max32 = (uint32_t)maxcode;
  const bool clean = k.mw == 0ull && (k.mw1 & 0x7fffull) == 0ull;      // positions p0 .. p0+78 are all real
  uint32_t sel;
  if (clean) {
    sel = 0xffffu;
    if (!whole) {
      sel = 0;
#pragma unroll
      for (int i = 0; i < KP_PPT; ++i) {
        const uint32_t code = (i ? __funnelshift_l(u1, u0, 2 * i) : u0) >> cs;
        if (code >= min32 && code <= max32) sel |= 1u << i;
      }
    }
  } else {
    sel = p0 < n ? kp_select(k, pl, mincode, maxcode, whole != 0) : 0u;
  }
  uint32_t rk[KP_PPT / 2];                               // rank inside the bin (< 4096), two per word
#pragma unroll
  for (int i = 0; i < KP_PPT / 2; ++i) rk[i] = 0;
#pragma unroll
  for (int i = 0; i < KP_PPT; ++i) {
    if ((sel >> i) & 1u) {
      const uint32_t code = (i ? __funnelshift_l(u1, u0, 2 * i) : u0) >> cs;
      const uint32_t r = atomicAdd(&h[(code - min32) >> shift], 1u);
      rk[i >> 1] |= r << ((i & 1) * 16);
    }
  }
  __syncthreads();
  const uint32_t total = kp_scan_bins<256>(h, lstart, nbins, misc);
  __syncthreads();
  // reservations first (four bins per thread); their round trip hides behind the staging below
  uint32_t res[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint32_t i = threadIdx.x + q * 256;
    const uint32_t cnt = i < nbins ? h[i] : 0u;
    res[q] = cnt ? atomicAdd(&gcursor[i], cnt) : 0u;
  }
  if (clean) {
    // the stream advanced by pl bases: the key of position p0+i is the 56 bits at bit offset 2 i of v0 v1 v2
    const unsigned ks = 2u * pl;
    const uint32_t v0 = __funnelshift_l(u1, u0, ks), v1 = __funnelshift_l(u2, u1, ks), v2 = __funnelshift_l(u3, u2, ks);
#pragma unroll
    for (int i = 0; i < KP_PPT; ++i) {
      if ((sel >> i) & 1u) {
        const uint32_t code = (i ? __funnelshift_l(u1, u0, 2 * i) : u0) >> cs;
        const uint32_t khi = i ? __funnelshift_l(v1, v0, 2 * i) : v0;
        const uint32_t klo = (i ? __funnelshift_l(v2, v1, 2 * i) : v1) & 0xffffff00u;
        ulonglong2 rec;
        rec.x = ((uint64_t)((63u << 26) | code) << 32) | (uint64_t)(uint32_t)(p0 + i);
        rec.y = ((uint64_t)khi << 32) | (uint64_t)klo;
        stage[lstart[(code - min32) >> shift] + ((rk[i >> 1] >> ((i & 1) * 16)) & 0xffffu)] = rec;
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < KP_PPT; ++i) {
      if ((sel >> i) & 1u) {
        const uint64_t code = kp_code(k, i, pl);
        const ulonglong2 rec = kp_record(k, i, p0, pl, code);
        stage[lstart[(uint32_t)((code - mincode) >> shift)] + ((rk[i >> 1] >> ((i & 1) * 16)) & 0xffffu)] = rec;
      }
    }
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint32_t i = threadIdx.x + q * 256;
    if (i < nbins) h[i] = res[q] - lstart[i];             // global index = h[bin] + local index (own bins only)
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < total; i += 256) {
    const ulonglong2 r = stage[i];
    const uint32_t bin = (((uint32_t)(r.x >> 32) & (uint32_t)REC_CODE_MASK) - min32) >> shift;
    __stcs(rec_out + h[bin] + i, r);
  }
}

// Level 2, from the records of one level-1 bin.  Block b handles KP_SUPER consecutive 4096-record chunks of one
// parent (blk_first = prefix sums of the parents' block counts); bin = child index inside the parent.  Two things
// keep the memory pipe busy although only two such blocks fit an SM: the records of chunk j+1 are requested before
// chunk j is written out, and the global reservations of a chunk are issued before its records are staged, so their
// round trip is hidden behind the staging.
#define KP_SUPER 4
__global__ void __launch_bounds__(512, 2) k_keyed_partition_rec(uint64_t mincode, int shift, uint32_t nbins,
                                                              const ulonglong2 *__restrict__ rec_in,
                                                              const uint32_t *__restrict__ seg_start,
                                                              const uint32_t *__restrict__ seg_end,
                                                              const uint32_t *__restrict__ blk_first, uint32_t nparents,
                                                              int parent_shift, uint32_t *__restrict__ gcursor,
                                                              ulonglong2 *__restrict__ rec_out) {
  extern __shared__ __align__(16) unsigned char kp_raw[];
  ulonglong2 *stage = (ulonglong2 *)kp_raw;
  uint32_t *h = (uint32_t *)(stage + KP_CHUNK);
  uint32_t *lstart = h + SP_MAXBINS;
  uint32_t *misc = lstart + SP_MAXBINS;                // [0..15] scan scratch, [8] parent, [9] block inside the parent
  if (threadIdx.x == 0) {
    uint32_t lo2 = 0, hi2 = nparents;                  // blk_first[lo2] <= blockIdx.x < blk_first[hi2]
    if (blockIdx.x >= blk_first[nparents]) {
      misc[8] = 0xffffffffu;
    } else {
      while (hi2 - lo2 > 1) {
        const uint32_t mid = (lo2 + hi2) >> 1;
        if (blk_first[mid] <= blockIdx.x) lo2 = mid; else hi2 = mid;
      }
      misc[8] = lo2;
      misc[9] = blockIdx.x - blk_first[lo2];
    }
  }
  __syncthreads();
  const uint32_t parent = misc[8];
  if (parent == 0xffffffffu) return;
  const uint32_t sbeg = seg_start[parent], send = seg_end[parent];
  const uint32_t rbeg0 = sbeg + misc[9] * (KP_CHUNK * KP_SUPER);
  const uint32_t rend = send - rbeg0 < KP_CHUNK * KP_SUPER ? send : rbeg0 + KP_CHUNK * KP_SUPER;
  const uint64_t binbase = (uint64_t)parent << (parent_shift - shift);
  const uint32_t c0 = (uint32_t)mincode;
  ulonglong2 r8[8];
  {
    const uint32_t rcnt = rend - rbeg0 < KP_CHUNK ? rend - rbeg0 : KP_CHUNK;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t j = k * 512 + threadIdx.x;
      r8[k] = j < rcnt ? __ldcs(rec_in + rbeg0 + j) : make_ulonglong2(0ull, 0ull);
    }
  }
  for (uint32_t rbeg = rbeg0; rbeg < rend; rbeg += KP_CHUNK) {
    const uint32_t rcnt = rend - rbeg < KP_CHUNK ? rend - rbeg : KP_CHUNK;
    __syncthreads();                                   // scan scratch read / previous chunk written out
    for (uint32_t i = threadIdx.x; i < nbins; i += 512) h[i] = 0;
    __syncthreads();
    uint32_t rk[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) rk[k] = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t j = k * 512 + threadIdx.x;
      if (j < rcnt) {
        const uint32_t bin = (uint32_t)(((((uint32_t)(r8[k].x >> 32) & (uint32_t)REC_CODE_MASK) - c0) >> shift) - binbase);
        const uint32_t q = atomicAdd(&h[bin], 1u);
        rk[k >> 1] |= q << ((k & 1) * 16);
      }
    }
    __syncthreads();
    const uint32_t total = kp_scan_bins<512>(h, lstart, nbins, misc);
    __syncthreads();
    // reservations first (two bins per thread), their results are needed only after the staging
    uint32_t res[2];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const uint32_t i = threadIdx.x + q * 512;
      const uint32_t cnt = i < nbins ? h[i] : 0u;
      res[q] = cnt ? atomicAdd(&gcursor[binbase + i], cnt) : 0u;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t j = k * 512 + threadIdx.x;
      if (j < rcnt) {
        const uint32_t bin = (uint32_t)(((((uint32_t)(r8[k].x >> 32) & (uint32_t)REC_CODE_MASK) - c0) >> shift) - binbase);
        stage[lstart[bin] + ((rk[k >> 1] >> ((k & 1) * 16)) & 0xffffu)] = r8[k];
      }
    }
    __syncthreads();                                   // every h[] was read by the reservation, every record is staged
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const uint32_t i = threadIdx.x + q * 512;
      if (i < nbins) h[i] = res[q] - lstart[i];        // global index = h[bin] + local index
    }
    if (rbeg + KP_CHUNK < rend) {                      // next chunk's records: in flight during the write-out
      const uint32_t nb = rbeg + KP_CHUNK;
      const uint32_t ncnt = rend - nb < KP_CHUNK ? rend - nb : KP_CHUNK;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint32_t j = k * 512 + threadIdx.x;
        r8[k] = j < ncnt ? __ldcs(rec_in + nb + j) : make_ulonglong2(0ull, 0ull);
      }
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < total; i += 512) {
      const ulonglong2 r = stage[i];
      const uint32_t bin = (uint32_t)(((((uint32_t)(r.x >> 32) & (uint32_t)REC_CODE_MASK) - c0) >> shift) - binbase);
      __stcs(rec_out + h[bin] + i, r);
    }
  }
}

// blk_first[0..nparents]: prefix sums of ceil(records/(KP_SUPER*KP_CHUNK)) per level-1 bin (single block)
__global__ void __launch_bounds__(256) k_keyed_chunk_table(const uint32_t *__restrict__ seg_start,
                                                           const uint32_t *__restrict__ seg_end, uint32_t nparents,
                                                           uint32_t *blk_first) {
  __shared__ uint32_t sm[8];
  uint32_t carry = 0;
  for (uint32_t base = 0; base < nparents; base += 256) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t v = i < nparents ? (seg_end[i] - seg_start[i] + KP_CHUNK * KP_SUPER - 1) / (KP_CHUNK * KP_SUPER) : 0;
    uint32_t tot;
    const uint32_t ex = block_excl_scan256(v, &tot, sm);
    if (i < nparents) blk_first[i] = carry + ex;
    carry += tot;
  }
  if (threadIdx.x == 0) blk_first[nparents] = carry;
}

// tiles.cuh's k_tile_sort for keyed records: insertion only moves (position, key) into the bucket's slots.
// keys_out (byte text): buckets the tile does not sort itself (more than 128 suffixes: the rule on protein) leave
// their keys beside their positions, keys_out[r] for out.sa[r], for k_sort_medium.
template <int BITS>
__global__ void __launch_bounds__(TS_THREADS, 2) k_tile_sort_keyed(Text<BITS> T, PartOut out, const uint32_t *__restrict__ lb,
                                                                   uint32_t *__restrict__ cursor,
                                                                   const ulonglong2 *__restrict__ rec,
                                                                   const uint32_t *__restrict__ gcursor, uint64_t mincode,
                                                                   uint64_t maxcode, int shift, uint32_t nbins, unsigned pl,
                                                                   uint32_t *__restrict__ fbbits, uint64_t tieoff,
                                                                   uint64_t *__restrict__ keys_out) {
  extern __shared__ __align__(16) unsigned char ts_raw[];
  uint64_t *s_key = (uint64_t *)ts_raw;
  uint32_t *s_pos = (uint32_t *)(s_key + TILE_CAP);
  uint32_t *s_lb = s_pos + TILE_CAP;
  uint32_t *s_cur = s_lb + TILE_MAXCODES + 1;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  unsigned long long sum = 0;
  uint32_t mx = 0;
  for (uint32_t g = blockIdx.x; g < nbins; g += gridDim.x) {
    const uint64_t c0 = mincode + ((uint64_t)g << shift);
    uint64_t c1 = c0 + (1ull << shift);
    if (c1 > maxcode + 1) c1 = maxcode + 1;
    const uint32_t ncf = (uint32_t)(c1 - c0);
    const uint32_t base = lb[c0];                       // global suffix-array index of the tile
    const uint32_t slots = lb[c1] - base;
    const uint32_t rs = base - out.partoffset;          // part-local; also the tile's first record
    const uint32_t cnt = gcursor[g] - rs;
    __syncthreads();                                    // previous tile fully consumed
    for (uint32_t i = threadIdx.x; i <= ncf; i += TS_THREADS) {
      const uint32_t v = lb[c0 + i] - base;
      s_lb[i] = v;
      if (i < ncf) s_cur[i] = v;
    }
    __syncthreads();
    if (slots > TILE_CAP) {
      // too many suffixes for shared memory: plain insertion, the global kernels sort these buckets
      for (uint32_t i = threadIdx.x; i < cnt; i += TS_THREADS) {
        const ulonglong2 a2 = __ldcs(rec + rs + i);
        const uint64_t a = a2.x;
        const uint32_t slot = atomicAdd(&s_cur[(uint32_t)(((a >> 32) & REC_CODE_MASK) - c0)], 1u);
        out.sa[rs + slot] = (uint32_t)a;
        if (keys_out) keys_out[rs + slot] = a2.y;
      }
      __syncthreads();
      for (uint32_t i = threadIdx.x; i < ncf; i += TS_THREADS) cursor[c0 + i] = base + s_cur[i];
      if (threadIdx.x == 0) atomicAdd(&out.stats->tile_leftover, 1u);
      continue;
    }
    for (uint32_t i0 = threadIdx.x; i0 < cnt; i0 += TS_THREADS * TS_BATCH) {
      ulonglong2 r[TS_BATCH];
#pragma unroll
      for (int u = 0; u < TS_BATCH; ++u) {
        const uint32_t i = i0 + u * TS_THREADS;
        r[u] = i < cnt ? __ldcs(rec + rs + i) : make_ulonglong2(0ull, 0ull);
      }
#pragma unroll
      for (int u = 0; u < TS_BATCH; ++u) {
        const uint32_t i = i0 + u * TS_THREADS;
        if (i < cnt) {
          const uint32_t slot = atomicAdd(&s_cur[(uint32_t)(((r[u].x >> 32) & REC_CODE_MASK) - c0)], 1u);
          s_pos[slot] = (uint32_t)r[u].x;
          s_key[slot] = r[u].y;
        }
      }
    }
    __syncthreads();
    for (uint32_t f = wid; f < ncf; f += TS_WARPS) {
      const uint32_t ls = s_lb[f];
      const uint32_t cf = s_cur[f] - ls;                // regular suffixes of bucket c0+f
      const uint32_t start = rs + ls;
      bool done = true;
      if (cf == 1) {
        if (lane == 0) put_sa(out, start, s_pos[ls]);
      } else if (cf > 1) {
        const uint32_t *src = s_pos + ls;
        const uint64_t *keys = s_key + ls;
        bool clean = true;                              // sorted and free of ties
        if (cf <= 32) clean = warp_sort_any<BITS, 1>(T, out, src, nullptr, nullptr, keys, start, cf, pl, lane, nullptr, nullptr, sum, mx);
        else if (cf <= 64) clean = warp_sort_any<BITS, 2>(T, out, src, nullptr, nullptr, keys, start, cf, pl, lane, nullptr, nullptr, sum, mx);
        else if (cf <= 96) clean = warp_sort_any<BITS, 3>(T, out, src, nullptr, nullptr, keys, start, cf, pl, lane, nullptr, nullptr, sum, mx);
        else if (cf <= SORT_MAXBUCKET) clean = warp_sort_any<BITS, 4>(T, out, src, nullptr, nullptr, keys, start, cf, pl, lane, nullptr, nullptr, sum, mx);
        else {                                          // too large for a warp: leave it to the global sorters
          for (uint32_t i = lane; i < cf; i += 32) {
            out.sa[start + i] = src[i];
            if (keys_out) keys_out[start + i] = keys[i];
          }
          done = false;
          if (lane == 0) atomicAdd(&out.stats->tile_leftover, 1u);
        }
        if (!clean && lane == 0) {                      // sorted on pl+W symbols, tie groups go to refinement
          const uint64_t rel = c0 + f - mincode;
          atomicOr(&fbbits[tieoff + (rel >> 5)], 1u << (rel & 31));
          atomicAdd(&out.stats->fb_elems, (unsigned long long)cf);
          atomicAdd(&out.stats->fb_buckets, 1u);
        }
      }
      if (lane == 0) cursor[c0 + f] = done ? base + ls : base + ls + cf;
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

// ---------------------------------------------------------------------------------------------
// Keyed records for byte text (sigma > 4: protein).  The block sorter k_sort_medium - where 98 % of a protein
// index's suffixes end up - fetched its keys from the 1 GB text at random: 125 GB of DRAM reads for 25 GB
// algorithmic at 1 G residues (profiles/r2_ncu_full_protein.txt).  Here too the level-1 partition has the 7 residues
// after the prefix in registers, so the key travels with the record; the tile sorter hands the keys of the buckets it
// only inserts to k_sort_medium through a key array beside the suffix array.
//   thread = 16 positions = 2 text words (+2 for the windows); codes are base-sigma numbers: Horner for the first
//   position, then one multiply-add per position (the identity holds modulo 2^32 whatever the digits are, so a
//   special only spoils the codes whose window contains it, and those are masked out).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t kb_special_bits(uint64_t w) {      // bit j = byte j (big-endian order) is >= 0xfe
  const uint64_t z = ((~w) >> 1) & 0x7f7f7f7f7f7f7f7full;
  const uint64_t real = (z + 0x7f7f7f7f7f7f7f7full) & 0x8080808080808080ull;
  const uint64_t t = (real ^ 0x8080808080808080ull) >> 7;
  return (uint32_t)((t * 0x8040201008040201ull) >> 56);
}

struct KbWin {
  uint64_t w0, w1, w2, w3;
  uint32_t sp;                 // special flags of the 32 bytes from p0 on
};

__device__ __forceinline__ KbWin kb_load(const Text<8> &T, uint64_t p0) {
  KbWin k;
  const uint64_t t = p0 >> 3;
  k.w0 = __ldg(T.w + t);
  k.w1 = __ldg(T.w + t + 1);
  k.w2 = __ldg(T.w + t + 2);
  k.w3 = __ldg(T.w + t + 3);
  k.sp = kb_special_bits(k.w0) | (kb_special_bits(k.w1) << 8) | (kb_special_bits(k.w2) << 16) | (kb_special_bits(k.w3) << 24);
  return k;
}

// byte j (0..31, compile-time constant after unrolling) of the window
__device__ __forceinline__ uint32_t kb_byte(const KbWin &k, int j) {
  const uint64_t w = j < 8 ? k.w0 : j < 16 ? k.w1 : j < 24 ? k.w2 : k.w3;
  return (uint32_t)(w >> (56 - 8 * (j & 7))) & 0xffu;
}

// codes of the 16 positions (garbage where the window holds a special: masked by the caller)
__device__ __forceinline__ void kb_codes(const KbWin &k, unsigned pl, uint32_t sigma, uint32_t top /* sigma^(pl-1) */,
                                         uint64_t v0, uint64_t v1, uint32_t (&code)[KP_PPT]) {
  uint32_t c = 0;
#pragma unroll
  for (int t = 0; t < 8; ++t)
    if ((unsigned)t < pl) c = c * sigma + kb_byte(k, t);
  code[0] = c;
#pragma unroll
  for (int i = 1; i < KP_PPT; ++i) {
    // leaves: digit i-1; enters: digit i-1+pl = byte i-1 of the stream advanced by pl bytes (v0 v1)
    const uint32_t in = (uint32_t)((i - 1 < 8 ? v0 : v1) >> (56 - 8 * ((i - 1) & 7))) & 0xffu;
    c = (c - kb_byte(k, i - 1) * top) * sigma + in;
    code[i] = c;
  }
}

__global__ void __launch_bounds__(256) k_keyed_partition_bytes(Text<8> T, CodeParams cp, uint64_t n, uint64_t mincode,
                                                                uint64_t maxcode, int shift, uint32_t nbins, int reps,
                                                                uint32_t *__restrict__ gcursor,
                                                                ulonglong2 *__restrict__ rec_out) {
  extern __shared__ __align__(16) unsigned char kp_raw[];
  ulonglong2 *stage = (ulonglong2 *)kp_raw;
  uint32_t *h = (uint32_t *)(stage + KP_CHUNK);
  uint32_t *lstart = h + SP_MAXBINS;                     // local bin starts, then the staging cursors
  uint32_t *misc = lstart + SP_MAXBINS;
  for (uint32_t i = threadIdx.x; i < nbins; i += 256) h[i] = 0;
  __syncthreads();
  const unsigned pl = cp.pl;                              // <= 8
  const uint32_t sigma = cp.sigma, top = (uint32_t)cp.pw[pl - 1];
  const uint32_t min32 = (uint32_t)mincode, span = (uint32_t)(maxcode - mincode), plmask = (1u << pl) - 1u;
  const uint64_t pbase = (uint64_t)blockIdx.x * reps * KP_CHUNK + (uint64_t)threadIdx.x * KP_PPT;
  uint64_t sb0 = 0, sb1 = 0;                             // 16 selection bits per repetition
#pragma unroll 1
  for (int rep = 0; rep < reps; ++rep) {
    const uint64_t p0 = pbase + (uint64_t)rep * KP_CHUNK;
    if (p0 >= n) break;
    const KbWin k = kb_load(T, p0);
    const unsigned sh = 8u * pl;
    const uint64_t v0 = pl == 8 ? k.w1 : (k.w0 << sh) | (k.w1 >> (64 - sh));
    const uint64_t v1 = pl == 8 ? k.w2 : (k.w1 << sh) | (k.w2 >> (64 - sh));
    uint32_t code[KP_PPT];
    kb_codes(k, pl, sigma, top, v0, v1, code);
    uint32_t sel = 0;
#pragma unroll
    for (int i = 0; i < KP_PPT; ++i) {
      if (((k.sp >> i) & plmask) == 0u && code[i] - min32 <= span) {
        sel |= 1u << i;
        atomicAdd(&h[(code[i] - min32) >> shift], 1u);
      }
    }
    if (rep < 4) sb0 |= (uint64_t)sel << (16 * rep); else sb1 |= (uint64_t)sel << (16 * (rep - 4));
  }
  __syncthreads();
  const uint32_t total = kp_scan_bins<256>(h, lstart, nbins, misc);
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < nbins; i += 256) {
    const uint32_t cnt = h[i];
    h[i] = cnt ? atomicAdd(&gcursor[i], cnt) - lstart[i] : 0u;      // global index = h[bin] + local index
  }
  __syncthreads();
  const bool staged = total <= KP_CHUNK;
#pragma unroll 1
  for (int rep = 0; rep < reps; ++rep) {
    const uint32_t sel = (uint32_t)((rep < 4 ? sb0 >> (16 * rep) : sb1 >> (16 * (rep - 4))) & 0xffffull);
    if (!sel) continue;
    const uint64_t p0 = pbase + (uint64_t)rep * KP_CHUNK;
    const KbWin k = kb_load(T, p0);
    const unsigned sh = 8u * pl;
    const uint64_t v0 = pl == 8 ? k.w1 : (k.w0 << sh) | (k.w1 >> (64 - sh));
    const uint64_t v1 = pl == 8 ? k.w2 : (k.w1 << sh) | (k.w2 >> (64 - sh));
    const uint64_t v2 = pl == 8 ? k.w3 : (k.w2 << sh) | (k.w3 >> (64 - sh));
    uint32_t code[KP_PPT];
    kb_codes(k, pl, sigma, top, v0, v1, code);
#pragma unroll
    for (int i = 0; i < KP_PPT; ++i) {
      if ((sel >> i) & 1u) {
        // the 8 bytes at position p0+i+pl; 7 of them are key symbols (make_key<8>, common.cuh)
        const uint64_t hi = i < 8 ? v0 : v1, lo = i < 8 ? v1 : v2;
        const uint64_t kw = (i & 7) ? (hi << (8 * (i & 7))) | (lo >> (64 - 8 * (i & 7))) : hi;
        const uint32_t a = (k.sp >> (i + pl)) & 0x7fu;
        const unsigned avail = a ? (unsigned)(__ffs(a) - 1) : 7u;
        uint64_t x = kw >> 8;
        if (avail < 7u) x |= (~0ull) >> (64 - (7u - avail) * 8u);
        ulonglong2 rec;
        rec.x = ((uint64_t)((63u << 26) | code[i]) << 32) | (uint64_t)(uint32_t)(p0 + i);
        rec.y = (x << 8) | (uint64_t)(7u - avail);
        const uint32_t bin = (code[i] - min32) >> shift;
        const uint32_t r = atomicAdd(&lstart[bin], 1u);
        if (staged) stage[r] = rec;
        else rec_out[h[bin] + r] = rec;
      }
    }
  }
  if (!staged) return;
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < total; i += 256) {
    const ulonglong2 r = stage[i];
    const uint32_t bin = (((uint32_t)(r.x >> 32) & (uint32_t)REC_CODE_MASK) - min32) >> shift;
    __stcs(rec_out + h[bin] + i, r);
  }
}

// ---------------------------------------------------------------------------------------------
// Tile sorter for keyed records on byte text.  With sigma = 20 and pl = 5 a bucket holds ~300 suffixes - too many
// for one warp network, and the block sorter that took them needs 22 warp instructions per suffix.  Inside the tile
// every bucket is therefore split once more by the first key symbol (sigma + 1 sub-buckets: the last one collects
// keys that start with a special), which brings the groups down to ~15 suffixes, and the warp networks sort those
// from the keys in shared memory.  Neighbours in different sub-buckets differ in their first symbol after the
// prefix, so the LCP at a sub-bucket border is exactly pl.
//   P1 histogram over (bucket, first symbol) - P2 prefix per bucket - P3 scatter (second read of the records: L2) -
//   P4 one warp per sub-bucket.  A bucket with a sub-bucket of more than 128 suffixes is left to the global sorters
//   (positions and keys written out unsorted), like every bucket of a tile that does not fit shared memory.
// ---------------------------------------------------------------------------------------------
#define TSUB_MAX 4096                    // (bucket, symbol) counters per tile
// dynamic shared memory: key[TILE_CAP] u64 | pos[TILE_CAP] | lb[TILE_MAXCODES+1] | sub[TSUB_MAX] | cur[TSUB_MAX] | big[TILE_MAXCODES] u8
#define TSUB_SMEM (TILE_CAP * 12 + (TILE_MAXCODES + 1) * 4 + 2 * TSUB_MAX * 4 + TILE_MAXCODES + 12)

__global__ void __launch_bounds__(TS_THREADS, 2) k_tile_subsort(Text<8> T, PartOut out, const uint32_t *__restrict__ lb,
                                                                uint32_t *__restrict__ cursor,
                                                                const ulonglong2 *__restrict__ rec,
                                                                const uint32_t *__restrict__ gcursor, uint64_t mincode,
                                                                uint64_t maxcode, int shift, uint32_t nbins, unsigned pl,
                                                                uint32_t sigma, uint32_t *__restrict__ fbbits, uint64_t tieoff,
                                                                uint64_t *__restrict__ keys_out) {
  extern __shared__ __align__(16) unsigned char ts_raw[];
  uint64_t *s_key = (uint64_t *)ts_raw;
  uint32_t *s_pos = (uint32_t *)(s_key + TILE_CAP);
  uint32_t *s_lb = s_pos + TILE_CAP;
  uint32_t *s_sub = s_lb + TILE_MAXCODES + 1;           // start slot of every (bucket, symbol) group
  uint32_t *s_cur = s_sub + TSUB_MAX;                   // counters, then cursors, then end slots
  unsigned char *s_big = (unsigned char *)(s_cur + TSUB_MAX);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const uint32_t S1 = sigma + 1;
  unsigned long long sum = 0;
  uint32_t mx = 0;
  for (uint32_t g = blockIdx.x; g < nbins; g += gridDim.x) {
    const uint64_t c0 = mincode + ((uint64_t)g << shift);
    uint64_t c1 = c0 + (1ull << shift);
    if (c1 > maxcode + 1) c1 = maxcode + 1;
    const uint32_t ncf = (uint32_t)(c1 - c0);
    const uint32_t base = lb[c0];
    const uint32_t slots = lb[c1] - base;
    const uint32_t rs = base - out.partoffset;
    const uint32_t cnt = gcursor[g] - rs;
    const uint32_t nsub = ncf * S1;
    __syncthreads();                                    // previous tile fully consumed
    for (uint32_t i = threadIdx.x; i <= ncf; i += TS_THREADS) s_lb[i] = lb[c0 + i] - base;
    for (uint32_t i = threadIdx.x; i < nsub; i += TS_THREADS) s_cur[i] = 0;
    for (uint32_t i = threadIdx.x; i < ncf; i += TS_THREADS) s_big[i] = 0;
    __syncthreads();
    if (slots > TILE_CAP) {
      // too many suffixes for shared memory: plain insertion, the global kernels sort these buckets
      for (uint32_t i = threadIdx.x; i < ncf; i += TS_THREADS) s_sub[i] = s_lb[i];
      __syncthreads();
      for (uint32_t i = threadIdx.x; i < cnt; i += TS_THREADS) {
        const ulonglong2 a = __ldcs(rec + rs + i);
        const uint32_t slot = atomicAdd(&s_sub[(uint32_t)(((a.x >> 32) & REC_CODE_MASK) - c0)], 1u);
        out.sa[rs + slot] = (uint32_t)a.x;
        keys_out[rs + slot] = a.y;
      }
      __syncthreads();
      for (uint32_t i = threadIdx.x; i < ncf; i += TS_THREADS) cursor[c0 + i] = base + s_sub[i];
      if (threadIdx.x == 0) atomicAdd(&out.stats->tile_leftover, 1u);
      continue;
    }
    // ---- P1: histogram over (bucket, first key symbol) ----
    for (uint32_t i0 = threadIdx.x; i0 < cnt; i0 += TS_THREADS * TS_BATCH) {
      ulonglong2 r[TS_BATCH];
#pragma unroll
      for (int u = 0; u < TS_BATCH; ++u) {
        const uint32_t i = i0 + u * TS_THREADS;
        r[u] = i < cnt ? __ldg(rec + rs + i) : make_ulonglong2(0ull, 0ull);
      }
#pragma unroll
      for (int u = 0; u < TS_BATCH; ++u) {
        const uint32_t i = i0 + u * TS_THREADS;
        if (i < cnt) {
          const uint32_t f = ((uint32_t)(r[u].x >> 32) & (uint32_t)REC_CODE_MASK) - (uint32_t)c0;
          const uint32_t sy = min((uint32_t)(r[u].y >> 56), sigma);
          atomicAdd(&s_cur[f * S1 + sy], 1u);
        }
      }
    }
    __syncthreads();
    // ---- P2: start slot of every group; a bucket with a group too large for a warp is given up as a whole ----
    for (uint32_t f = threadIdx.x; f < ncf; f += TS_THREADS) {
      uint32_t run = s_lb[f];
      bool big = false;
      for (uint32_t sy = 0; sy < S1; ++sy) {
        const uint32_t c = s_cur[f * S1 + sy];
        s_sub[f * S1 + sy] = run;
        s_cur[f * S1 + sy] = run;
        run += c;
        big |= c > SORT_MAXBUCKET;
      }
      s_big[f] = big ? 1 : 0;
    }
    __syncthreads();
    // ---- P3: scatter into the groups ----
    for (uint32_t i0 = threadIdx.x; i0 < cnt; i0 += TS_THREADS * TS_BATCH) {
      ulonglong2 r[TS_BATCH];
#pragma unroll
      for (int u = 0; u < TS_BATCH; ++u) {
        const uint32_t i = i0 + u * TS_THREADS;
        r[u] = i < cnt ? __ldcs(rec + rs + i) : make_ulonglong2(0ull, 0ull);
      }
#pragma unroll
      for (int u = 0; u < TS_BATCH; ++u) {
        const uint32_t i = i0 + u * TS_THREADS;
        if (i < cnt) {
          const uint32_t f = ((uint32_t)(r[u].x >> 32) & (uint32_t)REC_CODE_MASK) - (uint32_t)c0;
          const uint32_t sy = min((uint32_t)(r[u].y >> 56), sigma);
          const uint32_t slot = atomicAdd(&s_cur[f * S1 + sy], 1u);
          s_pos[slot] = (uint32_t)r[u].x;
          s_key[slot] = r[u].y;
        }
      }
    }
    __syncthreads();
    // ---- P4: one warp per group ----
    for (uint32_t q = wid; q < nsub; q += TS_WARPS) {
      const uint32_t f = q / S1;
      if (s_big[f]) continue;
      const uint32_t ls = s_sub[q];
      const uint32_t cf = s_cur[q] - ls;
      if (cf == 0) continue;
      const uint32_t start = rs + ls;
      bool clean = true;
      if (cf == 1) {
        if (lane == 0) put_sa(out, start, s_pos[ls]);
      } else {
        const uint32_t *src = s_pos + ls;
        const uint64_t *keys = s_key + ls;
        if (cf <= 32) clean = warp_sort_any<8, 1>(T, out, src, nullptr, nullptr, keys, start, cf, pl, lane, nullptr, nullptr, sum, mx, 8);
        else if (cf <= 64) clean = warp_sort_any<8, 2>(T, out, src, nullptr, nullptr, keys, start, cf, pl, lane, nullptr, nullptr, sum, mx, 8);
        else if (cf <= 96) clean = warp_sort_any<8, 3>(T, out, src, nullptr, nullptr, keys, start, cf, pl, lane, nullptr, nullptr, sum, mx, 8);
        else clean = warp_sort_any<8, 4>(T, out, src, nullptr, nullptr, keys, start, cf, pl, lane, nullptr, nullptr, sum, mx, 8);
      }
      if (lane == 0) {
        if (ls != s_lb[f]) {                            // first entry of a group that is not the bucket's first
          __stcs(out.lcp + start, (uint8_t)pl);
          sum += pl;
          mx = max(mx, (uint32_t)pl);
        }
        if (!clean) {                                   // sorted on pl+7 symbols, tie groups go to refinement
          const uint64_t rel = c0 + f - mincode;
          const uint32_t bit = 1u << (rel & 31);
          const uint32_t old = atomicOr(&fbbits[tieoff + (rel >> 5)], bit);
          if (!(old & bit)) {
            atomicAdd(&out.stats->fb_elems, (unsigned long long)(s_cur[f * S1 + sigma] - s_lb[f]));
            atomicAdd(&out.stats->fb_buckets, 1u);
          }
        }
      }
    }
    // ---- per bucket: what the later kernels need to know ----
    __syncthreads();
    for (uint32_t f = wid; f < ncf; f += TS_WARPS) {
      const uint32_t ls = s_lb[f];
      const uint32_t cf = s_cur[f * S1 + sigma] - ls;   // the last group's end = the bucket's regular suffixes
      if (s_big[f]) {
        for (uint32_t i = lane; i < cf; i += 32) {
          out.sa[rs + ls + i] = s_pos[ls + i];
          keys_out[rs + ls + i] = s_key[ls + i];
        }
        if (lane == 0) {
          cursor[c0 + f] = base + ls + cf;
          atomicAdd(&out.stats->tile_leftover, 1u);
        }
      } else if (lane == 0) {
        cursor[c0 + f] = base + ls;
      }
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

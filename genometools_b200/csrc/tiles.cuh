// Tile path: insertion and in-bucket sort for DNA without a scattered global pass.
//
//   k_stage_partition  groups (code,pos) records by leading code bits.  A block ranks its 8192
//                      records per bin in shared memory, reserves one range per bin with a single
//                      global atomic, stages the records bin by bin in shared memory and writes
//                      them out as contiguous runs (full sectors) - level 1 reads the text, level
//                      2 re-partitions the records of each level-1 bin.
//   k_tile_sort        one block per final bin (a few thousand suffixes of <= 1024 consecutive
//                      buckets): the bucket cursors and the suffix-array slice live in shared
//                      memory, records are dropped into their bucket with shared-memory atomics,
//                      then every bucket is sorted by a warp (same 32-bit / 64-bit networks as
//                      k_sort_buckets, reading positions from shared memory) and written out
//                      together with its LCP values.  Buckets it cannot finish (more than 128
//                      suffixes, or still tied after pl+28 symbols) are written unsorted and left
//                      to k_sort_buckets / k_sort_medium / the refinement path via cursor[].
// Included by gtsfx_b200.cu after the warp sorters.
#pragma once

#define REC_CODE_MASK 0x3ffffffull
__device__ __forceinline__ uint64_t rec_code(uint64_t r, int recL) { return recL ? ((r >> 32) & REC_CODE_MASK) : (r >> 32); }

#define SP_CHUNK 8192
#define SP_MAXBINS 1024

// dynamic shared memory: stage[SP_CHUNK] u64 | h[SP_MAXBINS] | lstart[SP_MAXBINS] | misc[16]
#define SP_SMEM (SP_CHUNK * 8 + SP_MAXBINS * 4 * 2 + 64)

// exclusive scan of h[0..nbins) into lstart[], 256 threads, nbins <= 1024; returns total
__device__ __forceinline__ uint32_t sp_scan_bins(const uint32_t *h, uint32_t *lstart, uint32_t nbins, uint32_t *sm8) {
  uint32_t v[4], s = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t i = threadIdx.x * 4 + k;
    v[k] = i < nbins ? h[i] : 0;
    s += v[k];
  }
  uint32_t tot;
  uint32_t ex = block_excl_scan256(s, &tot, sm8);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t i = threadIdx.x * 4 + k;
    if (i < nbins) lstart[i] = ex;
    ex += v[k];
  }
  return tot;
}

// FROM_TEXT: block b scans positions [b*8192, (b+1)*8192) of the text, bin = (code-mincode)>>shift.
// else     : block b handles one 8192-record chunk of one parent bin (blk_first = prefix sums of
//            the parents' chunk counts); bin = child index inside the parent.
template <int BITS, bool FROM_TEXT>
__global__ void __launch_bounds__(256) k_stage_partition(Text<BITS> T, CodeParams cp, uint64_t n, uint64_t mincode,
                                                          uint64_t maxcode, int shift, uint32_t nbins,
                                                          const uint64_t *__restrict__ rec_in,
                                                          const uint32_t *__restrict__ seg_start,
                                                          const uint32_t *__restrict__ seg_end,
                                                          const uint32_t *__restrict__ blk_first, uint32_t nparents,
                                                          int parent_shift, uint32_t *__restrict__ gcursor,
                                                          uint64_t *__restrict__ rec_out, int recL) {
  extern __shared__ __align__(16) unsigned char sp_raw[];
  uint64_t *stage = (uint64_t *)sp_raw;
  uint32_t *h = (uint32_t *)(stage + SP_CHUNK);
  uint32_t *lstart = h + SP_MAXBINS;
  uint32_t *misc = lstart + SP_MAXBINS;    // [0..7] scan scratch, [8] parent, [9] chunk
  for (uint32_t i = threadIdx.x; i < nbins; i += 256) h[i] = 0;
  uint32_t rk[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) rk[i] = 0;
  uint64_t binbase = 0;                     // first global bin this block can see
  // ---- pass 0: rank of every record inside its bin (block-local) ----
  uint64_t p0 = 0, mw = ~0ull, mw1 = ~0ull, hi = 0, lo = 0;
  uint32_t rbeg = 0, rcnt = 0;
  if (FROM_TEXT) {
    __syncthreads();
    const uint64_t t = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    p0 = t << 5;
    if (p0 < n) {
      mw = T.maskwin(p0);
      if (recL) mw1 = T.maskwin(p0 + 64);
      if (BITS == 2) { hi = __ldg(T.w + t); lo = __ldg(T.w + t + 1); }
    }
    const uint64_t plmask = (1ull << cp.pl) - 1ull;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      if (p0 < n && ((mw >> i) & plmask) == 0) {
        uint64_t win;
        if (BITS == 2) win = i ? (hi << (2 * i)) | (lo >> (64 - 2 * i)) : hi;
        else win = T.window(p0 + i);
        const uint64_t code = prefix_value<BITS>(win, cp.pl, cp);
        if (code >= mincode && code <= maxcode) {
          const uint32_t r = atomicAdd(&h[(uint32_t)((code - mincode) >> shift)], 1u);
          rk[i >> 1] |= r << ((i & 1) * 16);
        }
      }
    }
  } else {
    if (threadIdx.x == 0) {
      uint32_t lo2 = 0, hi2 = nparents;     // blk_first[lo2] <= blockIdx.x < blk_first[hi2]
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
    rbeg = sbeg + misc[9] * SP_CHUNK;
    rcnt = send - rbeg < SP_CHUNK ? send - rbeg : SP_CHUNK;
    binbase = (uint64_t)parent << (parent_shift - shift);
#pragma unroll
    for (int g8 = 0; g8 < 4; ++g8) {           // eight loads in flight per thread, not thirty-two
      uint64_t r8[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint32_t j = (g8 * 8 + k) * 256 + threadIdx.x;
        r8[k] = j < rcnt ? rec_in[rbeg + j] : 0;
      }
      asm volatile("" ::: "memory");
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int i = g8 * 8 + k;
        const uint32_t j = i * 256 + threadIdx.x;
        if (j < rcnt) {
          const uint32_t bin = (uint32_t)(((rec_code(r8[k], recL) - mincode) >> shift) - binbase);
          const uint32_t q = atomicAdd(&h[bin], 1u);
          rk[i >> 1] |= q << ((i & 1) * 16);
        }
      }
    }
  }
  __syncthreads();
  // ---- local bin starts, one global reservation per non-empty bin ----
  const uint32_t total = sp_scan_bins(h, lstart, nbins, misc);
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < nbins; i += 256) {
    const uint32_t cnt = h[i];
    // h[] becomes (global offset of the bin's run) - (local start): global index = h[bin] + local index
    h[i] = cnt ? atomicAdd(&gcursor[binbase + i], cnt) - lstart[i] : 0u;
  }
  __syncthreads();
  // ---- pass 1: stage the records bin by bin ----
  if (FROM_TEXT) {
    const uint64_t plmask = (1ull << cp.pl) - 1ull;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      if (p0 < n && ((mw >> i) & plmask) == 0) {
        uint64_t win;
        if (BITS == 2) win = i ? (hi << (2 * i)) | (lo >> (64 - 2 * i)) : hi;
        else win = T.window(p0 + i);
        const uint64_t code = prefix_value<BITS>(win, cp.pl, cp);
        if (code >= mincode && code <= maxcode) {
          const uint32_t bin = (uint32_t)((code - mincode) >> shift);
          const uint32_t r = (rk[i >> 1] >> ((i & 1) * 16)) & 0xffffu;
          uint64_t rec = (code << 32) | (uint64_t)(uint32_t)(p0 + i);
          if (recL) {
            const uint64_t m = i ? (mw >> i) | (mw1 << (64 - i)) : mw;      // specials ahead of p0+i
            const unsigned L = m ? (unsigned)(__ffsll((long long)m) - 1) : 63u;
            rec |= (uint64_t)(L < 63u ? L : 63u) << 58;
          }
          stage[lstart[bin] + r] = rec;
        }
      }
    }
  } else {
#pragma unroll
    for (int g8 = 0; g8 < 4; ++g8) {
      uint64_t r8[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint32_t j = (g8 * 8 + k) * 256 + threadIdx.x;
        r8[k] = j < rcnt ? __ldcs(rec_in + rbeg + j) : 0;   // second and last read of the chunk (L2 hit)
      }
      asm volatile("" ::: "memory");
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int i = g8 * 8 + k;
        const uint32_t j = i * 256 + threadIdx.x;
        if (j < rcnt) {
          const uint32_t bin = (uint32_t)(((rec_code(r8[k], recL) - mincode) >> shift) - binbase);
          const uint32_t q = (rk[i >> 1] >> ((i & 1) * 16)) & 0xffffu;
          stage[lstart[bin] + q] = r8[k];
        }
      }
    }
  }
  __syncthreads();
  // ---- write-out: consecutive threads, consecutive addresses inside each bin's run ----
  for (uint32_t i = threadIdx.x; i < total; i += 256) {
    const uint64_t r = stage[i];
    const uint32_t bin = (uint32_t)(((rec_code(r, recL) - mincode) >> shift) - binbase);
    __stcs(rec_out + h[bin] + i, r);
  }
}

// gcursor[g] = record offset of bin g = suffix-array offset of its first bucket (part-local)
__global__ void k_bins_init(const uint32_t *__restrict__ lb, uint64_t mincode, int shift, uint32_t nbins,
                            uint32_t partoffset, uint32_t *gcursor, uint32_t *binstart) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < nbins) {
    const uint32_t v = lb[mincode + ((uint64_t)j << shift)] - partoffset;
    gcursor[j] = v;
    if (binstart) binstart[j] = v;
  }
}

// blk_first[0..nparents]: prefix sums of ceil(records/8192) per level-1 bin (single block)
__global__ void __launch_bounds__(256) k_chunk_table(const uint32_t *__restrict__ seg_start,
                                                     const uint32_t *__restrict__ seg_end, uint32_t nparents,
                                                     uint32_t *blk_first) {
  __shared__ uint32_t sm[8];
  uint32_t carry = 0;
  for (uint32_t base = 0; base < nparents; base += 256) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t v = i < nparents ? (seg_end[i] - seg_start[i] + SP_CHUNK - 1) / SP_CHUNK : 0;
    uint32_t tot;
    const uint32_t ex = block_excl_scan256(v, &tot, sm);
    if (i < nparents) blk_first[i] = carry + ex;
    carry += tot;
  }
  if (threadIdx.x == 0) blk_first[nparents] = carry;
}

// ---------------------------------------------------------------------------------------------
#define TILE_CAP 6144          // suffix-array slots (regular + special-padded) a tile may span
#define TILE_MAXCODES 1024     // buckets per tile
#define TS_THREADS 512
#define TS_WARPS (TS_THREADS / 32)
#define TS_BATCH 4             // records per thread in flight while their text windows are gathered
// dynamic shared memory: key[TILE_CAP] u64 | pos[TILE_CAP] u32 | lb[TILE_MAXCODES+1] | cur[TILE_MAXCODES]
#define TS_SMEM (TILE_CAP * 12 + (2 * TILE_MAXCODES + 1) * 4)

template <int BITS>
__global__ void __launch_bounds__(TS_THREADS, 2) k_tile_sort(Text<BITS> T, PartOut out, const uint32_t *__restrict__ lb,
                                                             uint32_t *__restrict__ cursor, const uint64_t *__restrict__ rec,
                                                             const uint32_t *__restrict__ gcursor, uint64_t mincode,
                                                             uint64_t maxcode, int shift, uint32_t nbins, unsigned pl, int recL,
                                                             uint32_t *__restrict__ fbbits, uint64_t tieoff) {
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
        const uint64_t r = __ldcs(rec + rs + i);
        const uint32_t slot = atomicAdd(&s_cur[(uint32_t)(rec_code(r, recL) - c0)], 1u);
        out.sa[rs + slot] = (uint32_t)r;
      }
      __syncthreads();
      for (uint32_t i = threadIdx.x; i < ncf; i += TS_THREADS) cursor[c0 + i] = base + s_cur[i];
      if (threadIdx.x == 0) atomicAdd(&out.stats->tile_leftover, 1u);
      continue;
    }
    // Insertion and key gather in one sweep: every thread keeps TS_BATCH records in flight, so the text windows
    // of the whole tile are fetched with full memory-level parallelism; the warp sorters below read only
    // shared memory.
    for (uint32_t i0 = threadIdx.x; i0 < cnt; i0 += TS_THREADS * TS_BATCH) {
      uint64_t r[TS_BATCH], key[TS_BATCH];
#pragma unroll
      for (int u = 0; u < TS_BATCH; ++u) {
        const uint32_t i = i0 + u * TS_THREADS;
        r[u] = i < cnt ? __ldcs(rec + rs + i) : 0ull;
      }
#pragma unroll
      for (int u = 0; u < TS_BATCH; ++u) {
        const uint32_t i = i0 + u * TS_THREADS;
        key[u] = 0;
        if (i < cnt) {
          // buckets that cannot be finished here (more slots than a warp sorts) are only inserted: no gather
          const uint32_t f = (uint32_t)(rec_code(r[u], recL) - c0);
          if (BITS == 2 || s_lb[f + 1] - s_lb[f] <= SORT_MAXBUCKET) {   // DNA buckets are small by choice of pl: always keyed
            const uint64_t q = (uint64_t)(uint32_t)r[u] + pl;
            key[u] = recL ? make_key_known<BITS>(T, q, (unsigned)(r[u] >> 58) - pl) : make_key<BITS>(T, q);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < TS_BATCH; ++u) {
        const uint32_t i = i0 + u * TS_THREADS;
        if (i < cnt) {
          const uint32_t slot = atomicAdd(&s_cur[(uint32_t)(rec_code(r[u], recL) - c0)], 1u);
          s_pos[slot] = (uint32_t)r[u];
          s_key[slot] = key[u];
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
        const bool keyed = BITS == 2 || s_lb[f + 1] - ls <= SORT_MAXBUCKET;   // otherwise the insertion sweep skipped the keys
        if (keyed && cf <= 32) clean = warp_sort_any<BITS, 1>(T, out, src, nullptr, nullptr, keys, start, cf, pl, lane, nullptr, nullptr, sum, mx);
        else if (keyed && cf <= 64) clean = warp_sort_any<BITS, 2>(T, out, src, nullptr, nullptr, keys, start, cf, pl, lane, nullptr, nullptr, sum, mx);
        else if (keyed && cf <= 96) clean = warp_sort_any<BITS, 3>(T, out, src, nullptr, nullptr, keys, start, cf, pl, lane, nullptr, nullptr, sum, mx);
        else if (keyed && cf <= SORT_MAXBUCKET) clean = warp_sort_any<BITS, 4>(T, out, src, nullptr, nullptr, keys, start, cf, pl, lane, nullptr, nullptr, sum, mx);
        else {                                          // too large for a warp: leave it to the global sorters
          for (uint32_t i = lane; i < cf; i += 32) out.sa[start + i] = src[i];
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
// Sparse variant of the text-level partition: when the part covers only a fraction of the codes
// (one of P parts, one of N GPUs, one slice of the pipelined fetch) most positions are rejected,
// and a block of 8192 positions would reserve almost one global range per record.  Here a block
// scans `reps` x 8192 positions so that it still collects about 8192 records: pass 0 only counts,
// pass 1 recomputes the codes and takes its local rank from a second shared-memory cursor.  If a
// block nevertheless collects more than the staging buffer holds, it writes its records straight
// to their reserved ranges (correct, just not coalesced).
// dynamic shared memory: stage[SP_CHUNK] u64 | h[SP_MAXBINS] | lstart[SP_MAXBINS] | misc[16] | cur[SP_MAXBINS]
#define SP_SMEM_SPARSE (SP_SMEM + SP_MAXBINS * 4)

// bit i of the result = position p0+i starts a regular suffix whose code lies in [mincode,maxcode]
template <int BITS>
__device__ __forceinline__ uint32_t sp_select(const Text<BITS> &T, const CodeParams &cp, uint64_t p0, uint64_t mw,
                                              uint64_t hi, uint64_t lo, uint64_t plmask, uint64_t mincode, uint64_t maxcode) {
  uint32_t sel = 0;
#pragma unroll 8
  for (int i = 0; i < 32; ++i) {
    if (((mw >> i) & plmask) != 0) continue;
    uint64_t win;
    if (BITS == 2) win = i ? (hi << (2 * i)) | (lo >> (64 - 2 * i)) : hi;
    else win = T.window(p0 + i);
    const uint64_t code = prefix_value<BITS>(win, cp.pl, cp);
    if (code >= mincode && code <= maxcode) sel |= 1u << i;
  }
  return sel;
}

template <int BITS>
__global__ void __launch_bounds__(256) k_stage_partition_sparse(Text<BITS> T, CodeParams cp, uint64_t n, uint64_t mincode,
                                                                 uint64_t maxcode, int shift, uint32_t nbins, int reps,
                                                                 uint32_t *__restrict__ gcursor,
                                                                 uint64_t *__restrict__ rec_out, int recL) {
  extern __shared__ __align__(16) unsigned char sp_raw[];
  uint64_t *stage = (uint64_t *)sp_raw;
  uint32_t *h = (uint32_t *)(stage + SP_CHUNK);
  uint32_t *lstart = h + SP_MAXBINS;
  uint32_t *misc = lstart + SP_MAXBINS;
  uint32_t *cur = misc + 16;
  for (uint32_t i = threadIdx.x; i < nbins; i += 256) h[i] = 0;
  __syncthreads();
  const uint64_t plmask = (1ull << cp.pl) - 1ull;
  const uint64_t tbase = ((uint64_t)blockIdx.x * reps) * 256 + threadIdx.x;   // word index of repetition 0
  uint32_t selbits[8];                       // per repetition: which of the thread's 32 positions are in range
#pragma unroll
  for (int rep = 0; rep < 8; ++rep) {
    selbits[rep] = 0;
    if (rep < reps) {
      const uint64_t t = tbase + (uint64_t)rep * 256;
      const uint64_t p0 = t << 5;
      if (p0 < n) {
        const uint64_t mw = T.maskwin(p0);
        uint64_t hi = 0, lo = 0;
        if (BITS == 2) { hi = __ldg(T.w + t); lo = __ldg(T.w + t + 1); }
        uint32_t sel = sp_select<BITS>(T, cp, p0, mw, hi, lo, plmask, mincode, maxcode);
        selbits[rep] = sel;
        // only the selected positions (about 32/reps of them) reach the shared-memory atomics
        while (sel) {
          const int i = __ffs(sel) - 1;
          sel &= sel - 1;
          uint64_t win;
          if (BITS == 2) win = i ? (hi << (2 * i)) | (lo >> (64 - 2 * i)) : hi;
          else win = T.window(p0 + i);
          const uint64_t code = prefix_value<BITS>(win, cp.pl, cp);
          atomicAdd(&h[(uint32_t)((code - mincode) >> shift)], 1u);
        }
      }
    }
  }
  __syncthreads();
  const uint32_t total = sp_scan_bins(h, lstart, nbins, misc);
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < nbins; i += 256) {
    const uint32_t cnt = h[i];
    cur[i] = lstart[i];
    h[i] = cnt ? atomicAdd(&gcursor[i], cnt) - lstart[i] : 0u;   // global index = h[bin] + local index
  }
  __syncthreads();
  const bool staged = total <= SP_CHUNK;
#pragma unroll
  for (int rep = 0; rep < 8; ++rep) {
    uint32_t sel = selbits[rep];
    if (sel) {
      const uint64_t t = tbase + (uint64_t)rep * 256;
      const uint64_t p0 = t << 5;
      const uint64_t mw = T.maskwin(p0);
      const uint64_t mw1 = recL ? T.maskwin(p0 + 64) : ~0ull;
      uint64_t hi = 0, lo = 0;
      if (BITS == 2) { hi = __ldg(T.w + t); lo = __ldg(T.w + t + 1); }
      while (sel) {
        const int i = __ffs(sel) - 1;
        sel &= sel - 1;
        uint64_t win;
        if (BITS == 2) win = i ? (hi << (2 * i)) | (lo >> (64 - 2 * i)) : hi;
        else win = T.window(p0 + i);
        const uint64_t code = prefix_value<BITS>(win, cp.pl, cp);
        const uint32_t bin = (uint32_t)((code - mincode) >> shift);
        uint64_t rec = (code << 32) | (uint64_t)(uint32_t)(p0 + i);
        if (recL) {
          const uint64_t m = i ? (mw >> i) | (mw1 << (64 - i)) : mw;
          const unsigned L = m ? (unsigned)(__ffsll((long long)m) - 1) : 63u;
          rec |= (uint64_t)(L < 63u ? L : 63u) << 58;
        }
        const uint32_t r = atomicAdd(&cur[bin], 1u);
        if (staged) stage[r] = rec;
        else __stcs(rec_out + h[bin] + r, rec);
      }
    }
  }
  if (!staged) return;
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < total; i += 256) {
    const uint64_t r = stage[i];
    const uint32_t bin = (uint32_t)((rec_code(r, recL) - mincode) >> shift);
    __stcs(rec_out + h[bin] + i, r);
  }
}

// Hand-written device-wide primitives: a single-pass (decoupled look-back) exclusive scan over uint32 and a
// stable LSD radix sort of (uint64 key, uint32 value) pairs (histogram / scan / scatter per 8-bit digit).  They
// serve the bucket-table prefix sum and the refinement path, neither of which is the dominant kernel.
#pragma once
#include "common.cuh"

#define SC_THREADS 256

// exclusive scan of one value per thread across a 256-thread block; *total = block sum.
// sm must hold 8 words; contains two __syncthreads, safe to call repeatedly.
__device__ __forceinline__ uint32_t block_excl_scan256(uint32_t v, uint32_t *total, uint32_t *sm) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  uint32_t x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(GTB_FULL, x, o);
    if (lane >= o) x += y;
  }
  __syncthreads();            // previous users of sm are done
  if (lane == 31) sm[wid] = x;
  __syncthreads();
  uint32_t woff = 0, tot = 0;
#pragma unroll
  for (int k = 0; k < SC_THREADS / 32; ++k) {
    const uint32_t s = sm[k];
    if (k < wid) woff += s;
    tot += s;
  }
  *total = tot;
  return woff + x - v;
}

// ---------------------------------------------------------------------------------------------
// Single-pass exclusive scan (decoupled look-back): one launch instead of three.  Every block takes a
// ticket (tiles are therefore processed in dispatch order and a block only ever waits for blocks that
// already run), scans its 256*R values, publishes its aggregate, and warp 0 walks back over the
// predecessors' published words until it meets an inclusive prefix.
//   state word = epoch << 34 | flag << 32 | value;  flag 1 = tile aggregate, 2 = inclusive prefix.
// The epoch changes with every launch, so words of earlier scans read as "not ready" and the array
// never has to be cleared; the last ticket holder rewinds the ticket counter.
// Optionally also reduces max(in[]) into *maxout (atomicMax; caller zeroes it).
// ---------------------------------------------------------------------------------------------
#define SCAN_FLAG_AGG 1ull
#define SCAN_FLAG_PFX 2ull

template <int R>
__global__ void __launch_bounds__(SC_THREADS) k_scan_onepass(const uint32_t *in, uint32_t *out, uint64_t n, uint32_t ntiles,
                                                             unsigned long long *__restrict__ state,
                                                             unsigned int *__restrict__ ticket, uint64_t epoch,
                                                             int append_total, unsigned long long *__restrict__ total_out,
                                                             unsigned int *__restrict__ maxout) {
  __shared__ uint32_t sm[8];
  __shared__ uint32_t s_tile, s_excl;
  if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
  __syncthreads();
  const uint32_t tile = s_tile;
  const uint64_t base = (uint64_t)tile * (SC_THREADS * R) + (uint64_t)threadIdx.x * R;
  uint32_t v[R];
  if (base + R <= n) {
    if (R % 4 == 0 && (((uintptr_t)in) & 15) == 0) {
#pragma unroll
      for (int j = 0; j < R; j += 4) {
        const uint4 q = *reinterpret_cast<const uint4 *>(in + base + j);
        v[j] = q.x; v[j + 1] = q.y; v[j + 2] = q.z; v[j + 3] = q.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < R; ++j) v[j] = in[base + j];
    }
  } else {
#pragma unroll
    for (int j = 0; j < R; ++j) v[j] = base + j < n ? in[base + j] : 0u;
  }
  uint32_t tsum = 0, tmax = 0;
#pragma unroll
  for (int j = 0; j < R; ++j) {
    tsum += v[j];
    tmax = max(tmax, v[j]);
  }
  uint32_t tot;
  const uint32_t ex = block_excl_scan256(tsum, &tot, sm);
  if (maxout) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tmax = max(tmax, __shfl_xor_sync(GTB_FULL, tmax, o));
    if ((threadIdx.x & 31) == 0 && tmax) atomicMax(maxout, tmax);
  }
  const unsigned long long tag = (unsigned long long)epoch << 34;
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    if (tile == 0) {
      if (lane == 0) {
        atomicExch(&state[0], tag | (SCAN_FLAG_PFX << 32) | tot);
        s_excl = 0;
      }
    } else {
      if (lane == 0) atomicExch(&state[tile], tag | (SCAN_FLAG_AGG << 32) | tot);
      uint32_t excl = 0;
      int64_t look = (int64_t)tile - 1;
      while (true) {
        const int64_t idx = look - lane;
        unsigned long long st = tag | (SCAN_FLAG_PFX << 32);      // before tile 0: an inclusive prefix of 0
        if (idx >= 0) {
          do {
            st = *((volatile unsigned long long *)&state[idx]);
          } while ((st >> 34) != epoch || ((st >> 32) & 3ull) == 0);
        }
        const unsigned pmask = __ballot_sync(GTB_FULL, ((st >> 32) & 3ull) == SCAN_FLAG_PFX);
        const int firstp = pmask ? __ffs(pmask) - 1 : 31;
        uint32_t val = lane <= firstp ? (uint32_t)st : 0u;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(GTB_FULL, val, o);
        excl += val;
        if (pmask) break;
        look -= 32;
      }
      if (lane == 0) {
        atomicExch(&state[tile], tag | (SCAN_FLAG_PFX << 32) | (uint32_t)(excl + tot));
        s_excl = excl;
      }
    }
  }
  __syncthreads();
  uint32_t run = s_excl + ex;
  if (base + R <= n) {
    if (R % 4 == 0 && (((uintptr_t)out) & 15) == 0) {
#pragma unroll
      for (int j = 0; j < R; j += 4) {
        uint4 q;
        q.x = run; run += v[j];
        q.y = run; run += v[j + 1];
        q.z = run; run += v[j + 2];
        q.w = run; run += v[j + 3];
        *reinterpret_cast<uint4 *>(out + base + j) = q;
      }
    } else {
#pragma unroll
      for (int j = 0; j < R; ++j) { out[base + j] = run; run += v[j]; }
    }
  } else {
#pragma unroll
    for (int j = 0; j < R; ++j) {
      if (base + j < n) { out[base + j] = run; run += v[j]; }
    }
  }
  if (tile == ntiles - 1 && threadIdx.x == 0) {
    const unsigned long long total = (unsigned long long)s_excl + tot;     // sums stay below 2^32 by construction
    if (append_total) out[n] = (uint32_t)total;
    if (total_out) *total_out = total;
    *ticket = 0;                                   // every ticket of this launch has been handed out
  }
}

// ---------------------------------------------------------------------------------------------
// Stable LSD radix sort pass on 8 bits.
// ---------------------------------------------------------------------------------------------
#define RS_THREADS 256
#define RS_WARPS (RS_THREADS / 32)
#define RS_ITEMS 8                          // rounds of 32 consecutive elements per warp
#define RS_TILE (RS_THREADS * RS_ITEMS)

__global__ void __launch_bounds__(RS_THREADS) k_radix_hist(const uint64_t *__restrict__ keys, uint32_t n, int shift,
                                                           uint32_t *__restrict__ blockhist, uint32_t nblocks) {
  __shared__ uint32_t h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t base = blockIdx.x * RS_TILE;
#pragma unroll
  for (int r = 0; r < RS_ITEMS; ++r) {
    const uint32_t j = base + r * RS_THREADS + threadIdx.x;
    if (j < n) atomicAdd(&h[(unsigned)(keys[j] >> shift) & 255u], 1u);
  }
  __syncthreads();
  blockhist[(uint64_t)threadIdx.x * nblocks + blockIdx.x] = h[threadIdx.x];
}

// offsets = exclusive scan of blockhist laid out [digit][block]
__global__ void __launch_bounds__(RS_THREADS) k_radix_scatter(const uint64_t *__restrict__ keys_in,
                                                              const uint32_t *__restrict__ vals_in,
                                                              uint64_t *__restrict__ keys_out,
                                                              uint32_t *__restrict__ vals_out, uint32_t n, int shift,
                                                              const uint32_t *__restrict__ offsets, uint32_t nblocks) {
  __shared__ uint32_t wc[RS_WARPS][256];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int k = threadIdx.x; k < RS_WARPS * 256; k += RS_THREADS) (&wc[0][0])[k] = 0;
  __syncthreads();
  const uint32_t wbase = blockIdx.x * RS_TILE + wid * (32 * RS_ITEMS);
  uint64_t key[RS_ITEMS];
  uint32_t val[RS_ITEMS], rank[RS_ITEMS];
#pragma unroll
  for (int r = 0; r < RS_ITEMS; ++r) {
    const uint32_t j = wbase + r * 32 + lane;
    const bool valid = j < n;
    key[r] = valid ? keys_in[j] : 0;
    val[r] = valid ? vals_in[j] : 0;
    const unsigned d = (unsigned)(key[r] >> shift) & 255u;
    const unsigned tag = valid ? d : (256u + lane);
    const unsigned peers = __match_any_sync(GTB_FULL, tag);
    const unsigned before = __popc(peers & ((1u << lane) - 1u));
    uint32_t old = 0;
    if (valid) old = wc[wid][d];
    __syncwarp();
    if (valid && before == 0) wc[wid][d] = old + __popc(peers);
    __syncwarp();
    rank[r] = old + before;
  }
  __syncthreads();
  {
    const unsigned d = threadIdx.x;   // one digit per thread
    uint32_t run = offsets[(uint64_t)d * nblocks + blockIdx.x];
#pragma unroll
    for (int w2 = 0; w2 < RS_WARPS; ++w2) {
      const uint32_t t = wc[w2][d];
      wc[w2][d] = run;
      run += t;
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < RS_ITEMS; ++r) {
    const uint32_t j = wbase + r * 32 + lane;
    if (j < n) {
      const unsigned d = (unsigned)(key[r] >> shift) & 255u;
      const uint32_t o = wc[wid][d] + rank[r];
      keys_out[o] = key[r];
      vals_out[o] = val[r];
    }
  }
}

// Hand-written device-wide primitives: exclusive scan over uint32 and a stable LSD radix sort of
// (uint64 key, uint32 value) pairs.  Both are three-kernel reduce/scan/apply designs; they serve the
// bucket-table prefix sum and the refinement path, neither of which is the dominant kernel.
#pragma once
#include "common.cuh"

#define SC_THREADS 256
#define SC_ROUNDS 16
#define SC_TILE (SC_THREADS * SC_ROUNDS)

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

// rounds: tile = 256 * rounds counters per block (16 for large tables; 2 for small ones, where sixteen
// dependent block scans per block would leave the GPU idle)
__global__ void __launch_bounds__(SC_THREADS) k_scan_tilesums(const uint32_t *__restrict__ in, uint64_t n,
                                                              uint32_t *__restrict__ tilesums, int rounds) {
  __shared__ uint32_t sm[8];
  const uint64_t base = (uint64_t)blockIdx.x * SC_THREADS * rounds;
  uint32_t s = 0;
#pragma unroll 4
  for (int r = 0; r < rounds; ++r) {
    const uint64_t i = base + (uint64_t)r * SC_THREADS + threadIdx.x;
    if (i < n) s += in[i];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(GTB_FULL, s, o);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int k = 0; k < 8; ++k) t += sm[k];
    tilesums[blockIdx.x] = t;
  }
}

// single block: in-place exclusive scan of the tile sums, total written to *total_out (64-bit)
__global__ void __launch_bounds__(SC_THREADS) k_scan_spine(uint32_t *__restrict__ tilesums, uint32_t ntiles,
                                                           unsigned long long *__restrict__ total_out) {
  __shared__ uint32_t sm[8];
  unsigned long long carry = 0;
  for (uint32_t base = 0; base < ntiles; base += SC_THREADS) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t v = i < ntiles ? tilesums[i] : 0;
    uint32_t tot;
    const uint32_t ex = block_excl_scan256(v, &tot, sm);
    if (i < ntiles) tilesums[i] = (uint32_t)(carry + ex);
    carry += tot;
  }
  if (threadIdx.x == 0 && total_out) *total_out = carry;
}

// out[i] = sum_{j<i} in[j] (mod 2^32); out may alias in.  If n_out == n+1 the grand total is
// appended as out[n] (the bucket table's last entry).
__global__ void __launch_bounds__(SC_THREADS) k_scan_apply(const uint32_t *in, uint32_t *out, uint64_t n,
                                                           const uint32_t *__restrict__ tilesums,
                                                           int append_total, int rounds) {
  __shared__ uint32_t sm[8];
  const uint64_t base = (uint64_t)blockIdx.x * SC_THREADS * rounds;
  uint32_t carry = tilesums[blockIdx.x];
#pragma unroll 1
  for (int r = 0; r < rounds; ++r) {
    const uint64_t i = base + (uint64_t)r * SC_THREADS + threadIdx.x;
    const uint32_t v = i < n ? in[i] : 0;
    uint32_t tot;
    const uint32_t ex = block_excl_scan256(v, &tot, sm);
    if (i < n) out[i] = carry + ex;
    if (append_total && i == n - 1) out[n] = carry + ex + v;
    carry += tot;
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

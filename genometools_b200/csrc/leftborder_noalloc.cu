// gtb200_leftborder_partialsums_noalloc: the bucket-table prefix sum for callers that must not
// allocate.
//
// Who needs this: a host program that has handed the whole HBM of a device to its own memory
// pool, so that a cudaMalloc issued by a library can fail or fragment that pool; and a program
// that sits in the middle of a running pipeline, where the implicit device-wide synchronisation
// of cudaMalloc/cudaFree would stall every other stream.  The regular path (gtb200_count) sizes
// and allocates its tile-sum scratch per context; this entry point instead keeps that scratch in
// one fixed array that is part of the library image, so it never asks the driver for memory.
//
// Consequences the caller has to know:
//   * the array is GTB200_NOALLOC_MAX_CODES/4096+2 words (about 64 KB) and every GPU that loads
//     this library carries it until the process exits;
//   * table sizes above GTB200_NOALLOC_MAX_CODES are refused with an error, nothing is computed;
//   * the first call on each device looks up the array's device address (one
//     cudaGetSymbolAddress, no allocation); make that first call before any stream capture starts;
//   * all calls on one device share the array and it still holds what the previous call left
//     there.  The words that must start from a known value are rewritten at the beginning of
//     every call, but two calls whose kernels overlap in time on the same GPU (different
//     streams, different host threads) overwrite each other's partial sums: serialise them.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <mutex>

#include "../../include/gtsfx_b200.h"

#define NA_THREADS 256
#define NA_ROUNDS 16
#define NA_TILE (NA_THREADS * NA_ROUNDS)
#define NA_MAX_TILES ((GTB200_NOALLOC_MAX_CODES + NA_TILE - 1) / NA_TILE)
#define NA_MAX_DEVICES 64

// word 0: grand total of the last call; words 1..: one partial sum per tile of 4096 counters
static __device__ __align__(256) uint32_t g_na_scratch[NA_MAX_TILES + 2];

// In host code `g_na_scratch` is only a handle for the runtime, not a device address: the pointer
// the kernels use is obtained once per device with cudaGetSymbolAddress and cached here.
static uint32_t *g_na_ptr[NA_MAX_DEVICES];
static std::mutex g_na_mutex;

static __device__ __forceinline__ uint32_t na_block_excl_scan(uint32_t v, uint32_t *total, uint32_t *sm) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  uint32_t x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  __syncthreads();
  if (lane == 31) sm[wid] = x;
  __syncthreads();
  uint32_t woff = 0, tot = 0;
#pragma unroll
  for (int k = 0; k < NA_THREADS / 32; ++k) {
    const uint32_t s = sm[k];
    if (k < wid) woff += s;
    tot += s;
  }
  *total = tot;
  return woff + x - v;
}

static __global__ void __launch_bounds__(NA_THREADS) k_na_tilesums(const uint32_t *__restrict__ counts, uint64_t n,
                                                                  uint32_t *__restrict__ scratch) {
  __shared__ uint32_t sm[8];
  const uint64_t base = (uint64_t)blockIdx.x * NA_TILE;
  uint32_t s = 0;
  for (int r = 0; r < NA_ROUNDS; ++r) {
    const uint64_t i = base + (uint64_t)r * NA_THREADS + threadIdx.x;
    if (i < n) s += counts[i];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int k = 0; k < 8; ++k) t += sm[k];
    scratch[1 + blockIdx.x] = t;
  }
}

static __global__ void __launch_bounds__(NA_THREADS) k_na_spine(uint32_t *__restrict__ scratch, uint32_t ntiles) {
  __shared__ uint32_t sm[8];
  uint32_t carry = 0;
  for (uint32_t base = 0; base < ntiles; base += NA_THREADS) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t v = i < ntiles ? scratch[1 + i] : 0;
    uint32_t tot;
    const uint32_t ex = na_block_excl_scan(v, &tot, sm);
    if (i < ntiles) scratch[1 + i] = carry + ex;
    carry += tot;
  }
  if (threadIdx.x == 0) scratch[0] = carry;
}

static __global__ void __launch_bounds__(NA_THREADS) k_na_apply(const uint32_t *counts, uint32_t *leftborder, uint64_t n,
                                                               const uint32_t *__restrict__ scratch) {
  __shared__ uint32_t sm[8];
  const uint64_t base = (uint64_t)blockIdx.x * NA_TILE;
  uint32_t carry = scratch[1 + blockIdx.x];
  for (int r = 0; r < NA_ROUNDS; ++r) {
    const uint64_t i = base + (uint64_t)r * NA_THREADS + threadIdx.x;
    const uint32_t v = i < n ? counts[i] : 0;
    uint32_t tot;
    const uint32_t ex = na_block_excl_scan(v, &tot, sm);
    if (i < n) leftborder[i] = carry + ex;
    if (i == n - 1) leftborder[n] = carry + ex + v;
    carry += tot;
  }
}

static int na_fail(char *errbuf, size_t errbuf_len, const char *what, cudaError_t e) {
  if (errbuf && errbuf_len) {
    if (e == cudaSuccess) snprintf(errbuf, errbuf_len, "gtb200_leftborder_partialsums_noalloc: %s", what);
    else snprintf(errbuf, errbuf_len, "gtb200_leftborder_partialsums_noalloc: %s: %s", what, cudaGetErrorString(e));
  }
  return -1;
}

extern "C" int gtb200_leftborder_partialsums_noalloc(const uint32_t *counts_dev, uint32_t *leftborder_dev,
                                                     uint64_t numofallcodes, void *stream, char *errbuf,
                                                     size_t errbuf_len) {
  if (!counts_dev || !leftborder_dev) return na_fail(errbuf, errbuf_len, "null device pointer", cudaSuccess);
  if (numofallcodes == 0) return na_fail(errbuf, errbuf_len, "numofallcodes must be at least 1", cudaSuccess);
  if (numofallcodes > GTB200_NOALLOC_MAX_CODES) {
    if (errbuf && errbuf_len)
      snprintf(errbuf, errbuf_len,
               "gtb200_leftborder_partialsums_noalloc: %llu buckets exceed the fixed scratch (maximum %llu); "
               "use gtb200_count, which allocates",
               (unsigned long long)numofallcodes, (unsigned long long)GTB200_NOALLOC_MAX_CODES);
    return -1;
  }
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return na_fail(errbuf, errbuf_len, "cudaGetDevice", e);
  if (dev < 0 || dev >= NA_MAX_DEVICES) return na_fail(errbuf, errbuf_len, "device index out of range", cudaSuccess);
  uint32_t *scratch;
  {
    // set-up, once per device, with that device current: resolve the array's device address
    std::lock_guard<std::mutex> lock(g_na_mutex);
    if (!g_na_ptr[dev]) {
      uint32_t *p = nullptr;
      e = cudaGetSymbolAddress((void **)&p, g_na_scratch);
      if (e != cudaSuccess || !p) return na_fail(errbuf, errbuf_len, "cudaGetSymbolAddress(scratch)", e);
      g_na_ptr[dev] = p;
    }
    scratch = g_na_ptr[dev];
  }
  cudaStream_t s = (cudaStream_t)stream;
  const uint32_t ntiles = (uint32_t)((numofallcodes + NA_TILE - 1) / NA_TILE);
  // the array still holds the previous call's sums: reset the words this call relies on
  e = cudaMemsetAsync(scratch, 0, ((size_t)ntiles + 1) * sizeof(uint32_t), s);
  if (e != cudaSuccess) return na_fail(errbuf, errbuf_len, "clearing the scratch", e);
  k_na_tilesums<<<ntiles, NA_THREADS, 0, s>>>(counts_dev, numofallcodes, scratch);
  k_na_spine<<<1, NA_THREADS, 0, s>>>(scratch, ntiles);
  k_na_apply<<<ntiles, NA_THREADS, 0, s>>>(counts_dev, leftborder_dev, numofallcodes, scratch);
  e = cudaGetLastError();
  if (e != cudaSuccess) return na_fail(errbuf, errbuf_len, "kernel launch", e);
  return 0;
}

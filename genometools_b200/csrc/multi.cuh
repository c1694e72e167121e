// Several GPUs of one box behind one handle (single process, one host thread per device while a part is sorted).
//
//   open   : the sequence is uploaded to the first device only; its device image (packed text, special bitmap,
//            range table) is then copied to the other GPUs peer to peer over NVLink - the reference's encoded
//            sequence, broadcast once.
//   count  : every GPU counts the suffixes that start in its 1/ndev of the positions; the sigma^pl tables are summed
//            by one kernel per GPU that reads its slice of every table and writes the sums back to all of them over
//            peer memory (no host round trip, no library collective); ordering between the devices is by CUDA
//            events, never by spinning kernels.
//   part   : a part [mincode, maxcode] of the reference's -parts split is cut once more into ndev code ranges of
//            about equal suffix count; every GPU inserts, sorts and LCP-annotates its range and copies it straight
//            into its slice of the caller's host buffers, so the ranges arrive in bucket order by construction.
// Output is byte-identical for every number of devices (a shard's first LCP value comes from codes, k_part_first).
// Included at the end of gtsfx_b200.cu.
#pragma once
#include <algorithm>
#include <string>
#include <thread>

#define GTB200_MAXDEV 16

struct gtb200_multi {
  int ndev;
  gtb200_ctx *ctx[GTB200_MAXDEV];
  bool p2p;                         // every device can read and write every other device's memory
  bool counted;
  std::vector<uint32_t> h_lb;       // host copy of leftborder: the shard map of every part
  cudaEvent_t ev_a[GTB200_MAXDEV], ev_b[GTB200_MAXDEV];
  uint64_t peer_bytes;              // bytes that crossed NVLink so far (broadcast + table reduction)
};

struct TablePtrs { uint32_t *p[GTB200_MAXDEV]; };

// sums words [lo, hi) of ndev tables (lo a multiple of 4) and writes the sums back to all of them
__global__ void __launch_bounds__(256) k_table_allreduce(TablePtrs t, int ndev, uint64_t lo, uint64_t hi) {
  const uint64_t nvec = (hi - lo) / 4;
  for (uint64_t v = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (uint64_t)gridDim.x * blockDim.x) {
    uint4 s = make_uint4(0u, 0u, 0u, 0u);
    for (int d = 0; d < ndev; ++d) {
      const uint4 q = *reinterpret_cast<const uint4 *>(t.p[d] + lo + 4 * v);
      s.x += q.x; s.y += q.y; s.z += q.z; s.w += q.w;
    }
    for (int d = 0; d < ndev; ++d) *reinterpret_cast<uint4 *>(t.p[d] + lo + 4 * v) = s;
  }
  if (blockIdx.x == 0 && threadIdx.x < (hi - lo) % 4) {
    const uint64_t i = lo + 4 * nvec + threadIdx.x;
    uint32_t s = 0;
    for (int d = 0; d < ndev; ++d) s += t.p[d][i];
    for (int d = 0; d < ndev; ++d) t.p[d][i] = s;
  }
}

// a second context on `device` holding a copy of src's device image of the sequence
static int ctx_clone(gtb200_ctx *src, int device, gtb200_ctx **out, uint64_t *moved) {
  gtb200_ctx *c = ctx_create(device, src->cp.sigma, src->n, src->cp.pl);
  if (!c) return -1;
  c->ntextwords = src->ntextwords;
  c->nmaskwords = src->nmaskwords;
  c->nranges = src->nranges;
  c->specialchars = src->specialchars;
  c->nsort = src->nsort;
  auto fail = [&](void) { free_ctx(c); return -1; };
  const size_t tbytes = (c->ntextwords + c->nmaskwords) * 8;
  if (dmalloc(c, &c->d_text, tbytes) != cudaSuccess) { set_error("gtb200_multi_open: out of device memory on device %d", device); return fail(); }
  c->d_mask = c->d_text + c->ntextwords;
  if (dmalloc(c, &c->d_stats, sizeof(PartStats)) != cudaSuccess ||
      dmalloc(c, &c->d_ranges, (c->nranges + 1) * sizeof(gtb200_range)) != cudaSuccess ||
      dmalloc(c, &c->d_rangeoff, (c->nranges + 2) * 4) != cudaSuccess) {
    set_error("gtb200_multi_open: out of device memory on device %d", device);
    return fail();
  }
  {
    std::lock_guard<std::mutex> lock(g_pool_mu);
    if (!g_pinned_pool.empty()) { c->h_stats = (PartStats *)g_pinned_pool.back(); g_pinned_pool.pop_back(); }
  }
  if (!c->h_stats && cudaMallocHost(&c->h_stats, 256) != cudaSuccess) { set_error("gtb200_multi_open: cudaMallocHost failed"); return fail(); }
  cudaError_t e = cudaMemsetAsync(c->d_stats, 0, sizeof(PartStats), c->stream);
  if (e == cudaSuccess) e = cudaMemcpyPeerAsync(c->d_text, device, src->d_text, src->device, tbytes, c->stream);
  if (e == cudaSuccess && c->nranges)
    e = cudaMemcpyPeerAsync(c->d_ranges, device, src->d_ranges, src->device, c->nranges * sizeof(gtb200_range), c->stream);
  if (e == cudaSuccess) e = cudaMemcpyPeerAsync(c->d_rangeoff, device, src->d_rangeoff, src->device, (c->nranges + 2) * 4, c->stream);
  *moved += tbytes + c->nranges * sizeof(gtb200_range) + (c->nranges + 2) * 4;
  if (e == cudaSuccess && src->d_gather) {                 // small genomes: the sorters' gather table, rebuilt locally
    const uint64_t nrec = c->ntextwords - 1;
    if (dmalloc(c, &c->d_gather, nrec * sizeof(ulonglong2)) != cudaSuccess) { set_error("gtb200_multi_open: out of device memory"); return fail(); }
    if (c->bits == 2) k_build_gather<2><<<grid_for(nrec, 256), 256, 0, c->stream>>>(c->d_text, c->d_mask, nrec, c->d_gather);
    else k_build_gather<8><<<grid_for(nrec, 256), 256, 0, c->stream>>>(c->d_text, c->d_mask, nrec, c->d_gather);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) { set_error("gtb200_multi_open: copy to device %d failed: %s", device, cudaGetErrorString(e)); return fail(); }
  *out = c;
  return 0;
}

extern "C" void gtb200_multi_close(gtb200_multi *m) {
  if (!m) return;
  for (int d = 0; d < m->ndev; ++d) {
    if (m->ctx[d]) {
      cudaSetDevice(m->ctx[d]->device);
      if (m->ev_a[d]) cudaEventDestroy(m->ev_a[d]);
      if (m->ev_b[d]) cudaEventDestroy(m->ev_b[d]);
      free_ctx(m->ctx[d]);
    }
  }
  delete m;
}

extern "C" int gtb200_multi_open(gtb200_multi **out, const int *devices, int ndev, const gtb200_seqdesc *seq,
                                 unsigned prefixlength) {
  if (!out || !devices || !seq) return set_error("gtb200_multi_open: null argument");
  *out = nullptr;
  if (ndev < 1 || ndev > GTB200_MAXDEV) return set_error("gtb200_multi_open: %d devices (1..%d supported)", ndev, GTB200_MAXDEV);
  for (int a = 0; a < ndev; ++a)
    for (int b = a + 1; b < ndev; ++b)
      if (devices[a] == devices[b]) return set_error("gtb200_multi_open: device %d listed twice", devices[a]);
  gtb200_multi *m = new gtb200_multi();
  m->ndev = ndev;
  if (gtb200_open(&m->ctx[0], devices[0], seq, prefixlength)) { delete m; return -1; }
  // peer access for every ordered pair; without it the devices still work, each on its own copy (staged copies,
  // replicated counting)
  m->p2p = ndev > 1;
  for (int a = 0; a < ndev && ndev > 1; ++a) {
    cudaSetDevice(devices[a]);
    for (int b = 0; b < ndev; ++b) {
      if (a == b) continue;
      int can = 0;
      if (cudaDeviceCanAccessPeer(&can, devices[a], devices[b]) != cudaSuccess || !can) { m->p2p = false; continue; }
      const cudaError_t e = cudaDeviceEnablePeerAccess(devices[b], 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) m->p2p = false;
      (void)cudaGetLastError();
    }
  }
  {
    // the copies to the other GPUs run concurrently (context creation and the first allocations on a fresh device
    // take longer than the 1.2 GB copy itself)
    std::vector<std::thread> workers;
    int rcs[GTB200_MAXDEV] = {0};
    uint64_t moved[GTB200_MAXDEV] = {0};
    std::string msg[GTB200_MAXDEV];
    for (int d = 1; d < ndev; ++d)
      workers.emplace_back([&, d]() {
        rcs[d] = ctx_clone(m->ctx[0], devices[d], &m->ctx[d], &moved[d]);
        if (rcs[d]) msg[d] = gtb200_last_error();
      });
    for (auto &w : workers) w.join();
    for (int d = 1; d < ndev; ++d) {
      m->peer_bytes += moved[d];
      if (rcs[d]) {
        set_error("%s", msg[d].c_str());
        gtb200_multi_close(m);
        return -1;
      }
    }
  }
  // Several GPUs bring several PCIe links, and the host cores that widen 32-bit positions are then the narrower
  // resource (measured: 2 GPUs, 3.1 Gbp, 429 ms narrow against 401 ms on one GPU): copy the 64-bit image directly.
  if (ndev > 1 && !getenv("GTB200_NARROW"))
    for (int d = 0; d < ndev; ++d) m->ctx[d]->narrow = 0;
  for (int d = 0; d < ndev && ndev > 1; ++d) {
    cudaSetDevice(devices[d]);
    if (cudaEventCreateWithFlags(&m->ev_a[d], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&m->ev_b[d], cudaEventDisableTiming) != cudaSuccess) {
      gtb200_multi_close(m);
      return set_error("gtb200_multi_open: cudaEventCreate failed");
    }
  }
  *out = m;
  return 0;
}

extern "C" int gtb200_multi_devices(const gtb200_multi *m) { return m ? m->ndev : 0; }
extern "C" gtb200_ctx *gtb200_multi_ctx(gtb200_multi *m, int i) { return (m && i >= 0 && i < m->ndev) ? m->ctx[i] : nullptr; }
extern "C" uint64_t gtb200_multi_peer_bytes(const gtb200_multi *m) { return m ? m->peer_bytes : 0; }

static void fill_countinfo(const gtb200_ctx *c, gtb200_countinfo *info) {
  info->numofallcodes = c->numcodes;
  info->numofspecialcodes = c->numspecialcodes;
  info->numofdistpfxidx = c->numdist;
  info->numofsuffixestosort = c->nsort;
  info->specialcharacters = c->specialchars;
  info->largestbucket = c->largestbucket;
}

extern "C" int gtb200_multi_count(gtb200_multi *m, uint32_t *leftborder, uint32_t *countspecialcodes,
                                  uint32_t *distpfxidx, gtb200_countinfo *info) {
  if (!m) return set_error("gtb200_multi_count: null handle");
  gtb200_ctx *c0 = m->ctx[0];
  m->h_lb.resize(c0->numcodes + 1);
  gtb200_countinfo ci;
  if (m->ndev == 1 || !m->p2p) {
    // one device, or no peer access: every device counts the whole sequence for itself
    for (int d = m->ndev - 1; d >= 1; --d)
      if (gtb200_count(m->ctx[d], nullptr, nullptr, nullptr, nullptr)) return -1;
    if (gtb200_count(c0, m->h_lb.data(), countspecialcodes, distpfxidx, &ci)) return -1;
  } else {
    const uint64_t n = c0->n;
    const uint64_t shard = ((n / m->ndev + 63) / 64) * 64;
    const uint64_t W = (c0->numcodes + 2) + (c0->numspecialcodes + 1) + (c0->numdist + 1);
    TablePtrs tp;
    for (int d = 0; d < m->ndev; ++d) {
      gtb200_ctx *c = m->ctx[d];
      uint64_t lo = std::min(n, (uint64_t)d * shard), hi = d == m->ndev - 1 ? n : std::min(n, (uint64_t)(d + 1) * shard);
      if (lo >= hi) lo = hi = 0;                     // more devices than 64-position blocks: an empty shard
      void *tab = nullptr;
      if (gtb200_count_shard(c, lo, hi, d == 0 ? 1 : 0, &tab, nullptr)) return -1;
      tp.p[d] = (uint32_t *)tab;
      CK(cudaEventRecord(m->ev_a[d], c->stream));
    }
    const uint64_t slice = ((W / m->ndev + 3) / 4) * 4;
    for (int d = 0; d < m->ndev; ++d) {
      gtb200_ctx *c = m->ctx[d];
      CK(cudaSetDevice(c->device));
      for (int e = 0; e < m->ndev; ++e)
        if (e != d) CK(cudaStreamWaitEvent(c->stream, m->ev_a[e], 0));     // every table is complete
      const uint64_t lo = std::min(W, (uint64_t)d * slice), hi = d == m->ndev - 1 ? W : std::min(W, (uint64_t)(d + 1) * slice);
      if (hi > lo) {
        ProfScope ps(c, KC_COUNT);
        k_table_allreduce<<<148 * 4, 256, 0, c->stream>>>(tp, m->ndev, lo, hi);
        CK(cudaGetLastError());
        m->peer_bytes += (hi - lo) * 4 * 2 * (uint64_t)(m->ndev - 1);
      }
      CK(cudaEventRecord(m->ev_b[d], c->stream));
    }
    for (int d = 0; d < m->ndev; ++d) {
      gtb200_ctx *c = m->ctx[d];
      CK(cudaSetDevice(c->device));
      for (int e = 0; e < m->ndev; ++e)
        if (e != d) CK(cudaStreamWaitEvent(c->stream, m->ev_b[e], 0));     // every slice has been written back
    }
    for (int d = m->ndev - 1; d >= 1; --d)
      if (gtb200_count_finish(m->ctx[d], nullptr, nullptr, nullptr, nullptr)) return -1;
    if (gtb200_count_finish(c0, m->h_lb.data(), countspecialcodes, distpfxidx, &ci)) return -1;
  }
  if (leftborder) memcpy(leftborder, m->h_lb.data(), (c0->numcodes + 1) * sizeof(uint32_t));
  if (info) *info = ci;
  m->counted = true;
  return 0;
}

// the per-GPU cut of a part: pure host logic, csrc/host_logic.c (tested on the CPU)
extern "C" int gtb200_split_range(const uint32_t *leftborder, uint64_t mincode, uint64_t maxcode, int nshards,
                                  uint64_t *lo, uint64_t *hi);
static int multi_split(const gtb200_multi *m, uint64_t mincode, uint64_t maxcode, uint64_t *lo, uint64_t *hi) {
  return gtb200_split_range(m->h_lb.data(), mincode, maxcode, m->ndev, lo, hi);
}

extern "C" int gtb200_multi_sort_part(gtb200_multi *m, uint64_t mincode, uint64_t maxcode, void *suftab,
                                      int suftab_is_uint32, uint8_t *lcptab, gtb200_largelcp *llv,
                                      uint64_t llv_capacity, gtb200_partinfo *info) {
  if (!m || !info) return set_error("gtb200_multi_sort_part: null argument");
  if (m->ndev == 1)
    return gtb200_sort_part(m->ctx[0], mincode, maxcode, suftab, suftab_is_uint32, lcptab, llv, llv_capacity, info);
  if (!m->counted) return set_error("gtb200_multi_sort_part: call gtb200_multi_count first");
  if (mincode > maxcode || maxcode >= m->ctx[0]->numcodes)
    return set_error("gtb200_multi_sort_part: code range [%llu,%llu] invalid", (unsigned long long)mincode,
                     (unsigned long long)maxcode);
  uint64_t lo[GTB200_MAXDEV], hi[GTB200_MAXDEV];
  const int k = multi_split(m, mincode, maxcode, lo, hi);
  const uint32_t *lb = m->h_lb.data();
  const uint64_t base = lb[mincode];
  gtb200_partinfo pi[GTB200_MAXDEV];
  int rc[GTB200_MAXDEV];
  std::string msg[GTB200_MAXDEV];
  std::vector<std::thread> workers;
  const size_t esz = suftab_is_uint32 ? 4 : 8;
  for (int j = 0; j < k; ++j) {
    workers.emplace_back([&, j]() {
      const uint64_t off = (uint64_t)lb[lo[j]] - base;
      rc[j] = gtb200_sort_part(m->ctx[j], lo[j], hi[j], suftab ? (char *)suftab + off * esz : nullptr, suftab_is_uint32,
                               lcptab ? lcptab + off : nullptr, nullptr, 0, &pi[j]);
      if (rc[j]) msg[j] = gtb200_last_error();
    });
  }
  for (auto &w : workers) w.join();
  for (int j = 0; j < k; ++j)
    if (rc[j]) return set_error("device %d: %s", m->ctx[j]->device, msg[j].c_str());
  memset(info, 0, sizeof *info);
  info->partoffset = base;
  info->width = (uint64_t)lb[maxcode + 1] - base;
  info->longest = UINT64_MAX;
  for (int j = 0; j < k; ++j) {
    if (pi[j].numoflargelcpvalues) {
      if (llv) {
        if (info->numoflargelcpvalues + pi[j].numoflargelcpvalues > llv_capacity)
          return set_error("gtb200_multi_sort_part: more than %llu large lcp values", (unsigned long long)llv_capacity);
        if (gtb200_fetch_part(m->ctx[j], nullptr, 0, nullptr, llv + info->numoflargelcpvalues,
                              llv_capacity - info->numoflargelcpvalues))
          return -1;
      }
      info->numoflargelcpvalues += pi[j].numoflargelcpvalues;
    }
    if (pi[j].maxbranchdepth > info->maxbranchdepth) info->maxbranchdepth = pi[j].maxbranchdepth;
    info->lcptabsum += pi[j].lcptabsum;
    if (pi[j].longest != UINT64_MAX) info->longest = pi[j].longest;
    info->fallback_suffixes += pi[j].fallback_suffixes;
    info->fallback_rounds = std::max(info->fallback_rounds, pi[j].fallback_rounds);
  }
  return 0;
}

extern "C" int gtb200_multi_tail(gtb200_multi *m, uint64_t first, uint64_t count, void *suftab, int suftab_is_uint32) {
  if (!m) return set_error("gtb200_multi_tail: null handle");
  return gtb200_tail(m->ctx[0], first, count, suftab, suftab_is_uint32);
}

// Page-locks a host buffer the caller owns (the GtSuffixsortspace array, the lcp buffer), so that the copies of
// finished slices really overlap the sorting of the next ones; pageable memory makes every copy synchronous.
extern "C" int gtb200_host_register(void *ptr, size_t bytes) {
  if (!ptr || !bytes) return 0;
  const cudaError_t e = cudaHostRegister(ptr, bytes, cudaHostRegisterPortable);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return set_error("cudaHostRegister(%zu bytes): %s", bytes, cudaGetErrorString(e));
  }
  return 0;
}

extern "C" void gtb200_host_unregister(void *ptr) {
  if (ptr && cudaHostUnregister(ptr) != cudaSuccess) (void)cudaGetLastError();
}

// Smallest free device memory over the handle's GPUs (blocks cached by this library count as free): what the host
// side needs to choose a number of parts that fits (the reference derives parts from -memlimit the same way,
// src/match/sfx-partssuf.c:449-527).
extern "C" int gtb200_multi_free_memory(gtb200_multi *m, uint64_t *free_bytes, uint64_t *total_bytes) {
  if (!m || !free_bytes || !total_bytes) return set_error("gtb200_multi_free_memory: null argument");
  uint64_t fmin = ~0ull, tmin = ~0ull;
  for (int d = 0; d < m->ndev; ++d) {
    size_t f = 0, t = 0;
    CK(cudaSetDevice(m->ctx[d]->device));
    CK(cudaMemGetInfo(&f, &t));
    {
      std::lock_guard<std::mutex> lock(g_pool_mu);
      for (auto &b : g_pool[m->ctx[d]->device]) f += b.sz;
    }
    fmin = std::min<uint64_t>(fmin, f);
    tmin = std::min<uint64_t>(tmin, t);
  }
  *free_bytes = fmin;
  *total_bytes = tmin;
  return 0;
}

"""Probe: do two half-range parts on two streams of one GPU overlap (partition of one with tile sort of the other)?"""
import ctypes as C, sys, time, threading
sys.path.insert(0, '/root/repo')
import numpy as np, torch
import bench
from genometools_b200 import encseq, sfx
L = sfx.load_library()
wl = sys.argv[1] if len(sys.argv) > 1 else "human"
enc, _ = bench.make_workload(wl, 1)
n = int(enc.shape[0])
tw = encseq.twobit_pack(enc); rg = encseq.special_ranges(enc).astype(np.uint64)
del enc
pl = sfx.recommended_prefixlength(4, n)
def open_ctx():
    d = sfx.SeqDesc(); d.numofchars, d.totallength, d.memory = 4, n, 0
    d.twobitencoding, d.numofwords = tw.ctypes.data, tw.shape[0]
    d.specialranges, d.numofspecialranges = rg.ctypes.data, rg.shape[0]
    c = C.c_void_p(); sfx._check(L.gtb200_open(C.byref(c), 0, C.byref(d), pl)); return c
lb = np.zeros(4 ** pl + 1, dtype=np.uint32)
A = open_ctx(); B = open_ctx()
info = sfx.CountInfo()
sfx._check(L.gtb200_count(A, lb.ctypes.data, None, None, C.byref(info)))
sfx._check(L.gtb200_count(B, None, None, None, None))
ncodes = 4 ** pl
tot = int(lb[-1])
def run(ctx, lo, hi, out):
    pi = sfx.PartInfo(); sfx._check(L.gtb200_sort_part_device(ctx, lo, hi, C.byref(pi))); out.append(int(pi.width))
for K in (2, 4):
    # K ranges alternating between the two contexts, each context processes its ranges in order
    bounds = [int(np.searchsorted(lb, tot * k // K)) for k in range(K + 1)]; bounds[0] = 0; bounds[-1] = ncodes
    for rep in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        def worker(ctx, idxs, out):
            for k in idxs: run(ctx, bounds[k], bounds[k + 1] - 1, out)
        oa, ob = [], []
        ta = threading.Thread(target=worker, args=(A, range(0, K, 2), oa)); tb = threading.Thread(target=worker, args=(B, range(1, K, 2), ob))
        ta.start(); tb.start(); ta.join(); tb.join(); torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print(f"{wl} K={K} two contexts concurrently: {dt*1e3:.1f} ms  (suffixes {sum(oa)+sum(ob)})", flush=True)
    for rep in range(2):
        torch.cuda.synchronize(); t0 = time.perf_counter(); o = []
        for k in range(K): run(A, bounds[k], bounds[k + 1] - 1, o)
        torch.cuda.synchronize(); print(f"{wl} K={K} one context sequentially: {(time.perf_counter()-t0)*1e3:.1f} ms", flush=True)
for rep in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter(); o = []
    run(A, 0, ncodes - 1, o); torch.cuda.synchronize(); print(f"{wl} whole range one context: {(time.perf_counter()-t0)*1e3:.1f} ms", flush=True)

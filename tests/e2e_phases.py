"""Where the end-to-end time goes (GPU box, not collected by pytest): host buffers in and out through the C ABI,
phase by phase, like bench.py's e2e leg.   python tests/e2e_phases.py [workload] [steps]"""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from genometools_b200 import encseq, sfx  # noqa: E402

L = sfx.load_library()
workload = sys.argv[1] if len(sys.argv) > 1 else "ecoli"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
enc, alphabet = bench.make_workload(workload, 1)
n = int(enc.shape[0])
words = encseq.twobit_pack(enc)
ranges = encseq.special_ranges(enc).astype(np.uint64)
nspecial = int((enc >= 254).sum())
del enc
words_h = torch.from_numpy(words.view(np.int64)).pin_memory()
ranges_h = torch.from_numpy(ranges.view(np.int64).reshape(-1)).pin_memory()
suf_h = torch.empty(n + 1, dtype=torch.int64).pin_memory()
lcp_h = torch.empty(n + 1, dtype=torch.uint8).pin_memory()
pl = sfx.recommended_prefixlength(4, n)
lb_h = torch.empty(4 ** pl + 1, dtype=torch.int32).pin_memory()
csc_h = torch.empty(4 ** (pl - 1), dtype=torch.int32).pin_memory()
dist_h = torch.empty(sum(4 ** i for i in range(1, pl - 1)) + 1, dtype=torch.int32).pin_memory()


def step(T):
    t0 = time.perf_counter()
    desc = sfx.SeqDesc()
    desc.numofchars, desc.totallength, desc.memory = 4, n, 0
    desc.twobitencoding, desc.numofwords = words_h.data_ptr(), words.shape[0]
    desc.specialranges, desc.numofspecialranges = ranges_h.data_ptr(), ranges.shape[0]
    ctx = C.c_void_p()
    sfx._check(L.gtb200_open(C.byref(ctx), 0, C.byref(desc), pl))
    t1 = time.perf_counter()
    info = sfx.CountInfo()
    sfx._check(L.gtb200_count(ctx, lb_h.data_ptr(), csc_h.data_ptr(), dist_h.data_ptr(), C.byref(info)))
    t2 = time.perf_counter()
    pi = sfx.PartInfo()
    sfx._check(L.gtb200_sort_part(ctx, 0, 4 ** pl - 1, suf_h.data_ptr(), 0, lcp_h.data_ptr(), None, 0, C.byref(pi)))
    t3 = time.perf_counter()
    sfx._check(L.gtb200_tail(ctx, 0, int(info.specialcharacters) + 1, suf_h.data_ptr() + 8 * int(pi.width), 0))
    t4 = time.perf_counter()
    L.gtb200_close(ctx)
    t5 = time.perf_counter()
    T.append((t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4))


T = []
for _ in range(2):
    step(T)
T = []
for _ in range(steps):
    step(T)
a = np.array(T) * 1e3
print("%s n=%d NARROW=%s threads=%s: open %.1f count %.1f sort_part %.1f tail %.1f close %.1f total %.1f ms (min total %.1f)"
      % (workload, n, os.environ.get("GTB200_NARROW", "default"), os.environ.get("GTB200_WIDEN_THREADS", "default"),
         *a.mean(0), a.sum(1).mean(), a.sum(1).min()))

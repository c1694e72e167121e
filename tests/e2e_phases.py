import sys, time, ctypes as C
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from genometools_b200 import sfx, synth, encseq
L = sfx.load_library()
n = 4_600_000
enc = synth.random_dna(n, seed=42)
words = encseq.twobit_pack(enc)
ranges = encseq.special_ranges(enc)
words_h = torch.from_numpy(words.view(np.int64)).pin_memory()
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
    desc.specialranges, desc.numofspecialranges = None, 0
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
for _ in range(5): step(T)
T = []
for _ in range(20): step(T)
a = np.array(T) * 1e3
print("open %.3f count %.3f sort_part %.3f tail %.3f close %.3f total %.3f ms" % (*a.mean(0), a.sum(1).mean()))

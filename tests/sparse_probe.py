import ctypes as C, sys, time
sys.path.insert(0, '/root/repo')
import numpy as np, torch
import bench
from genometools_b200 import encseq, sfx
L = sfx.load_library()
enc, _ = bench.make_workload("human", 1)
n = int(enc.shape[0]); tw = encseq.twobit_pack(enc); rg = encseq.special_ranges(enc).astype(np.uint64); del enc
pl = sfx.recommended_prefixlength(4, n)
d = sfx.SeqDesc(); d.numofchars, d.totallength, d.memory = 4, n, 0
d.twobitencoding, d.numofwords = tw.ctypes.data, tw.shape[0]
d.specialranges, d.numofspecialranges = rg.ctypes.data, rg.shape[0]
c = C.c_void_p(); sfx._check(L.gtb200_open(C.byref(c), 0, C.byref(d), pl))
sfx._check(L.gtb200_count(c, None, None, None, None))
nc = 4 ** pl
for rep in range(3):
    pi = sfx.PartInfo(); torch.cuda.synchronize(); t0 = time.perf_counter()
    sfx._check(L.gtb200_sort_part_device(c, nc // 8 * 3, nc // 8 * 4 - 1, C.byref(pi))); torch.cuda.synchronize()
    print("1/8 range:", (time.perf_counter() - t0) * 1e3, "ms width", pi.width, flush=True)

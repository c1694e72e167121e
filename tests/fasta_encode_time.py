"""How long FASTA -> encoded sequence takes on the device (GPU box, not collected by pytest):
    python tests/fasta_encode_time.py [workload]
The reference's encoder needs 58 s for the 3.1 Gbp genome (tests/golden/scale_fingerprints.json, "computing sequence encoding")."""
import ctypes as C
import io
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import bench  # noqa: E402
from genometools_b200 import sfx  # noqa: E402
from tests_scale_common import fast_write_fasta  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "human"
enc, alphabet = bench.make_workload(wl, 1)
path = "/tmp/fasta_encode_time.fna"
fast_write_fasta(path, enc, alphabet)
want_n, want_sp = int(enc.shape[0]), int((enc >= 254).sum())
del enc
data = np.fromfile(path, dtype=np.uint8)
os.unlink(path)
L = sfx.load_library()
for rep in range(3):
    ctx, fi = C.c_void_p(), sfx.FastaInfo()
    t0 = time.perf_counter()
    sfx._check(L.gtb200_open_fasta(C.byref(ctx), 0, data.ctypes.data, data.shape[0], 1 if alphabet == "protein" else 0, 0, C.byref(fi)))
    dt = time.perf_counter() - t0
    assert int(fi.totallength) == want_n and int(fi.specialcharacters) == want_sp
    L.gtb200_close(ctx)
    print(f"{wl}: {data.shape[0]} bytes of FASTA (pageable host memory) -> context with packed sequence, bitmap, ranges: "
          f"{dt*1e3:.1f} ms ({int(fi.numofsequences)} sequences, totallength {int(fi.totallength)})", flush=True)

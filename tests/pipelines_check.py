"""Every pipeline of the library once, on inputs small enough for instrumented kernels, each result compared with the
oracle (a 20-second check of a freshly built .so on a GPU box; also the workload for `compute-sanitizer --tool memcheck
python tests/pipelines_check.py [group ...]` where that tool may be run - it is closed on the B200 pool this was
developed on, so no sanitizer log is committed).  Not a pytest module.

groups: direct refine tile keyed protein fasta count narrow noalloc (default: all)."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from genometools_b200 import encseq, sfx, synth  # noqa: E402
from oracle import oracle  # noqa: E402


def same(got, want, what):
    for ext in ("suf", "lcp", "llv", "bck"):
        if got[ext] != want[ext]:
            raise AssertionError(f"{what}: .{ext} differs from the oracle")
    print("ok", what, flush=True)


class env:
    def __init__(self, **kv):
        self.kv = kv

    def __enter__(self):
        self.saved = {k: os.environ.get(k) for k in self.kv}
        os.environ.update(self.kv)

    def __exit__(self, *a):
        for k, v in self.saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def g_direct():
    enc = synth.multi_dna_with_nruns(120_000, nseq=6, seed=99)
    want = oracle.build(enc, 4)
    for parts in (1, 3):
        same(sfx.suffixerator(enc, "dna", parts=parts), want, f"direct/parts{parts}")
    prot = synth.random_protein(300, seed=5)
    same(sfx.suffixerator(prot, "protein"), oracle.build(prot, 20), "direct/protein")


def g_refine():
    rep = synth.repeat_rich_dna(unit=1500, copies=5, seed=98)
    want = oracle.build(rep, 4)
    for doubling in ("1", "0"):
        with env(GTB200_DOUBLING=doubling):
            same(sfx.suffixerator(rep, "dna"), want, f"refine/doubling{doubling}")
    same(sfx.suffixerator(rep, "dna", parts=2), want, "refine/parts2")


def g_tile():
    enc = synth.multi_dna_with_nruns(200_000, nseq=5, seed=97)
    want = oracle.build(enc, 4, 7)
    for levels in ("1", "2"):
        with env(GTB200_TILE="2", GTB200_TILE_LEVELS=levels):
            same(sfx.suffixerator(enc, "dna", prefixlength=7), want, f"tile/levels{levels}")


def g_keyed():
    enc = synth.multi_dna_with_nruns(200_000, nseq=5, seed=96)
    want = oracle.build(enc, 4, 7)
    for levels in ("1", "2"):
        for parts in (1, 3):            # parts > 1: the sparse level-1 partition
            with env(GTB200_TILE="2", GTB200_KEYED="2", GTB200_TILE_LEVELS=levels):
                same(sfx.suffixerator(enc, "dna", prefixlength=7, parts=parts), want, f"keyed/levels{levels}/parts{parts}")


def g_protein():
    prot = synth.random_protein(900, seed=95)
    want = oracle.build(prot, 20, 3)
    for subsort in ("1", "0"):
        with env(GTB200_TILE="2", GTB200_KEYED="2", GTB200_TILE_LEVELS="2", GTB200_SUBSORT=subsort):
            same(sfx.suffixerator(prot, "protein", prefixlength=3), want, f"keyed/protein/subsort{subsort}")


def g_fasta():
    enc = synth.multi_dna_with_nruns(60_000, nseq=4, seed=94)
    text = b"".join(b">s%d x\n" % k + b"\n".join(s[i:i + 70] for i in range(0, len(s), 70)) + b"\n"
                    for k, s in enumerate(encseq.decode_to_fasta_bytes(enc, "dna")))
    same(sfx.suffixerator(None, "dna", fasta=text), oracle.build(enc, 4), "fasta/dna")


def g_count():
    enc = synth.multi_dna_with_nruns(150_000, nseq=3, seed=93)
    want = oracle.build(enc, 4)
    with env(GTB200_FASTCOUNT="2"):
        for parts in (1, 2):
            same(sfx.suffixerator(enc, "dna", parts=parts), want, f"count/funnel/parts{parts}")


def g_narrow():
    enc = synth.multi_dna_with_nruns(4_300_000, nseq=3, seed=92)
    with env(GTB200_NARROW="2", GTB200_WIDEN_THREADS="3"):
        a = sfx.suffixerator(enc, "dna")
    with env(GTB200_NARROW="0"):
        b = sfx.suffixerator(enc, "dna")
    for ext in ("suf", "lcp", "llv", "bck"):
        assert a[ext] == b[ext], f"narrow transfer: .{ext} differs from the direct copy"
    print("ok narrow/4.3M", flush=True)


def g_noalloc():
    import torch
    L = sfx.load_library()
    counts = torch.randint(0, 9, (4 ** 8,), dtype=torch.int32, device="cuda:0")
    out = torch.zeros(4 ** 8 + 1, dtype=torch.int32, device="cuda:0")
    err = C.create_string_buffer(256)
    L.gtb200_leftborder_partialsums_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_char_p, C.c_size_t]
    rc = L.gtb200_leftborder_partialsums_noalloc(counts.data_ptr(), out.data_ptr(), 4 ** 8, None, err, len(err))
    assert rc == 0, err.value.decode()
    torch.cuda.synchronize()
    want = torch.cumsum(counts.to(torch.int64), 0)
    assert int(out[0]) == 0 and torch.equal(out[1:].to(torch.int64), want)
    print("ok noalloc", flush=True)


GROUPS = {"direct": g_direct, "refine": g_refine, "tile": g_tile, "keyed": g_keyed, "protein": g_protein, "fasta": g_fasta,
          "count": g_count, "narrow": g_narrow, "noalloc": g_noalloc}

if __name__ == "__main__":
    for name in (sys.argv[1:] or list(GROUPS)):
        GROUPS[name]()
    print("pipelines_check: all groups done")

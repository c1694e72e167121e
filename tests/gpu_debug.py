"""Run on the GPU box for quick diagnosis: walks a ladder of cases, prints the first mismatch of each."""
import sys, os, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from genometools_b200 import sfx, synth, encseq
from oracle import oracle
import importlib.util
_spec = importlib.util.spec_from_file_location('gt_conftest', os.path.join(os.path.dirname(os.path.abspath(__file__)), 'conftest.py'))
_cf = importlib.util.module_from_spec(_spec); _spec.loader.exec_module(_cf)
golden_names, load_golden = _cf.golden_names, _cf.load_golden

def cmp(got, want, label):
    bad = []
    for ext, w in (("bck", 4), ("suf", 8), ("lcp", 1), ("llv", 16)):
        if got[ext] != want[ext]:
            a = np.frombuffer(got[ext], dtype=np.uint8); b = np.frombuffer(want[ext], dtype=np.uint8)
            if a.size != b.size:
                bad.append(f"{ext}: size {a.size} vs {b.size}")
                continue
            d = np.unique(np.flatnonzero(a != b) // w)
            dt = {4: np.uint32, 8: np.uint64, 1: np.uint8, 16: np.uint64}[w]
            ga = np.frombuffer(got[ext], dtype=dt); wa = np.frombuffer(want[ext], dtype=dt)
            k = int(d[0]) * (2 if w == 16 else 1)
            bad.append(f"{ext}: {d.size} diffs, first@{int(d[0])} got {ga[k:k+6].tolist()} want {wa[k:k+6].tolist()}")
    for k in ("longest", "maxbranchdepth", "largelcpvalues", "averagelcp"):
        if got[k] != want[k]:
            bad.append(f"{k}: {got[k]} vs {want[k]}")
    print(("OK   " if not bad else "FAIL ") + label + (" fb=%d rounds=%d" % (got["fallback_suffixes"], got["fallback_rounds"])) + ("" if not bad else " :: " + " | ".join(bad)), flush=True)
    return not bad

def run(label, enc, alphabet, pl, want, **kw):
    try:
        got = sfx.suffixerator(enc, alphabet, prefixlength=pl, **kw)
        return cmp(got, want, label)
    except Exception as e:
        print("EXC  " + label + " :: " + repr(e)[:300], flush=True)
        traceback.print_exc()
        return False

if __name__ == "__main__":
    print("devices", sfx.load_library().gtb200_device_count(), flush=True)
    for name in golden_names():
        g = load_golden(name)
        run(name + "/p1", g["enc"], g["alphabet"], g["prefixlength"], g)
        run(name + "/p3", g["enc"], g["alphabet"], g["prefixlength"], g, parts=3)
        if g["alphabet"] == "dna":
            run(name + "/bytes", g["enc"], g["alphabet"], g["prefixlength"], g, use_twobit=False)
        os.environ["GTB200_MAXBUCKET"] = "0"
        run(name + "/refine", g["enc"], g["alphabet"], g["prefixlength"], g, parts=2)
        del os.environ["GTB200_MAXBUCKET"]
    enc = synth.random_dna(200_000, seed=11)
    want = oracle.build(enc, 4)
    run("dna200k", enc, "dna", want["prefixlength"], want)
    run("dna200k/p4", enc, "dna", want["prefixlength"], want, parts=4)

"""Parity at BASELINE.json's full sizes (run on the GPU box, not collected by pytest):

    python tests/scale_check.py --workload chr1 --reference      # ~4 min of reference CPU time

Builds the synthetic genome, runs the CUDA path through the C ABI, and compares its .suf/.lcp/.llv/.bck
with the UNMODIFIED reference's files (oracle/_ref/gt_ref) by SHA-256.  Without --reference only the
size-independent properties are checked (permutation, tail, sampled neighbour order/LCP, parts
neutrality)."""
import argparse
import hashlib
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import make_workload  # noqa: E402
from genometools_b200 import encseq, sfx  # noqa: E402


def sha(b):
    return hashlib.sha256(b).hexdigest()[:16]


def fast_write_fasta(path, enc, alphabet, width=80):
    seqs = encseq.decode_to_fasta_bytes(enc, alphabet)
    with open(path, "wb") as fh:
        for k, s in enumerate(seqs):
            fh.write(b">seq%d\n" % k)
            a = np.frombuffer(s, dtype=np.uint8)
            full = (a.shape[0] // width) * width
            if full:
                body = np.empty((full // width, width + 1), dtype=np.uint8)
                body[:, :width] = a[:full].reshape(-1, width)
                body[:, width] = 10
                fh.write(body.tobytes())
            if a.shape[0] > full:
                fh.write(a[full:].tobytes() + b"\n")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="chr1")
    ap.add_argument("--scale", type=int, default=1)
    ap.add_argument("--parts", type=int, default=1)
    ap.add_argument("--shrink", type=int, default=1, help="1/D of the named size (protein at 1/16 is 62 M residues)")
    ap.add_argument("--reference", action="store_true")
    ap.add_argument("--samples", type=int, default=20000)
    args = ap.parse_args()
    t0 = time.time()
    import bench
    bench.SHRINK = max(1, args.shrink)
    if args.scale > 8:
        raise SystemExit("--scale multiplies the workload (N-GPU weak scaling); use --shrink to run a fraction")
    enc, alphabet = make_workload(args.workload, args.scale)
    n = enc.shape[0]
    print(f"workload {args.workload}: n={n} specials={int((enc >= 254).sum())} gen {time.time()-t0:.1f}s", flush=True)
    t0 = time.time()
    got = sfx.suffixerator(enc, alphabet, parts=args.parts)
    t_gpu = time.time() - t0
    res = {"n": n, "gpu_wall_s": round(t_gpu, 2), "prefixlength": got["prefixlength"],
           "fallback_suffixes": got["fallback_suffixes"], "fallback_rounds": got["fallback_rounds"],
           "sha": {k: sha(got[k]) for k in ("suf", "lcp", "llv", "bck")},
           "longest": got["longest"], "maxbranchdepth": got["maxbranchdepth"], "averagelcp": got["averagelcp"],
           "largelcpvalues": got["largelcpvalues"]}
    print(json.dumps(res), flush=True)
    # size-independent properties
    suf = np.frombuffer(got["suf"], dtype=np.uint64)
    lcp = np.frombuffer(got["lcp"], dtype=np.uint8)
    seen = np.zeros(n + 1, dtype=bool)
    seen[suf] = True
    assert seen.all(), "not a permutation"
    del seen
    m = n - got["specialcharacters"]
    assert (np.diff(suf[m:].astype(np.int64)) > 0).all() and suf[-1] == n and (lcp[m:] == 0).all()
    sp = np.concatenate([np.flatnonzero(enc >= 254), [n]])
    rng = np.random.default_rng(0)
    for r in rng.integers(1, m, size=args.samples):
        a, b = int(suf[r - 1]), int(suf[r])
        la = int(sp[np.searchsorted(sp, a)]) - a
        lb = int(sp[np.searchsorted(sp, b)]) - b
        lim = min(la, lb, 600)
        x, y = enc[a:a + lim], enc[b:b + lim]
        neq = np.flatnonzero(x != y)
        l = int(neq[0]) if neq.size else lim
        if l < lim:
            assert x[l] < y[l], f"order violated at {r}"
        elif lim < 600:
            assert la > lb or (la == lb and a < b), f"special order violated at {r}"
        if lim < 600 or l < lim:
            assert lcp[r] == min(l, 255), f"lcp wrong at {r}"
    print(f"properties ok ({args.samples} sampled neighbour pairs)", flush=True)
    if args.reference:
        from oracle import oracle
        with tempfile.TemporaryDirectory() as tmp:
            fa = os.path.join(tmp, "w.fna")
            fast_write_fasta(fa, enc, alphabet)
            t0 = time.time()
            ref = oracle.run_reference(fa, alphabet, showtime=True)
            print(f"reference wall {time.time()-t0:.1f}s times {ref['times']}", flush=True)
        ok = True
        for k in ("suf", "lcp", "llv", "bck"):
            same = got[k] == ref[k]
            ok &= same
            print(f"  .{k}: {'IDENTICAL' if same else 'DIFFERENT'} ({len(ref[k])} bytes, sha {sha(ref[k])})", flush=True)
        prj = ref["prj"]
        for k in ("longest", "maxbranchdepth", "largelcpvalues", "averagelcp"):
            same = str(got[k]) == prj[k]
            ok &= same
            print(f"  prj {k}: {got[k]} vs {prj[k]} {'ok' if same else 'MISMATCH'}", flush=True)
        print("REFERENCE PARITY", "PASS" if ok else "FAIL", flush=True)
        sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()

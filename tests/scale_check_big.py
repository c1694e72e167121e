"""Byte-identity against the unmodified reference at BASELINE.json's largest DNA size (3.1 Gbp), streaming:
nothing larger than one part is held in host memory; both sides are compared by SHA-256 of the file images.

    python tests/scale_check_big.py --workload human --parts 8         # ~20 min of reference CPU time
"""
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
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from tests_scale_common import fast_write_fasta  # noqa: E402


def file_sha(path, chunk=1 << 26):
    h = hashlib.sha256()
    with open(path, "rb") as fh:
        while True:
            b = fh.read(chunk)
            if not b:
                break
            h.update(b)
    return h.hexdigest()[:16], os.path.getsize(path)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="human")
    ap.add_argument("--parts", type=int, default=8)
    ap.add_argument("--refparts", type=int, default=1)
    args = ap.parse_args()
    t0 = time.time()
    enc, alphabet = make_workload(args.workload, 1)
    n = enc.shape[0]
    print(f"workload {args.workload}: n={n} gen {time.time()-t0:.1f}s", flush=True)
    sigma = encseq.alphabet_size(alphabet)
    t0 = time.time()
    it = sfx.Sfxiterator(sigma, n, None, args.parts, twobit=encseq.twobit_pack(enc), ranges=encseq.special_ranges(enc))
    hs, hl, hv = hashlib.sha256(), hashlib.sha256(), hashlib.sha256()
    nbytes = {"suf": 0, "lcp": 0, "llv": 0}
    while True:
        r = it.next()
        if r is None:
            break
        suf, lcp, llv, special = r
        hs.update(suf.data)
        nbytes["suf"] += suf.nbytes
        if special:
            z = np.zeros(suf.shape[0], dtype=np.uint8)
            hl.update(z.data)
            nbytes["lcp"] += z.nbytes
        else:
            hl.update(lcp.data)
            nbytes["lcp"] += lcp.nbytes
            if llv.size:
                hv.update(llv.data)
                nbytes["llv"] += llv.nbytes
    bck = it.bcktab_bytes()
    got = {"suf": (hs.hexdigest()[:16], nbytes["suf"]), "lcp": (hl.hexdigest()[:16], nbytes["lcp"]),
           "llv": (hv.hexdigest()[:16], nbytes["llv"]), "bck": (hashlib.sha256(bck).hexdigest()[:16], len(bck))}
    stats = {"longest": it.longest(), "maxbranchdepth": it.maxbranchdepth, "largelcpvalues": it.numoflargelcpvalues,
             "averagelcp": "%.2f" % (it.lcptabsum / (n + 1)), "prefixlength": it.prefixlength,
             "fallback_suffixes": it.fallback_suffixes}
    it.delete()
    print(json.dumps({"gpu_wall_s": round(time.time() - t0, 1), "parts": args.parts, "files": got, **stats}), flush=True)
    from oracle import oracle
    with tempfile.TemporaryDirectory(dir="/tmp") as tmp:
        fa = os.path.join(tmp, "w.fna")
        t0 = time.time()
        fast_write_fasta(fa, enc, alphabet)
        del enc
        print(f"fasta written {time.time()-t0:.1f}s", flush=True)
        import subprocess
        cmd = [oracle.GT_REF, "suffixerator", "-" + alphabet, "-suf", "-lcp", "-bck", "-tis", "-v", "-parts", str(args.refparts),
               "-indexname", os.path.join(tmp, "ref"), "-db", fa]
        t0 = time.time()
        p = subprocess.run(cmd, capture_output=True, text=True, env=dict(os.environ, GT_ENV_OPTIONS="-showtime"))
        print(f"reference rc={p.returncode} wall {time.time()-t0:.1f}s", flush=True)
        print("\n".join(l for l in p.stdout.splitlines() if l.startswith("# TIME")), flush=True)
        if p.returncode != 0:
            print(p.stderr[-2000:])
            sys.exit(2)
        ok = True
        for k in ("suf", "lcp", "llv", "bck"):
            ref = file_sha(os.path.join(tmp, "ref." + k))
            same = ref == got[k]
            ok &= same
            print(f"  .{k}: {'IDENTICAL' if same else 'DIFFERENT'} ref {ref} gpu {got[k]}", flush=True)
        prj = oracle.parse_prj(open(os.path.join(tmp, "ref.prj")).read())
        for k in ("longest", "maxbranchdepth", "largelcpvalues", "averagelcp", "prefixlength"):
            same = str(stats[k]) == prj[k]
            ok &= same
            print(f"  prj {k}: {stats[k]} vs {prj[k]} {'ok' if same else 'MISMATCH'}", flush=True)
    print("REFERENCE PARITY", "PASS" if ok else "FAIL", flush=True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()

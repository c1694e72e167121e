"""north_star's target sentence as one run (GPU box, not collected by pytest):

    python tests/multi_scale_check.py --workload human --devices 0-7 --parts 8

builds the index of the 3.1 Gbp genome through gtb200_multi_* on the listed GPUs - every part cut into one bucket
range per GPU - streams the parts through SHA-256 and compares with the fingerprints of the UNMODIFIED reference's
.suf/.lcp/.llv/.bck (tests/golden/scale_fingerprints.json).  Prints one JSON line with the phase times."""
import argparse
import ctypes as C
import hashlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from genometools_b200 import encseq, sfx  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="human")
    ap.add_argument("--devices", default="0-7")
    ap.add_argument("--parts", type=int, default=8)
    args = ap.parse_args()
    lo, _, hi = args.devices.partition("-")
    devices = list(range(int(lo), int(hi) + 1)) if hi else [int(x) for x in lo.split(",")]
    want = json.load(open(os.path.join(ROOT, "tests", "golden", "scale_fingerprints.json")))[args.workload]
    t0 = time.time()
    enc, alphabet = bench.make_workload(args.workload, 1)
    n = int(enc.shape[0])
    tw, rg = (encseq.twobit_pack(enc), encseq.special_ranges(enc)) if alphabet == "dna" else (None, None)
    t_gen = time.time() - t0
    assert hashlib.sha256(enc.data).hexdigest() == want["enc_sha256"]
    t0 = time.time()
    if alphabet == "dna":
        it = sfx.Sfxiterator(4, n, None, args.parts, twobit=tw, ranges=rg, devices=devices)
    else:
        it = sfx.Sfxiterator(20, n, None, args.parts, enc=enc, devices=devices)
    t_open_count = time.time() - t0
    del enc
    h = {k: hashlib.sha256() for k in ("suf", "lcp", "llv")}
    size = {k: 0 for k in ("suf", "lcp", "llv")}
    t_sort = t_hash = 0.0
    while True:
        t0 = time.time()
        r = it.next()
        t_sort += time.time() - t0
        if r is None:
            break
        suf, lcp, llv, special = r
        t0 = time.time()
        h["suf"].update(suf.data)
        size["suf"] += suf.nbytes
        if special:
            z = np.zeros(suf.shape[0], dtype=np.uint8)
            h["lcp"].update(z.data)
            size["lcp"] += z.nbytes
        else:
            h["lcp"].update(lcp.data)
            size["lcp"] += lcp.nbytes
            if llv.size:
                h["llv"].update(llv.data)
                size["llv"] += llv.nbytes
        t_hash += time.time() - t0
    bck = it.bcktab_bytes()
    got = {k: {"bytes": size[k], "sha256": h[k].hexdigest()} for k in h}
    got["bck"] = {"bytes": len(bck), "sha256": hashlib.sha256(bck).hexdigest()}
    L = sfx.load_library()
    peer = int(L.gtb200_multi_peer_bytes(it._multi))
    prj = {"longest": str(it.longest()), "largelcpvalues": str(it.numoflargelcpvalues),
           "averagelcp": "%.2f" % (it.lcptabsum / (n + 1)), "maxbranchdepth": str(it.maxbranchdepth)}
    it.delete()
    ok = all(got[k] == want["files"][k] for k in got) and all(prj[k] == want["prj"][k] for k in prj)
    print(json.dumps({"workload": args.workload, "totallength": n, "devices": devices, "parts": args.parts,
                      "identical_to_reference": ok, "files": got, "prj": prj,
                      "seconds": {"synthesis": round(t_gen, 1), "open+broadcast+count": round(t_open_count, 2),
                                  "sort_all_parts_incl_d2h": round(t_sort, 2), "sha256_on_host": round(t_hash, 1)},
                      "suffixes_per_s_sort_incl_d2h": (n + 1) / t_sort, "nvlink_bytes": peer,
                      "reference_seconds_1_thread": want["reference"]["times_s"].get("overall")}))
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()

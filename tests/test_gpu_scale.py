"""Byte identity at BASELINE.json's FULL sizes (configs[2], [3], [4]) under `pytest -m gpu`.

tests/golden/scale_fingerprints.json holds size + SHA-256 of the .suf/.lcp/.llv/.bck files and the .prj
statistics that the UNMODIFIED reference (oracle/_ref/gt_ref, `gt suffixerator -dna|-protein -suf -lcp -bck -tis`)
wrote for the seeded synthetic inputs (generated once in the build container by
tests/golden/make_scale_fingerprints.py: 250 Mbp 79 s, 1 G residues ~10 min, 3.1 Gbp ~20 min of one core).
Here the same inputs are regenerated from their seeds, the index is built through the C ABI on the GPU, part by
part like gt_Sfxiterator_next, and the streamed hashes must match: 28 GB of output compared without shipping it.
"""
import hashlib
import json
import os

import numpy as np
import pytest

from genometools_b200 import encseq, sfx

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FP_PATH = os.path.join(ROOT, "tests", "golden", "scale_fingerprints.json")


def _fingerprints():
    with open(FP_PATH) as fh:
        return json.load(fh)


def _build_and_hash(name, parts):
    import bench
    enc, alphabet = bench.make_workload(name, 1)
    n = int(enc.shape[0])
    sigma = encseq.alphabet_size(alphabet)
    enc_sha = hashlib.sha256(enc.data).hexdigest()
    # GTB200_TEST_DEVICES=0,1,...: the same check through gtb200_multi_* (every part spread over the listed GPUs)
    devices = [int(x) for x in os.environ["GTB200_TEST_DEVICES"].split(",")] if os.environ.get("GTB200_TEST_DEVICES") else None
    if alphabet == "dna":
        it = sfx.Sfxiterator(sigma, n, None, parts, twobit=encseq.twobit_pack(enc), ranges=encseq.special_ranges(enc),
                             devices=devices)
    else:
        it = sfx.Sfxiterator(sigma, n, None, parts, enc=enc, devices=devices)
    del enc
    h = {k: hashlib.sha256() for k in ("suf", "lcp", "llv")}
    size = {k: 0 for k in ("suf", "lcp", "llv")}
    try:
        while True:
            r = it.next()
            if r is None:
                break
            suf, lcp, llv, special = r
            h["suf"].update(suf.data)
            size["suf"] += suf.nbytes
            if special:                                   # outmany0lcpvalues (sfx-lcpvalues.c:435-451)
                left = suf.shape[0]
                zeros = np.zeros(min(left, 1 << 26), dtype=np.uint8)
                while left:
                    k = min(left, zeros.shape[0])
                    h["lcp"].update(zeros[:k].data)
                    left -= k
                size["lcp"] += suf.shape[0]
            else:
                h["lcp"].update(lcp.data)
                size["lcp"] += lcp.nbytes
                if llv.size:
                    h["llv"].update(llv.data)
                    size["llv"] += llv.nbytes
        bck = it.bcktab_bytes()
        got = {k: {"bytes": size[k], "sha256": h[k].hexdigest()} for k in h}
        got["bck"] = {"bytes": len(bck), "sha256": hashlib.sha256(bck).hexdigest()}
        prj = {"totallength": str(n), "specialcharacters": str(it.specialcharacters), "longest": str(it.longest()),
               "prefixlength": str(it.prefixlength), "largelcpvalues": str(it.numoflargelcpvalues),
               "averagelcp": "%.2f" % (it.lcptabsum / (n + 1)), "maxbranchdepth": str(it.maxbranchdepth)}
    finally:
        it.delete()
    return enc_sha, got, prj


@pytest.mark.parametrize("name,parts", [("chr1", 1), ("protein", 2), ("human", 8)])
def test_full_size_index_matches_reference_fingerprint(name, parts):
    fp = _fingerprints()
    assert name in fp, f"{FP_PATH} has no entry for {name}: run tests/golden/make_scale_fingerprints.py {name}"
    want = fp[name]
    enc_sha, got, prj = _build_and_hash(name, parts)
    assert enc_sha == want["enc_sha256"], "synthetic input differs from the one the reference indexed (numpy RNG drift?)"
    for ext in ("bck", "suf", "lcp", "llv"):
        assert got[ext] == want["files"][ext], f"{name}: .{ext} differs from the reference's file: {got[ext]} vs {want['files'][ext]}"
    for k, v in want["prj"].items():
        assert prj[k] == v, f"{name}: .prj {k} = {prj[k]}, reference {v}"

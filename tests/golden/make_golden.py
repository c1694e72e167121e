"""Generates tests/golden/*.npz with the UNMODIFIED reference (oracle/_ref/gt_ref, built by
oracle/Makefile from /root/reference).  Run in the build container only:

    make -C oracle && python tests/golden/make_golden.py

Each .npz holds the input (encoded sequence as GtEncseq would hold it: codes 0..sigma-1, 254/255
specials) and the bytes of the files `gt suffixerator -dna|-protein -suf -lcp -bck -tis` wrote
(.suf .lcp .llv .bck) plus the .prj statistics.  The reference's own fixtures used here are the
ones testsuite/gt_suffixerator_include.rb:119-143,292-309 lists; the synthetic ones come from
genometools_b200/synth.py with fixed seeds.
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from genometools_b200 import encseq, synth  # noqa: E402
from oracle import oracle  # noqa: E402

TESTDATA = "/root/reference/testdata"
OUT = os.path.dirname(os.path.abspath(__file__))

CASES = [
    # name, source, alphabet, extra kwargs
    ("atinsert", ("file", "Atinsert.fna"), "dna", {}),
    ("atinsert_pl6_parts3", ("file", "Atinsert.fna"), "dna", {"pl": 6, "parts": 3}),
    ("duplicate", ("file", "Duplicate.fna"), "dna", {}),
    ("tttn", ("file", "TTTN.fna"), "dna", {}),
    ("verysmall", ("file", "Verysmall.fna"), "dna", {}),
    ("randomn", ("file", "RandomN.fna"), "dna", {}),
    ("random160", ("file", "Random160.fna"), "dna", {}),
    ("sw100k1", ("file", "sw100K1.fsa"), "protein", {}),
    ("sw100k1_pl2", ("file", "sw100K1.fsa"), "protein", {"pl": 2}),
    ("repeat_rich", ("synth", lambda: synth.repeat_rich_dna()), "dna", {}),
    ("repeat_rich_pl5", ("synth", lambda: synth.repeat_rich_dna(seed=11)), "dna", {"pl": 5}),
    ("multi_nruns_60k", ("synth", lambda: synth.multi_dna_with_nruns(60_000, nseq=7, seed=3)), "dna", {}),
    ("protein_synth_40k", ("synth", lambda: synth.random_protein(120, seed=9)), "protein", {"pl": 2}),
]


def main() -> None:
    for name, (kind, src), alphabet, kw in CASES:
        with tempfile.TemporaryDirectory() as tmp:
            if kind == "file":
                fasta = os.path.join(TESTDATA, src)
                enc = encseq.encode_fasta(fasta, alphabet)
            else:
                enc = src()
                fasta = os.path.join(tmp, name + ".fna")
                encseq.write_fasta(fasta, encseq.decode_to_fasta_bytes(enc, alphabet))
                assert np.array_equal(encseq.encode_fasta(fasta, alphabet), enc)
            ref = oracle.run_reference(fasta, alphabet, **kw)
        prj = ref["prj"]
        np.savez_compressed(
            os.path.join(OUT, name + ".npz"),
            enc=enc, alphabet=alphabet, prefixlength=int(prj["prefixlength"]),
            suf=np.frombuffer(ref["suf"], dtype=np.uint8), lcp=np.frombuffer(ref["lcp"], dtype=np.uint8),
            llv=np.frombuffer(ref["llv"], dtype=np.uint8), bck=np.frombuffer(ref["bck"], dtype=np.uint8),
            longest=int(prj["longest"]), maxbranchdepth=int(prj["maxbranchdepth"]),
            largelcpvalues=int(prj["largelcpvalues"]), averagelcp=prj["averagelcp"],
            specialcharacters=int(prj["specialcharacters"]), totallength=int(prj["totallength"]))
        print(f"{name}: n={prj['totallength']} pl={prj['prefixlength']} llv={prj['largelcpvalues']} "
              f"maxdepth={prj['maxbranchdepth']}")


if __name__ == "__main__":
    main()

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
    # the rest of the reference's DNA fixture list (testsuite/gt_suffixerator_include.rb:119-143)
    ("arabidopsis_c99826", ("file", "Arabidopsis-C99826.fna"), "dna", {}),
    ("atinsert_seqrange_13_17_rev", ("file", "Atinsert_seqrange_13-17_rev.fna"), "dna", {}),
    ("atinsert_seqrange_3_7", ("file", "Atinsert_seqrange_3-7.fna"), "dna", {}),
    ("atinsert_single_3", ("file", "Atinsert_single_3.fna"), "dna", {}),
    ("atinsert_single_3_rev", ("file", "Atinsert_single_3_rev.fna"), "dna", {}),
    ("copysorttest", ("file", "Copysorttest.fna"), "dna", {}),
    ("ecoli_section1", ("file", "Ecoli-section1.fna"), "dna", {}),
    ("ecoli_section2", ("file", "Ecoli-section2.fna"), "dna", {}),
    ("random_small", ("file", "Random-Small.fna"), "dna", {}),
    ("random", ("file", "Random.fna"), "dna", {}),
    ("random159", ("file", "Random159.fna"), "dna", {}),
    ("reads1", ("file", "Reads1.fna"), "dna", {}),
    ("reads2", ("file", "Reads2.fna"), "dna", {}),
    ("reads3", ("file", "Reads3.fna"), "dna", {}),
    ("repfind_example", ("file", "Repfind-example.fna"), "dna", {}),
    ("small", ("file", "Small.fna"), "dna", {}),
    ("smalldup", ("file", "Smalldup.fna"), "dna", {}),
    ("ttt_small", ("file", "TTT-small.fna"), "dna", {}),
    ("trna_glutamine", ("file", "trna_glutamine.fna"), "dna", {}),
    ("sw100k2", ("file", "sw100K2.fsa"), "protein", {}),
    # file lists (:148-153, :292-309): several -db arguments, one separator between files
    ("dna_filelist", ("files", ["Atinsert.fna", "Duplicate.fna", "Small.fna", "Verysmall.fna", "Random159.fna",
                                "Random160.fna"]), "dna", {"parts": 2}),
    ("protein_filelist", ("files", ["sw100K1.fsa", "sw100K2.fsa"]), "protein", {"parts": 3}),
    # symbol maps (:17-56 run the matrix with -smap TransDNA / TransProt11): TransDNA is the DNA alphabet, TransProt11
    # has 11 classes (sigma = 11: neither 4 nor 20), and a made-up 3-letter map puts a non-power-of-two alphabet on
    # the 2-bit text (sigma = 3)
    ("sw100k1_transprot11", ("file", "sw100K1.fsa"), "smap:/root/reference/gtdata/trans/TransProt11", {}),
    ("protein_filelist_transprot11", ("files", ["sw100K1.fsa", "sw100K2.fsa"]),
     "smap:/root/reference/gtdata/trans/TransProt11", {"parts": 2}),
    ("atinsert_transdna", ("file", "Atinsert.fna"), "smap:/root/reference/gtdata/trans/TransDNA", {}),
    ("three_letter_map", ("synthmap", lambda: [np.random.default_rng(33 + k).choice(
        np.array([0, 1, 2, 254], dtype=np.uint8), size=3_000 + 1_700 * k, p=[0.4, 0.3, 0.28, 0.02]) for k in range(6)]),
     "smap:abc", {}),
]

ABC_MAP = "aA\nbB\ncC\nxX\n"       # three letters + one wildcard class


def parse_smap(path_or_text):
    """Symbol map as gt_alphabet_new_from_file reads it (src/core/alphabet.c): one line per class, the characters of
    the class first (an optional display character after a blank), '#' comments; the LAST line is the wildcard class."""
    text = open(path_or_text).read() if os.path.exists(path_or_text) else path_or_text
    lines = [l for l in text.splitlines() if l.strip() and not l.startswith("#")]
    table = np.full(256, 253, dtype=np.uint8)
    for k, line in enumerate(lines):
        code = 254 if k == len(lines) - 1 else k
        for ch in line.split(" ")[0]:
            table[ord(ch)] = code
    return table, len(lines) - 1


def main() -> None:
    only = set(sys.argv[1:])
    for name, (kind, src), alphabet, kw in CASES:
        if only and name not in only:
            continue
        sigma, extra = None, []
        with tempfile.TemporaryDirectory() as tmp:
            if alphabet.startswith("smap:"):
                mapname = alphabet[5:]
                mapfile = mapname
                if mapname == "abc":
                    mapfile = os.path.join(tmp, "abc.map")
                    open(mapfile, "w").write(ABC_MAP)
                table, sigma = parse_smap(mapfile)
                extra = ["-smap", mapfile]
                if kind == "synthmap":
                    letters = np.frombuffer(b"abc", dtype=np.uint8)
                    out = []
                    for k, part in enumerate(src()):
                        body = np.where(part < 3, letters[np.minimum(part, 2)], ord("x")).astype(np.uint8)
                        out.append(b">s%d\n" % k + body.tobytes() + b"\n")
                    fasta = os.path.join(tmp, name + ".fna")
                    open(fasta, "wb").write(b"".join(out))
                    files = [fasta]
                else:
                    files = [os.path.join(TESTDATA, f) for f in ([src] if kind == "file" else src)]
                seqs = [q for f in files for q in encseq.read_fasta(f)]
                enc = np.concatenate([np.concatenate([table[np.frombuffer(q, dtype=np.uint8)], [255]]) for q in seqs])[:-1].astype(np.uint8)
                assert not (enc == 253).any(), "character outside the symbol map"
                ref = oracle.run_reference(files, "dna", extra=extra, drop_alphabet_flag=True, **kw)
            elif kind == "files":
                files = [os.path.join(TESTDATA, f) for f in src]
                enc = encseq.encode_sequences([q for f in files for q in encseq.read_fasta(f)], alphabet)
                ref = oracle.run_reference(files, alphabet, **kw)
            elif kind == "file":
                fasta = os.path.join(TESTDATA, src)
                enc = encseq.encode_fasta(fasta, alphabet)
                ref = oracle.run_reference(fasta, alphabet, **kw)
            else:
                enc = src()
                fasta = os.path.join(tmp, name + ".fna")
                encseq.write_fasta(fasta, encseq.decode_to_fasta_bytes(enc, alphabet))
                assert np.array_equal(encseq.encode_fasta(fasta, alphabet), enc)
                ref = oracle.run_reference(fasta, alphabet, **kw)
        prj = ref["prj"]
        assert int(prj["totallength"]) == enc.shape[0] and int(prj["specialcharacters"]) == int((enc >= 254).sum()), name
        np.savez_compressed(
            os.path.join(OUT, name + ".npz"),
            enc=enc, alphabet=alphabet, sigma=(sigma or encseq.alphabet_size(alphabet)), prefixlength=int(prj["prefixlength"]),
            suf=np.frombuffer(ref["suf"], dtype=np.uint8), lcp=np.frombuffer(ref["lcp"], dtype=np.uint8),
            llv=np.frombuffer(ref["llv"], dtype=np.uint8), bck=np.frombuffer(ref["bck"], dtype=np.uint8),
            longest=int(prj["longest"]), maxbranchdepth=int(prj["maxbranchdepth"]),
            largelcpvalues=int(prj["largelcpvalues"]), averagelcp=prj["averagelcp"],
            specialcharacters=int(prj["specialcharacters"]), totallength=int(prj["totallength"]))
        print(f"{name}: n={prj['totallength']} pl={prj['prefixlength']} llv={prj['largelcpvalues']} "
              f"maxdepth={prj['maxbranchdepth']}")


if __name__ == "__main__":
    main()

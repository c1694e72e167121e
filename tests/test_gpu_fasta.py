"""FASTA text straight to the device (gtb200_open_fasta): the index built from the raw
bytes of the file(s) must be byte-identical to the one the UNMODIFIED reference builds from the same file(s) - which
pins the device-side parse (descriptions, white space, '>' handling, symbol map, separators) to
gt_sequence_buffer_fasta_advance + the reference's alphabet."""
import os
import subprocess

import numpy as np
import pytest

from genometools_b200 import encseq, sfx, synth

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GT_REF = os.path.join(ROOT, "oracle", "_ref", "gt_ref")


def _fasta(enc, alphabet, width=80, eol=b"\n", lower=False, desc=lambda k: b"seq%d some text" % k):
    out = []
    for k, s in enumerate(encseq.decode_to_fasta_bytes(enc, alphabet)):
        if lower:
            s = s.lower()
        out.append(b">" + desc(k) + eol)
        if width:
            out += [s[i:i + width] + eol for i in range(0, len(s), width)]
        else:
            out.append(s + eol)
    return b"".join(out)


CASES = {
    "dna_wrapped": lambda: ([_fasta(synth.multi_dna_with_nruns(300_000, nseq=7, seed=301), "dna")], "dna"),
    "dna_unwrapped_long_lines": lambda: ([_fasta(synth.multi_dna_with_nruns(400_000, nseq=3, seed=302), "dna", width=0)], "dna"),
    "dna_crlf_lowercase_blank_lines": lambda: ([_fasta(synth.multi_dna_with_nruns(100_000, nseq=5, seed=303), "dna", width=61,
                                                       eol=b"\r\n", lower=True).replace(b"\r\n>", b"\r\n\r\n \t\r\n>")], "dna"),
    "dna_gt_inside_description": lambda: ([_fasta(synth.multi_dna_with_nruns(50_000, nseq=9, seed=304), "dna", width=70,
                                                  desc=lambda k: b"s%d >not a new sequence> x" % k)], "dna"),
    "dna_two_files": lambda: ([_fasta(synth.multi_dna_with_nruns(80_000, nseq=2, seed=305), "dna"),
                               _fasta(synth.random_dna(33_333, seed=306), "dna", width=50)], "dna"),
    "dna_tiny": lambda: ([b">a\nACGTN\n>b\nTT\n"], "dna"),
    "protein": lambda: ([_fasta(synth.random_protein(700, seed=307), "protein", width=60)], "protein"),
}


@pytest.mark.parametrize("parts", [1, 3])
@pytest.mark.parametrize("case", sorted(CASES))
def test_index_from_fasta_text_matches_reference(case, parts, tmp_path):
    assert os.path.isfile(GT_REF), f"{GT_REF} missing (build in the container first)"
    files, alphabet = CASES[case]()
    paths = []
    for k, data in enumerate(files):
        p = tmp_path / f"in{k}.fna"
        p.write_bytes(data)
        paths.append(str(p))
    subprocess.run([GT_REF, "suffixerator", "-" + alphabet, "-suf", "-lcp", "-bck", "-tis", "-indexname", "ref", "-db"] + paths,
                   cwd=str(tmp_path), check=True, capture_output=True)
    got = sfx.suffixerator(None, alphabet, parts=parts, fasta=b"".join(files))
    for ext in ("suf", "lcp", "llv", "bck"):
        want = (tmp_path / f"ref.{ext}").read_bytes()
        assert got[ext] == want, f"{case}: .{ext} built from FASTA text differs from the reference's"
    prj = dict(l.split("=", 1) for l in (tmp_path / "ref.prj").read_text().splitlines() if "=" in l)
    assert str(got["prefixlength"]) == prj["prefixlength"] and str(got["specialcharacters"]) == prj["specialcharacters"]


def test_fasta_errors_are_reported():
    with pytest.raises(sfx.Gtb200Error, match="illegal character at byte 9"):
        sfx.suffixerator(None, "dna", fasta=b">x\nACGTAC!GT\n")
    with pytest.raises(sfx.Gtb200Error, match="no sequences"):
        sfx.suffixerator(None, "dna", fasta=b"ACGT\nACGT\n")
    with pytest.raises(sfx.Gtb200Error, match="illegal character"):
        sfx.suffixerator(None, "protein", fasta=b">p\nLVIFKRED1\n")

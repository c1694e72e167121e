"""CPU tests that pin the oracle (oracle/sfx_oracle.c): against the committed golden vectors the
unmodified reference produced, and - when oracle/_ref/gt_ref is present - against the reference
itself on seeded random inputs, including `-parts` neutrality
(testsuite/gt_suffixerator_include.rb:58-69)."""
import os
import tempfile

import numpy as np
import pytest

from genometools_b200 import encseq, synth
from oracle import oracle


def test_oracle_matches_golden(golden):
    o = oracle.build(golden["enc"], golden["sigma"], golden["prefixlength"])
    for ext in ("suf", "lcp", "llv", "bck"):
        assert o[ext] == golden[ext], f"{golden['name']}.{ext} differs from the reference"
    assert o["longest"] == golden["longest"]
    assert o["maxbranchdepth"] == golden["maxbranchdepth"]
    assert o["largelcpvalues"] == golden["largelcpvalues"]
    assert o["averagelcp"] == golden["averagelcp"]


@pytest.mark.parametrize("sigma,n,pl", [(4, 11817, 4), (4, 4_600_000, 8), (4, 250_000_023, 11),
                                        (4, 3_100_000_000, 13), (20, 1_003_000_000, 5), (4, 9, 1),
                                        (20, 5234, 1), (4, 772376, 7)])
def test_prefixlength_rule(sigma, n, pl):
    # SURVEY.md section 8 table; the first, sixth, seventh and eighth rows were read off reference logs
    assert oracle.recommended_prefixlength(sigma, n) == pl


@pytest.mark.skipif(not oracle.have_reference(), reason="oracle/_ref/gt_ref not built")
@pytest.mark.parametrize("case", ["dna_random", "dna_nruns_parts3", "protein", "repeats_parts2"])
def test_oracle_matches_reference_binary(case):
    if case == "dna_random":
        enc, alphabet, kw = synth.random_dna(30_000, seed=1), "dna", {}
    elif case == "dna_nruns_parts3":
        enc, alphabet, kw = synth.multi_dna_with_nruns(40_000, nseq=5, seed=2), "dna", {"parts": 3}
    elif case == "protein":
        enc, alphabet, kw = synth.random_protein(60, seed=4), "protein", {"pl": 2}
    else:
        enc, alphabet, kw = synth.repeat_rich_dna(unit=300, copies=5, seed=5), "dna", {"parts": 2}
    sigma = encseq.alphabet_size(alphabet)
    with tempfile.TemporaryDirectory() as tmp:
        fa = os.path.join(tmp, "x.fna")
        encseq.write_fasta(fa, encseq.decode_to_fasta_bytes(enc, alphabet))
        ref = oracle.run_reference(fa, alphabet, **kw)
    o = oracle.build(enc, sigma, int(ref["prj"]["prefixlength"]))
    for ext in ("suf", "lcp", "llv", "bck"):
        assert o[ext] == ref[ext], ext
    assert str(o["longest"]) == ref["prj"]["longest"]
    assert o["averagelcp"] == ref["prj"]["averagelcp"]
    assert str(o["maxbranchdepth"]) == ref["prj"]["maxbranchdepth"]


def test_encoder_roundtrip_and_ranges():
    enc = synth.multi_dna_with_nruns(5_000, nseq=3, seed=8)
    seqs = encseq.decode_to_fasta_bytes(enc, "dna")
    assert np.array_equal(encseq.encode_sequences(seqs, "dna"), enc)
    rng = encseq.special_ranges(enc)
    sp = np.zeros(len(enc), dtype=bool)
    for a, b in rng:
        assert b > a
        sp[int(a):int(b)] = True
    assert np.array_equal(sp, enc >= 254)
    assert (rng[1:, 0] > rng[:-1, 1]).all()  # maximal and ascending
    tw = encseq.twobit_pack(enc)
    i = np.flatnonzero(~sp)
    got = (tw[i // 32] >> (62 - 2 * (i % 32)).astype(np.uint64)) & np.uint64(3)
    assert np.array_equal(got.astype(np.uint8), enc[i])


@pytest.mark.skipif(not oracle.have_reference(), reason="oracle/_ref/gt_ref not built")
@pytest.mark.parametrize("direction,readmode", [("rev", 1), ("cpl", 2), ("rcl", 3)])
def test_readmode_semantics_against_reference(direction, readmode):
    """`-dir rev|cpl|rcl` indexes the reversed / complemented sequence (specials stay specials): pins the
    transformation the GPU tests apply before asking the oracle."""
    enc = synth.multi_dna_with_nruns(20_000, nseq=3, seed=5)
    with tempfile.TemporaryDirectory() as tmp:
        fa = os.path.join(tmp, "x.fna")
        encseq.write_fasta(fa, encseq.decode_to_fasta_bytes(enc, "dna"))
        ref = oracle.run_reference(fa, "dna", extra=["-dir", direction])
    e = enc[::-1].copy() if readmode in (1, 3) else enc.copy()
    if readmode in (2, 3):
        e = np.where(e < 4, 3 - e, e).astype(np.uint8)
    o = oracle.build(e, 4, int(ref["prj"]["prefixlength"]))
    for ext in ("suf", "lcp", "llv", "bck"):
        assert o[ext] == ref[ext], ext
    assert ref["prj"]["readmode"] == str(readmode)

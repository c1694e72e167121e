"""The drop-in itself: oracle/_ref/gt_b200 is the reference's `gt suffixerator` with ONLY the Sfxiterator
translation unit replaced (integration/sfx_suffixer_b200.c over the C ABI) - option parser, GtEncseq, file
writers and .prj code are the unmodified reference.  Its index files must be byte-identical to those of the
unmodified binary oracle/_ref/gt_ref, and the reference's own checker (`gt dev sfxmap`) must accept them."""
import os
import subprocess

import numpy as np
import pytest

from genometools_b200 import encseq, synth

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GT_REF = os.path.join(ROOT, "oracle", "_ref", "gt_ref")
GT_B200 = os.path.join(ROOT, "oracle", "_ref", "gt_b200")


@pytest.fixture(autouse=True)
def _binaries_must_exist():
    """These tests only run with -m gpu, i.e. on the GPU box, where the prebuilt binaries travel with the snapshot
    (oracle/_ref is git-ignored, not gpurun-ignored).  Their absence is a broken build, not a reason to skip."""
    for b in (GT_REF, GT_B200):
        if not (os.path.isfile(b) and os.access(b, os.X_OK)):
            pytest.fail(f"{b} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` in the build "
                        "container (make -C oracle ref; make -C integration) before sending the tree to the GPU box")

CASES = {
    "dna_nruns": (lambda: synth.multi_dna_with_nruns(300_000, nseq=6, seed=81), "dna", []),
    "dna_nruns_parts3": (lambda: synth.multi_dna_with_nruns(300_000, nseq=6, seed=82), "dna", ["-parts", "3"]),
    "dna_repeats_pl5": (lambda: synth.genome_like_dna(250_000, seed=83), "dna", ["-pl", "5"]),
    "dna_suftabuint": (lambda: synth.random_dna(100_000, seed=84), "dna", ["-suftabuint"]),
    "dna_dir_rcl": (lambda: synth.multi_dna_with_nruns(150_000, nseq=5, seed=88), "dna", ["-dir", "rcl", "-parts", "2"]),
    "dna_dir_rev_bwt": (lambda: synth.genome_like_dna(120_000, seed=89), "dna", ["-dir", "rev", "-bwt"]),
    "dna_bwt_parts2": (lambda: synth.multi_dna_with_nruns(120_000, nseq=4, seed=87), "dna", ["-bwt", "-parts", "2"]),
    # options that change how the reference gets there, not what comes out (the drop-in ignores the strategy)
    "dna_mirrored": (lambda: synth.multi_dna_with_nruns(90_000, nseq=4, seed=90), "dna", ["-mirrored"]),
    "dna_mirrored_parts2": (lambda: synth.genome_like_dna(80_000, seed=91), "dna", ["-mirrored", "-parts", "2"]),
    "dna_memlimit": (lambda: synth.multi_dna_with_nruns(200_000, nseq=3, seed=92), "dna", ["-memlimit", "2MB"]),
    "dna_dc32": (lambda: synth.genome_like_dna(150_000, seed=93), "dna", ["-dc", "32"]),
    "dna_cmpcharbychar": (lambda: synth.multi_dna_with_nruns(100_000, nseq=3, seed=94), "dna", ["-cmpcharbychar"]),
    "dna_sat_uchar": (lambda: synth.multi_dna_with_nruns(100_000, nseq=3, seed=95), "dna", ["-sat", "uchar"]),
    "dna_sat_direct": (lambda: synth.multi_dna_with_nruns(100_000, nseq=3, seed=96), "dna", ["-sat", "direct"]),
    "dna_sat_bit": (lambda: synth.multi_dna_with_nruns(100_000, nseq=3, seed=97), "dna", ["-sat", "bit"]),
    "dna_lcpdist": (lambda: synth.genome_like_dna(60_000, seed=98), "dna", ["-lcpdist"]),
    "protein_parts2": (lambda: synth.random_protein(400, seed=85), "protein", ["-parts", "2"]),
    # symbol maps (testsuite/gt_suffixerator_include.rb:17-56 runs its matrix with -smap TransDNA / TransProt11):
    # sigma = 11 on byte text, and a three-letter map (sigma = 3: base-3 codes on the 2-bit text)
    "smap_transprot11": (lambda: synth.random_protein(300, seed=99), "smap:prot11", ["-parts", "2"]),
    "smap_three_letters": (lambda: None, "smap:abc", []),
}

TRANSPROT11 = "LVIFJ i\nKR +\nED -\nAG s\nST o\nNQ n\nYW a\nP p\nH h\nM m\nC c\nXUBZO*- x\n"   # gtdata/trans/TransProt11
ABC_MAP = "aA\nbB\ncC\nxX\n"


def _run(binary, args, cwd):
    p = subprocess.run([binary] + args, cwd=cwd, capture_output=True, text=True)
    assert p.returncode == 0, f"{os.path.basename(binary)} {' '.join(args)} failed:\n{p.stdout}\n{p.stderr}"
    return p.stdout


@pytest.mark.parametrize("case", sorted(CASES))
def test_gt_suffixerator_dropin_is_byte_identical(case, tmp_path):
    gen, alphabet, extra = CASES[case]
    enc = gen()
    fasta = str(tmp_path / "in.fna")
    if alphabet.startswith("smap:"):
        mapfile = str(tmp_path / "symbols.map")
        if alphabet == "smap:prot11":
            open(mapfile, "w").write(TRANSPROT11)
            encseq.write_fasta(fasta, encseq.decode_to_fasta_bytes(enc, "protein"))
        else:
            open(mapfile, "w").write(ABC_MAP)
            rng = np.random.default_rng(100)
            with open(fasta, "wb") as fh:
                for k in range(7):
                    body = rng.choice(np.frombuffer(b"abcx", dtype=np.uint8), size=20_000 + 3_000 * k, p=[0.4, 0.3, 0.28, 0.02])
                    fh.write(b">s%d\n" % k + body.tobytes() + b"\n")
        alpha_args = ["-smap", mapfile]
    else:
        encseq.write_fasta(fasta, encseq.decode_to_fasta_bytes(enc, alphabet))
        alpha_args = ["-" + alphabet]
    common = ["suffixerator"] + alpha_args + ["-suf", "-lcp", "-bck", "-tis", "-des", "-sds", "-ssp"] + extra
    out_ref = _run(GT_REF, common + ["-indexname", "ref", "-db", fasta], str(tmp_path))
    out_b200 = _run(GT_B200, common + ["-indexname", "b200", "-db", fasta], str(tmp_path))
    if "-lcpdist" in extra:                    # the LCP histogram the reference prints on stdout
        assert out_ref == out_b200
    exts = ("suf", "lcp", "llv", "bck", "esq", "prj") + (("bwt",) if "-bwt" in extra else ())
    for ext in exts:
        a = open(tmp_path / f"ref.{ext}", "rb").read()
        b = open(tmp_path / f"b200.{ext}", "rb").read()
        assert a == b, f".{ext} differs between the reference and the B200 drop-in ({len(a)} vs {len(b)} bytes)"
    if "-suftabuint" not in extra and "-dir" not in extra and "-mirrored" not in extra:
        # the reference's own checker on the drop-in's index
        _run(GT_REF, ["sfxmap", "-tis", "-suf", "-lcp", "-bck", "-esa", "b200"], str(tmp_path))


def test_dropin_reports_device_errors_through_gterror(tmp_path):
    fasta = str(tmp_path / "in.fna")
    encseq.write_fasta(fasta, encseq.decode_to_fasta_bytes(synth.random_dna(2000, seed=86), "dna"))
    env = dict(os.environ, GTB200_DEVICE="99")
    p = subprocess.run([GT_B200, "suffixerator", "-dna", "-suf", "-indexname", "x", "-db", fasta], cwd=str(tmp_path),
                       capture_output=True, text=True, env=env)
    assert p.returncode != 0 and "device 99 out of range" in p.stderr


@pytest.mark.parametrize("extra", [[], ["-parts", "3"]])
def test_packedindex_builds_through_the_dropin(extra, tmp_path):
    """The second in-tree caller of the Sfxiterator API: `gt packedindex mkindex` builds the packed (FM) index by
    iterating over the suffix array part by part (src/match/eis-suffixerator-interface.c:254,400).  With the drop-in
    behind it the block-compressed index (.bdx) and the project file must come out byte-identical."""
    enc = synth.multi_dna_with_nruns(400_000, nseq=5, seed=131)
    fasta = str(tmp_path / "in.fna")
    encseq.write_fasta(fasta, encseq.decode_to_fasta_bytes(enc, "dna"))
    args = ["packedindex", "mkindex", "-dna", "-tis", "-db", fasta] + extra
    _run(GT_REF, args + ["-indexname", "ref"], str(tmp_path))
    _run(GT_B200, args + ["-indexname", "b200"], str(tmp_path))
    for ext in ("bdx", "prj", "esq"):
        a = open(tmp_path / f"ref.{ext}", "rb").read()
        b = open(tmp_path / f"b200.{ext}", "rb").read()
        assert a == b, f"packedindex .{ext} differs between the reference and the B200 drop-in"

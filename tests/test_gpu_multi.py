"""Several GPUs inside the product (gtb200_multi_*, and the drop-in with GTB200_DEVICES): every part is cut into one
bucket range per GPU, sorted concurrently, and must come back byte-identical to the single-GPU / reference output.
Needs at least two visible GPUs (`gpurun --gpus 2`); skipped - loudly - on a one-GPU box."""
import os
import subprocess

import numpy as np
import pytest

from genometools_b200 import encseq, sfx, synth
from oracle import oracle

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GT_REF = os.path.join(ROOT, "oracle", "_ref", "gt_ref")
GT_B200 = os.path.join(ROOT, "oracle", "_ref", "gt_b200")


def _ndev():
    return int(sfx.load_library().gtb200_device_count())


needs_two = pytest.mark.skipif(_ndev() < 2, reason="multi-GPU test: fewer than 2 CUDA devices visible")


def _devices():
    return list(range(min(_ndev(), 8)))


def _same(got, want, label):
    for ext in ("bck", "suf", "lcp", "llv"):
        assert got[ext] == want[ext], f"{label}: .{ext} differs"
    for k in ("longest", "maxbranchdepth", "largelcpvalues", "averagelcp"):
        assert got[k] == want[k], f"{label}: {k} {got[k]} != {want[k]}"


CASES = {
    "dna_nruns_2m": lambda: (synth.multi_dna_with_nruns(2_000_000, nseq=7, seed=201), "dna", None),
    "dna_repeats": lambda: (synth.genome_like_dna(400_000, seed=202), "dna", None),
    "dna_pl3_tiny_parts": lambda: (synth.multi_dna_with_nruns(50_000, nseq=3, seed=203), "dna", 3),
    "protein_150k": lambda: (synth.random_protein(450, seed=204), "protein", None),
    "dna_all_special": lambda: (np.array([254, 255, 254], dtype=np.uint8), "dna", 1),
}


@needs_two
@pytest.mark.parametrize("parts", [1, 3])
@pytest.mark.parametrize("case", sorted(CASES))
def test_multi_device_parts_match_oracle(case, parts):
    enc, alphabet, pl = CASES[case]()
    want = oracle.build(enc, encseq.alphabet_size(alphabet), pl)
    got = sfx.suffixerator(enc, alphabet, parts=parts, prefixlength=want["prefixlength"], devices=_devices())
    _same(got, want, f"multi/{case}/parts{parts}")


@needs_two
def test_multi_device_tile_and_keyed_paths(monkeypatch):
    """Large parts through the tile path (8- and 16-byte records) and the sliced host copies, two devices."""
    enc = synth.multi_dna_with_nruns(6_000_000, nseq=8, seed=205)
    want = oracle.build(enc, 4)
    monkeypatch.setenv("GTB200_TILE", "2")
    for keyed in ("0", "2"):
        monkeypatch.setenv("GTB200_KEYED", keyed)
        got = sfx.suffixerator(enc, "dna", devices=_devices()[:2])
        _same(got, want, f"multi/tile/keyed{keyed}")


@needs_two
@pytest.mark.parametrize("extra", [[], ["-parts", "3", "-bwt"], ["-suftabuint"]])
def test_dropin_with_several_devices_is_byte_identical(extra, tmp_path):
    for b in (GT_REF, GT_B200):
        assert os.path.isfile(b), f"{b} missing (build in the container first)"
    enc = synth.multi_dna_with_nruns(1_500_000, nseq=6, seed=206)
    fasta = str(tmp_path / "in.fna")
    encseq.write_fasta(fasta, encseq.decode_to_fasta_bytes(enc, "dna"))
    common = ["suffixerator", "-dna", "-suf", "-lcp", "-bck", "-tis"] + extra
    subprocess.run([GT_REF] + common + ["-indexname", "ref", "-db", fasta], cwd=str(tmp_path), check=True, capture_output=True)
    env = dict(os.environ, GTB200_DEVICES="0-%d" % (min(_ndev(), 8) - 1))
    p = subprocess.run([GT_B200] + common + ["-indexname", "b200", "-db", fasta], cwd=str(tmp_path), capture_output=True,
                       text=True, env=env)
    assert p.returncode == 0, p.stderr
    for ext in ("suf", "lcp", "llv", "bck", "prj") + (("bwt",) if "-bwt" in extra else ()):
        a = open(tmp_path / f"ref.{ext}", "rb").read()
        b = open(tmp_path / f"b200.{ext}", "rb").read()
        assert a == b, f".{ext} differs between the reference and the multi-GPU drop-in"

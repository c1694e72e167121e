"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/gtsfx_b200.h declares,
its pure-host helpers agree with the oracle / the reference, and without a GPU it fails loudly
instead of falling back to a CPU path."""
import os
import re

import numpy as np
import pytest

from genometools_b200 import sfx, synth
from oracle import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _have_gpu():
    try:
        return sfx.load_library().gtb200_device_count() > 0
    except Exception:
        return False


def test_library_exports_header_symbols():
    L = sfx.load_library()
    header = open(os.path.join(ROOT, "include", "gtsfx_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(gtb200_[a-z0-9_]+)\s*\(", header)))
    assert declared, "no declarations found"
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/gtsfx_b200.h but not exported"
    assert set(sfx.EXPORTS) <= set(declared)


@pytest.mark.parametrize("sigma,n", [(4, 11817), (4, 4_600_000), (4, 250_000_023), (4, 3_100_000_000),
                                     (20, 1_003_000_000), (4, 9), (20, 5234), (4, 772376), (4, 1), (20, 40519)])
def test_prefixlength_matches_oracle(sigma, n):
    assert sfx.recommended_prefixlength(sigma, n) == oracle.recommended_prefixlength(sigma, n)


def test_suftabparts_cover_and_balance():
    rng = np.random.default_rng(0)
    counts = rng.poisson(3.0, size=4 ** 6).astype(np.uint32)
    counts[100:400] = 0
    lb = np.concatenate([[0], np.cumsum(counts)]).astype(np.uint32)
    total = int(lb[-1])
    for parts in (1, 2, 3, 7, 8, 64):
        p = sfx.suftabparts(parts, lb, 6)
        assert 1 <= len(p) <= parts
        assert p[0][0] == 0 and p[-1][1] == 4 ** 6 - 1
        assert sum(w for _, _, w in p) == total
        for (a0, b0, _), (a1, _, _) in zip(p[:-1], p[1:]):
            assert a1 == b0 + 1
        for a, b, w in p:
            assert w == int(lb[b + 1]) - int(lb[a]) and w > 0
            assert abs(w - total / parts) <= counts.max() + 1
    assert len(sfx.suftabparts(4, lb, 1)) == 1  # prefixlength 1 forces one part (sfx-partssuf.c:216-221)


@pytest.mark.skipif(not oracle.have_reference(), reason="oracle/_ref/gt_ref not built")
def test_suftabparts_widths_match_reference_log(tmp_path):
    from genometools_b200 import encseq
    enc = synth.multi_dna_with_nruns(50_000, nseq=4, seed=12)
    fa = str(tmp_path / "x.fna")
    encseq.write_fasta(fa, encseq.decode_to_fasta_bytes(enc, "dna"))
    ref = oracle.run_reference(fa, "dna", parts=3)
    widths = [int(m) for m in re.findall(r"widthofpart\[\d+\]=(\d+)", ref["log"])]
    o = oracle.build(enc, 4, int(ref["prj"]["prefixlength"]))
    mine = sfx.suftabparts(3, o["leftborder"], o["prefixlength"])
    assert [w for _, _, w in mine] == widths


@pytest.mark.skipif(_have_gpu(), reason="a GPU is present")
def test_no_cpu_fallback_without_gpu():
    enc = synth.random_dna(1000, seed=3)
    with pytest.raises(sfx.Gtb200Error) as ei:
        sfx.suffixerator(enc, "dna")
    assert "no CPU fallback" in str(ei.value) or "CUDA" in str(ei.value)


def test_split_range_covers_the_part_in_balanced_contiguous_ranges():
    """gtb200_split_range (the per-GPU cut inside gtb200_multi_sort_part) is host logic: every code of the part lies in
    exactly one range, ranges are ordered, none is without suffixes, and no range exceeds its fair share by more than
    the bucket that crosses the target."""
    rng = np.random.default_rng(7)
    for trial in range(40):
        ncodes = int(rng.integers(1, 5000))
        sizes = rng.integers(0, 50, size=ncodes).astype(np.uint32)
        sizes[rng.random(ncodes) < 0.3] = 0
        if trial % 5 == 0:
            sizes[:] = 0
            sizes[int(rng.integers(0, ncodes))] = 1000          # everything in one bucket
        lb = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint32)
        mincode = int(rng.integers(0, ncodes))
        maxcode = int(rng.integers(mincode, ncodes))
        for nshards in (1, 2, 3, 8, 16):
            ranges = sfx.split_range(lb, mincode, maxcode, nshards)
            total = int(lb[maxcode + 1]) - int(lb[mincode])
            if total == 0:
                assert len(ranges) <= 1
                continue
            assert 1 <= len(ranges) <= nshards
            assert ranges[0][0] == mincode and ranges[-1][1] == maxcode
            for (a, b), (c, d) in zip(ranges, ranges[1:]):
                assert c == b + 1
            widths = [int(lb[b + 1]) - int(lb[a]) for a, b in ranges]
            assert sum(widths) == total and all(w > 0 for w in widths[:-1])
            fair = total / nshards
            assert max(widths) <= 2 * fair + int(sizes.max()) + 1


def test_device_list_parser():
    import ctypes as C
    L = sfx.load_library()
    L.gtb200_parse_devices.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.c_int]
    out = (C.c_int * 16)()
    for text, want in ((b"0", [0]), (b"0,1,4-7", [0, 1, 4, 5, 6, 7]), (b"3-3", [3]), (b"0-15", list(range(16))),
                       (b"0-99", list(range(16))), (b"", []), (b"x", []), (b"1,", [1]), (b"2,x", [2]), (b"5-2", [])):
        k = L.gtb200_parse_devices(text, out, 16)
        assert [out[i] for i in range(k)] == want, (text, [out[i] for i in range(k)])


def test_widening_of_positions_is_exact_for_every_alignment():
    import ctypes as C
    L = sfx.load_library()
    L.gtb200_widen_u32_u64.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
    L.gtb200_widen_u32_u64.restype = None
    rng = np.random.default_rng(3)
    src = rng.integers(0, 2 ** 32, size=5000, dtype=np.uint64).astype(np.uint32)
    for off_s in (0, 1, 3):
        for off_d in (0, 1, 2, 3):
            for n in (0, 1, 7, 8, 9, 31, 1000, 4093):
                dst = np.full(n + 8, 0xdeadbeefdeadbeef, dtype=np.uint64)
                L.gtb200_widen_u32_u64(src[off_s:].ctypes.data, dst[off_d:].ctypes.data, n)
                assert np.array_equal(dst[off_d:off_d + n], src[off_s:off_s + n].astype(np.uint64))
                assert (dst[:off_d] == 0xdeadbeefdeadbeef).all() and (dst[off_d + n:] == 0xdeadbeefdeadbeef).all()

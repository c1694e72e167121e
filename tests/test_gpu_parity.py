"""GPU parity tests (run with -m gpu on the B200 box): the CUDA path, called through the C ABI,
must reproduce the reference's .suf/.lcp/.llv/.bck byte for byte.  Checkers: the committed golden
vectors (made by the unmodified reference), the plain-C oracle on seeded inputs, and - at sizes
the oracle cannot reach in seconds - size-independent properties (sortedness, permutation, LCP
consistency on samples, parts neutrality)."""
import numpy as np
import pytest

from genometools_b200 import encseq, sfx, synth
from oracle import oracle

pytestmark = pytest.mark.gpu


def load_golden_by_name(name):
    import os
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name + ".npz"))
    return {"enc": z["enc"], "prefixlength": int(z["prefixlength"])}


def _compare(got, want, label):
    for ext in ("bck", "suf", "lcp", "llv"):
        if got[ext] != want[ext]:
            a = np.frombuffer(got[ext], dtype=np.uint8)
            b = np.frombuffer(want[ext], dtype=np.uint8)
            msg = f"{label}.{ext}: sizes {a.size} vs {b.size}"
            if a.size == b.size:
                width = {"suf": 8, "llv": 16, "bck": 4, "lcp": 1}[ext]
                d = np.flatnonzero(a != b) // width
                first = int(d[0])
                msg += f"; {np.unique(d).size} differing entries, first at {first}"
                if ext == "suf":
                    ga = np.frombuffer(got[ext], dtype=np.uint64)
                    wa = np.frombuffer(want[ext], dtype=np.uint64)
                    msg += f" got {ga[first:first+4]} want {wa[first:first+4]}"
                if ext == "lcp":
                    msg += f" got {a[first:first+6]} want {b[first:first+6]}"
            pytest.fail(msg)
    for k in ("longest", "maxbranchdepth", "largelcpvalues", "averagelcp"):
        assert got[k] == want[k], f"{label}: {k} {got[k]} != {want[k]}"


def test_golden_vectors(golden):
    for parts in (1, 3):
        got = sfx.suffixerator(golden["enc"], golden["alphabet"], parts=parts, prefixlength=golden["prefixlength"],
                               sigma=golden["sigma"])
        _compare(got, golden, f"{golden['name']}/parts{parts}")
    if golden["alphabet"] == "dna":   # byte input path (ranges derived on the device)
        got = sfx.suffixerator(golden["enc"], "dna", prefixlength=golden["prefixlength"], use_twobit=False)
        _compare(got, golden, f"{golden['name']}/bytes")


def test_golden_vectors_refinement_only(golden, monkeypatch):
    """Every bucket forced through the refinement path (segmented radix sort on text windows)."""
    monkeypatch.setenv("GTB200_MAXBUCKET", "0")
    got = sfx.suffixerator(golden["enc"], golden["alphabet"], parts=2, prefixlength=golden["prefixlength"],
                           sigma=golden["sigma"])
    _compare(got, golden, f"{golden['name']}/refine")
    if golden["maxbranchdepth"] > 0:   # some bucket holds at least two suffixes
        assert got["fallback_suffixes"] > 0


CASES = {
    "dna_200k": lambda: (synth.random_dna(200_000, seed=11), "dna", None),
    "dna_nruns_300k": lambda: (synth.multi_dna_with_nruns(300_000, nseq=9, seed=12), "dna", None),
    "dna_nruns_300k_pl3": lambda: (synth.multi_dna_with_nruns(300_000, nseq=9, seed=13), "dna", 3),
    "dna_repeats": lambda: (synth.repeat_rich_dna(unit=4000, copies=8, seed=14), "dna", None),
    "dna_skewed": lambda: (np.random.default_rng(15).choice(4, size=150_000, p=[0.7, 0.1, 0.1, 0.1]).astype(np.uint8), "dna", None),
    "protein_60k": lambda: (synth.random_protein(180, seed=16), "protein", None),
    "protein_60k_pl3": lambda: (synth.random_protein(180, seed=17), "protein", 3),
    "dna_tiny_special_ends": lambda: (np.array([254, 0, 1, 255, 2, 3, 3, 254, 254, 1, 255], dtype=np.uint8), "dna", 2),
    "dna_all_special": lambda: (np.array([254, 255, 254], dtype=np.uint8), "dna", 1),
    "dna_single": lambda: (np.array([2], dtype=np.uint8), "dna", 1),
    "dna_polyT": lambda: (np.full(3000, 3, dtype=np.uint8), "dna", 4),
}


@pytest.mark.parametrize("case", sorted(CASES))
@pytest.mark.parametrize("parts", [1, 4])
def test_against_oracle(case, parts):
    enc, alphabet, pl = CASES[case]()
    sigma = encseq.alphabet_size(alphabet)
    want = oracle.build(enc, sigma, pl)
    got = sfx.suffixerator(enc, alphabet, parts=parts, prefixlength=want["prefixlength"])
    _compare(got, want, f"{case}/parts{parts}")


@pytest.mark.parametrize("parts", [1, 3])
def test_twolevel_insertion_forced(parts, monkeypatch):
    """The partition-then-scatter insertion used for parts larger than L2, forced on a small input."""
    monkeypatch.setenv("GTB200_TWOLEVEL_BYTES", "0")
    enc = synth.multi_dna_with_nruns(400_000, nseq=7, seed=31)
    want = oracle.build(enc, 4, 7)
    got = sfx.suffixerator(enc, "dna", parts=parts, prefixlength=7)
    _compare(got, want, f"twolevel/parts{parts}")
    prot = synth.random_protein(400, seed=32)
    want = oracle.build(prot, 20, 3)
    got = sfx.suffixerator(prot, "protein", parts=parts, prefixlength=3)
    _compare(got, want, f"twolevel-protein/parts{parts}")


TILE_CASES = {
    "nruns": lambda: (synth.multi_dna_with_nruns(500_000, nseq=7, seed=51), None),
    "nruns_pl5": lambda: (synth.multi_dna_with_nruns(300_000, nseq=3, seed=52), 5),
    "skewed": lambda: (np.random.default_rng(53).choice(4, size=400_000, p=[0.75, 0.1, 0.1, 0.05]).astype(np.uint8), 7),
    "repeats": lambda: (synth.repeat_rich_dna(unit=5000, copies=10, seed=54), None),
    "polyT_mixed": lambda: (np.concatenate([np.full(20_000, 3, dtype=np.uint8), synth.random_dna(100_000, seed=55)]), 6),
}


@pytest.mark.parametrize("levels", ["1", "2"])
@pytest.mark.parametrize("parts", [1, 3])
@pytest.mark.parametrize("case", sorted(TILE_CASES))
def test_tile_path_forced(case, parts, levels, monkeypatch):
    """The staged-partition + shared-memory tile sorter, forced on small inputs (one and two partition
    levels), including buckets it must hand over to the global sorters and the refinement path."""
    monkeypatch.setenv("GTB200_TILE", "2")
    monkeypatch.setenv("GTB200_TILE_LEVELS", levels)
    enc, pl = TILE_CASES[case]()
    want = oracle.build(enc, 4, pl)
    got = sfx.suffixerator(enc, "dna", parts=parts, prefixlength=want["prefixlength"])
    _compare(got, want, f"tile/{case}/parts{parts}/levels{levels}")


@pytest.mark.parametrize("levels", ["1", "2"])
@pytest.mark.parametrize("parts", [1, 3])
@pytest.mark.parametrize("case", sorted(TILE_CASES))
def test_tile_path_keyed_records(case, parts, levels, monkeypatch):
    """The tile path with 16-byte records that carry their sort key (what genomes larger than L2 use), forced on
    small inputs: the keys built from three text words in the level-1 partition must equal the gathered ones,
    including keys cut by specials, sparse code ranges (parts) and buckets handed to the other sorters."""
    monkeypatch.setenv("GTB200_TILE", "2")
    monkeypatch.setenv("GTB200_KEYED", "2")
    monkeypatch.setenv("GTB200_TILE_LEVELS", levels)
    enc, pl = TILE_CASES[case]()
    want = oracle.build(enc, 4, pl)
    got = sfx.suffixerator(enc, "dna", parts=parts, prefixlength=want["prefixlength"])
    _compare(got, want, f"keyed/{case}/parts{parts}/levels{levels}")


def test_tile_path_keyed_records_sliced_fetch(monkeypatch):
    """Keyed records under the pipelined fetch: the part is cut into code ranges whose level-1 records pass through
    the image buffer while earlier slices of the image are already being copied out."""
    monkeypatch.setenv("GTB200_TILE", "2")
    monkeypatch.setenv("GTB200_KEYED", "2")
    enc = synth.multi_dna_with_nruns(5_000_000, nseq=8, seed=97)
    want = oracle.build(enc, 4)
    for parts in (1, 2):
        _compare(sfx.suffixerator(enc, "dna", parts=parts), want, f"keyed/sliced/parts{parts}")


@pytest.mark.parametrize("levels", ["1", "2"])
def test_tile_path_forced_protein(levels, monkeypatch):
    """The same path on byte text (sigma = 20, codes are base-20 numbers): most buckets are larger than a warp
    sorts, so the tile sorter only inserts them (no key gather) and the block sorter finishes them."""
    monkeypatch.setenv("GTB200_TILE", "2")
    monkeypatch.setenv("GTB200_TILE_LEVELS", levels)
    for nseq, pl, parts in ((1500, 3, 1), (3000, 2, 2), (400, 4, 1), (900, 2, 1), (500, 2, 3)):
        prot = synth.random_protein(nseq, seed=60 + nseq % 7)
        want = oracle.build(prot, 20, pl)
        got = sfx.suffixerator(prot, "protein", prefixlength=pl, parts=parts)
        _compare(got, want, f"tile/protein/{nseq}/pl{pl}/levels{levels}")


@pytest.mark.parametrize("subsort", ["1", "0"])
@pytest.mark.parametrize("levels", ["1", "2"])
def test_tile_path_keyed_records_protein(levels, subsort, monkeypatch):
    """Keyed records on byte text: base-20 codes rolled along 16 positions per thread, 7-residue keys cut by
    separators/wildcards, and the keys of buckets the tile only inserts handed to the block sorter beside the positions."""
    monkeypatch.setenv("GTB200_TILE", "2")
    monkeypatch.setenv("GTB200_KEYED", "2")
    monkeypatch.setenv("GTB200_SUBSORT", subsort)      # 1: buckets split by the first key symbol inside the tile (default)
    monkeypatch.setenv("GTB200_TILE_LEVELS", levels)
    for nseq, pl, parts in ((1500, 3, 1), (3000, 2, 2), (400, 4, 1), (900, 2, 1), (500, 2, 3), (2500, 5, 1)):
        prot = synth.random_protein(nseq, seed=160 + nseq % 7)
        want = oracle.build(prot, 20, pl)
        got = sfx.suffixerator(prot, "protein", prefixlength=pl, parts=parts)
        _compare(got, want, f"keyed/protein/{nseq}/pl{pl}/levels{levels}")
    g = load_golden_by_name("protein_filelist_transprot11")       # sigma = 11
    z = np.load(__import__("os").path.join(__import__("os").path.dirname(__file__), "golden", "protein_filelist_transprot11.npz"))
    want = oracle.build(z["enc"], 11, 3)
    _compare(sfx.suffixerator(z["enc"], "smap", prefixlength=3, sigma=11), want, f"keyed/transprot11/levels{levels}")


@pytest.mark.parametrize("levels", ["1", "2"])
def test_tile_path_medium_buckets_dna(levels, monkeypatch):
    """Buckets of 129..1024 suffixes inside a tile are sorted by one warp with shared-memory merges
    (warp_sort_medium): all chunk counts, equal-prefix runs (fix-up), runs too long for the fix-up (handed to the
    global sorters) and specials inside the keys."""
    monkeypatch.setenv("GTB200_TILE", "2")
    monkeypatch.setenv("GTB200_TILE_LEVELS", levels)
    rng = np.random.default_rng(311)
    for n, pl in ((40_000, 4), (100_000, 4), (200_000, 4), (250_000, 5)):      # ~160 / 390 / 780 / 240 per bucket
        enc = rng.integers(0, 4, size=n, dtype=np.uint8)
        enc[rng.integers(0, n, size=n // 500)] = 254
        want = oracle.build(enc, 4, pl)
        _compare(sfx.suffixerator(enc, "dna", prefixlength=pl), want, f"tilemed/{n}/pl{pl}/levels{levels}")
    rep = np.concatenate([np.tile(rng.integers(0, 4, size=41, dtype=np.uint8), 60), rng.integers(0, 4, size=90_000, dtype=np.uint8),
                          np.tile(rng.integers(0, 4, size=7, dtype=np.uint8), 50), rng.integers(0, 4, size=20_000, dtype=np.uint8)]).astype(np.uint8)
    want = oracle.build(rep, 4, 4)
    _compare(sfx.suffixerator(rep, "dna", prefixlength=4, parts=2), want, f"tilemed/runs/levels{levels}")


@pytest.mark.parametrize("doubling", ["0", "1"])
@pytest.mark.parametrize("case", ["repeat_rich", "duplicate", "genome_like", "polyT", "maxbucket0"])
def test_refinement_modes(case, doubling, monkeypatch):
    """Deep repeats through both refinement strategies: prefix doubling on ranks + Phi/Kasai LCP (whole-array
    parts) and text-window rounds (what sharded parts use)."""
    monkeypatch.setenv("GTB200_DOUBLING", doubling)
    if case == "repeat_rich":
        enc, pl = synth.repeat_rich_dna(unit=6000, copies=9, seed=61), None
    elif case == "duplicate":
        g = load_golden_by_name("duplicate")
        enc, pl = g["enc"], g["prefixlength"]
    elif case == "genome_like":
        enc, pl = synth.genome_like_dna(600_000, seed=62), None
    elif case == "polyT":
        enc, pl = np.concatenate([np.full(9000, 3, dtype=np.uint8), synth.random_dna(5000, seed=63),
                                  np.full(7000, 3, dtype=np.uint8)]), 5
    else:
        monkeypatch.setenv("GTB200_MAXBUCKET", "0")     # every bucket reaches refinement unsorted
        enc, pl = synth.genome_like_dna(200_000, seed=64), None
    want = oracle.build(enc, 4, pl)
    got = sfx.suffixerator(enc, "dna", prefixlength=want["prefixlength"])
    _compare(got, want, f"refine/{case}/doubling{doubling}")
    assert got["fallback_suffixes"] > 0


@pytest.mark.parametrize("extend", ["2", "1", "0"])
def test_long_exact_repeat_in_sharded_parts(extend, monkeypatch):
    """A 100 kb exact duplication (one copy cut by a separator) and a 3 kb unit in 5 tandem copies, sorted in parts:
    parts cannot use rank doubling, so their text-window rounds either jump over the windows a whole group shares
    (GTB200_EXTEND=1, k_fb_extend_count) or order every member by where it leaves a pivot member of its group
    (GTB200_EXTEND=2, k_fb_keys_pivot); GTB200_EXTEND=0 is the round-per-28-bases path they replace.  Same bytes every
    way; the number of rounds drops from repeat length / 28 to a handful per part."""
    monkeypatch.setenv("GTB200_EXTEND", extend)
    rng = np.random.default_rng(5)
    u = rng.integers(0, 4, size=100_000, dtype=np.uint8)
    t = rng.integers(0, 4, size=3_000, dtype=np.uint8)
    enc = np.concatenate([rng.integers(0, 4, size=50_000, dtype=np.uint8), u, rng.integers(0, 4, size=70_000, dtype=np.uint8),
                          u, np.array([254], dtype=np.uint8), rng.integers(0, 4, size=20_000, dtype=np.uint8),
                          np.tile(t, 5), np.array([255], dtype=np.uint8), rng.integers(0, 4, size=9_000, dtype=np.uint8)])
    want = oracle.build(enc, 4)
    for parts in (2, 5):
        got = sfx.suffixerator(enc, "dna", parts=parts)
        _compare(got, want, f"longrepeat/extend{extend}/parts{parts}")
        if extend != "0":
            assert got["fallback_rounds"] <= 20 * parts, got["fallback_rounds"]
        else:
            assert got["fallback_rounds"] >= 100_000 // 28
    monkeypatch.setenv("GTB200_DOUBLING", "0")          # a whole-array part on text-window rounds
    got = sfx.suffixerator(enc, "dna")
    _compare(got, want, f"longrepeat/extend{extend}/whole")
    monkeypatch.delenv("GTB200_DOUBLING")
    prot = np.concatenate([synth.random_protein(40, seed=8), synth.random_protein(40, seed=8)[:9000], synth.random_protein(30, seed=9)])
    _compare(sfx.suffixerator(prot, "protein", parts=3), oracle.build(prot, 20), f"longrepeat/extend{extend}/protein")
    # members of one group that end at different depths: a homopolymer run, a 7-mer in 3,000 tandem copies, a
    # run cut by a wildcard.  Only the pivot rounds bring these below run length / 28 rounds.
    runs = np.concatenate([rng.integers(0, 4, size=30_000, dtype=np.uint8), np.full(20_000, 3, dtype=np.uint8),
                           rng.integers(0, 4, size=30_000, dtype=np.uint8), np.tile(rng.integers(0, 4, size=7, dtype=np.uint8), 3000),
                           np.array([2, 255], dtype=np.uint8), np.full(5_000, 0, dtype=np.uint8), np.array([255], dtype=np.uint8),
                           rng.integers(0, 4, size=10_000, dtype=np.uint8)])
    want = oracle.build(runs, 4)
    for parts in (1, 3):
        if parts == 1:
            monkeypatch.setenv("GTB200_DOUBLING", "0")
        got = sfx.suffixerator(runs, "dna", parts=parts)
        _compare(got, want, f"runs/extend{extend}/parts{parts}")
        if extend == "2":
            assert got["fallback_rounds"] <= 80 * parts, got["fallback_rounds"]


def test_config2_ecoli_size_full_parity():
    """BASELINE.json configs[1]: 4.6 Mbp random DNA, single sequence, seed 42, pl = 8."""
    enc = synth.random_dna(4_600_000, seed=42)
    want = oracle.build(enc, 4)
    assert want["prefixlength"] == 8
    got = sfx.suffixerator(enc, "dna")
    _compare(got, want, "config2")
    assert got["fallback_suffixes"] == 0


def _check_properties(enc, got, samples=20000, seed=0):
    n = enc.shape[0]
    suf = np.frombuffer(got["suf"], dtype=np.uint64)
    lcp = np.frombuffer(got["lcp"], dtype=np.uint8)
    assert suf.shape[0] == n + 1 and lcp.shape[0] == n + 1
    seen = np.zeros(n + 1, dtype=bool)
    seen[suf] = True
    assert seen.all(), "suffix array is not a permutation"
    nsp = int((enc >= 254).sum())
    m = n - nsp
    assert (enc[suf[:m].astype(np.int64)] < 254).all()
    tail = suf[m:]
    assert (np.diff(tail.astype(np.int64)) > 0).all() and tail[-1] == n
    assert (lcp[m:] == 0).all() and lcp[0] == 0
    # sampled neighbour pairs: order and lcp by direct comparison
    rng = np.random.default_rng(seed)
    idx = rng.integers(1, m, size=min(samples, max(m - 1, 1)))
    sp = np.concatenate([np.flatnonzero(enc >= 254), [n]])
    for r in idx:
        a, b = int(suf[r - 1]), int(suf[r])
        la = int(sp[np.searchsorted(sp, a)]) - a
        lb = int(sp[np.searchsorted(sp, b)]) - b
        lim = min(la, lb, 400)
        x, y = enc[a:a + lim], enc[b:b + lim]
        neq = np.flatnonzero(x != y)
        l = int(neq[0]) if neq.size else lim
        if l < lim:
            assert x[l] < y[l], f"order violated at {r}"
        elif lim < 400:
            assert la > lb or (la == lb and a < b), f"special order violated at {r}"
        if lim < 400 or l < lim:
            assert lcp[r] == min(l, 255), f"lcp wrong at {r}: {lcp[r]} vs {l}"


def test_config3_shape_scaled_properties_and_parts():
    """configs[2] shape (24 sequences, N-runs) at 20 Mbp: too slow for the naive oracle in a test,
    so check order/LCP properties and that sharding by bucket range is output-neutral."""
    enc = synth.multi_dna_with_nruns(20_000_000, nseq=24, seed=250)
    one = sfx.suffixerator(enc, "dna", parts=1)
    _check_properties(enc, one)
    eight = sfx.suffixerator(enc, "dna", parts=8)
    for ext in ("suf", "lcp", "llv", "bck"):
        assert one[ext] == eight[ext], f"parts=8 changes .{ext}"
    for k in ("longest", "maxbranchdepth", "largelcpvalues", "averagelcp"):
        assert one[k] == eight[k]


def test_device_resident_input_matches_host_input():
    """GTB200_MEM_DEVICE descriptors (what a rank holds after the NCCL broadcast): same bytes out."""
    import torch
    enc = synth.multi_dna_with_nruns(200_000, nseq=4, seed=41)
    want = oracle.build(enc, 4)
    tw = torch.from_numpy(encseq.twobit_pack(enc).view(np.int64)).cuda()
    rg = torch.from_numpy(encseq.special_ranges(enc).view(np.int64).reshape(-1)).cuda()
    it = sfx.Sfxiterator(4, enc.shape[0], None, 2, twobit=tw.data_ptr(), ranges=rg.data_ptr(), numofwords=tw.numel(),
                         numofranges=rg.numel() // 2, device_ptrs=True)
    sufs, lcps = [], []
    while True:
        r = it.next()
        if r is None:
            break
        sufs.append(r[0])
        lcps.append(r[1] if r[1] is not None else np.zeros(r[0].shape[0], dtype=np.uint8))
    it.delete()
    assert np.concatenate(sufs).tobytes() == want["suf"]
    assert np.concatenate(lcps).tobytes() == want["lcp"]
    # byte-encoded protein resident on the device
    prot = synth.random_protein(300, seed=42)
    wantp = oracle.build(prot, 20, 3)
    pe = torch.from_numpy(prot).cuda()
    it = sfx.Sfxiterator(20, prot.shape[0], 3, 1, enc=pe.data_ptr(), device_ptrs=True)
    suf, lcp, llv, _ = it.next()
    tail = it.next()[0]
    it.delete()
    assert np.concatenate([suf, tail]).tobytes() == wantp["suf"]


def test_medium_buckets_block_sorter():
    """Buckets of 129..1024 suffixes take the block-level sorter (k_sort_medium), not the refinement path: all
    three chunk widths (buckets around 200, 400 and 900 suffixes), equal 20-bit prefixes (the fix-up pass) and
    larger buckets next to them (refinement)."""
    rng = np.random.default_rng(43)
    for n, tag in ((200_000, "p1024"), (100_000, "p512"), (50_000, "p256")):
        enc = rng.integers(0, 4, size=n, dtype=np.uint8)
        want = oracle.build(enc, 4, 4)        # 256 buckets of ~780 / ~390 / ~195 suffixes
        got = sfx.suffixerator(enc, "dna", prefixlength=4)
        _compare(got, want, f"medium/dna/{tag}")
        assert got["fallback_suffixes"] == 0
    # runs of equal prefixes: a tandem repeat puts many suffixes with the same next 10 bases into one bucket
    rep = np.concatenate([np.tile(rng.integers(0, 4, size=37, dtype=np.uint8), 40), rng.integers(0, 4, size=60_000, dtype=np.uint8),
                          np.tile(rng.integers(0, 4, size=11, dtype=np.uint8), 30)]).astype(np.uint8)
    want = oracle.build(rep, 4, 4)
    got = sfx.suffixerator(rep, "dna", prefixlength=4)
    _compare(got, want, "medium/dna/runs")
    enc = rng.integers(0, 4, size=600_000, dtype=np.uint8)
    want = oracle.build(enc, 4, 4)            # ~2300 per bucket: beyond the block sorter, refined
    got = sfx.suffixerator(enc, "dna", prefixlength=4)
    _compare(got, want, "medium/dna/large")
    prot = synth.random_protein(1500, seed=44)
    want = oracle.build(prot, 20, 2)          # 400 buckets of ~1200 residues, non-uniform: both sides of the limit
    got = sfx.suffixerator(prot, "protein", prefixlength=2, parts=2)
    _compare(got, want, "medium/protein")
    prot = synth.random_protein(600, seed=45)
    want = oracle.build(prot, 20, 2)          # ~500 per bucket
    got = sfx.suffixerator(prot, "protein", prefixlength=2)
    _compare(got, want, "medium/protein500")


def test_sharded_count_equals_replicated_count():
    """gtb200_count_shard on two position shards (as two ranks would run it) + summed tables + gtb200_count_finish
    gives the bucket tables of gtb200_count; the sort that follows is unchanged."""
    import ctypes as C
    import torch

    class DevArray:
        def __init__(self, ptr, nwords):
            self.__cuda_array_interface__ = {"shape": (nwords,), "typestr": "<i4", "data": (ptr, False), "version": 3}

    enc = synth.multi_dna_with_nruns(700_000, nseq=5, seed=91)
    want = oracle.build(enc, 4)
    n = enc.shape[0]
    L = sfx.load_library()
    its = [sfx.Sfxiterator(4, n, None, 1, twobit=encseq.twobit_pack(enc), ranges=encseq.special_ranges(enc)) for _ in range(2)]
    cut = (n // 2 + 63) // 64 * 64
    tabs = []
    for r, (lo, hi) in enumerate([(0, cut), (cut, n)]):
        ptr, nw = C.c_void_p(), C.c_uint64()
        sfx._check(L.gtb200_count_shard(its[r].ctx, lo, hi, 1 if r == 0 else 0, C.byref(ptr), C.byref(nw)))
        tabs.append(torch.as_tensor(DevArray(ptr.value, nw.value), device="cuda"))
    torch.cuda.synchronize()
    total = tabs[0] + tabs[1]                  # what the all-reduce leaves on every rank
    for r in range(2):
        tabs[r].copy_(total)
    torch.cuda.synchronize()
    for r in range(2):
        lb = np.zeros(want["leftborder"].shape[0], dtype=np.uint32)
        csc = np.zeros(want["countspecialcodes"].shape[0], dtype=np.uint32)
        dst = np.zeros(max(1, want["distpfxidx"].shape[0]), dtype=np.uint32)
        info = sfx.CountInfo()
        sfx._check(L.gtb200_count_finish(its[r].ctx, lb.ctypes.data, csc.ctypes.data, dst.ctypes.data, C.byref(info)))
        assert np.array_equal(lb, want["leftborder"]) and np.array_equal(csc, want["countspecialcodes"])
        assert np.array_equal(dst[:want["distpfxidx"].shape[0]], want["distpfxidx"])
        assert info.numofsuffixestosort == want["nonspecial"]
    # rank 1's context sorts the second half of the buckets from the summed tables
    parts = sfx.suftabparts(2, want["leftborder"], want["prefixlength"])
    lo, hi, width = parts[1]
    pi = sfx.PartInfo()
    suf = np.empty(width, dtype=np.uint64)
    lcp = np.empty(width, dtype=np.uint8)
    sfx._check(L.gtb200_sort_part(its[1].ctx, lo, hi, suf.ctypes.data, 0, lcp.ctypes.data, None, 0, C.byref(pi)))
    off = int(pi.partoffset)
    assert np.array_equal(suf, want["suf_array"][off:off + width])
    assert lcp.tobytes() == want["lcp"][off:off + width]
    for it in its:
        it.delete()


@pytest.mark.parametrize("case", ["dna_nruns_300k", "dna_nruns_300k_pl3", "dna_tiny_special_ends", "dna_all_special", "dna_single", "dna_polyT"])
def test_funnel_shift_counting_kernel(case, monkeypatch):
    """k_count_dna (16 positions per thread, codes by funnel shifts; what genomes of >= 32 M positions use) forced on
    small inputs: windows with specials, sequence ends inside a thread's positions, sharded counting (parts)."""
    monkeypatch.setenv("GTB200_FASTCOUNT", "2")
    enc, alphabet, pl = CASES[case]()
    want = oracle.build(enc, encseq.alphabet_size(alphabet), pl)
    for parts in (1, 3):
        _compare(sfx.suffixerator(enc, alphabet, parts=parts, prefixlength=want["prefixlength"]), want, f"fastcount/{case}/{parts}")


@pytest.mark.parametrize("narrow", ["1", "0"])
def test_narrow_transfer_matches(narrow, monkeypatch):
    """A GtUword suftab of a large part crosses PCIe as 32-bit positions and is widened by host threads
    (GTB200_NARROW=1, the default) - same bytes as copying the 64-bit image, also when refinement redoes the part
    and with an odd thread count."""
    monkeypatch.setenv("GTB200_NARROW", "2" if narrow == "1" else "0")     # 2: also for parts below 32 M suffixes
    monkeypatch.setenv("GTB200_WIDEN_THREADS", "3")
    enc = synth.multi_dna_with_nruns(6_000_000, nseq=5, seed=191)
    _compare(sfx.suffixerator(enc, "dna"), oracle.build(enc, 4), f"narrow{narrow}/plain")
    rep = synth.genome_like_dna(5_500_000, seed=192)
    got = sfx.suffixerator(rep, "dna", parts=1)
    _compare(got, oracle.build(rep, 4), f"narrow{narrow}/repeats")
    assert got["fallback_suffixes"] > 0


@pytest.mark.parametrize("pipeline", ["0", "1"])
def test_pipelined_fetch_matches(pipeline, monkeypatch):
    """gtb200_sort_part cuts a large part into code ranges and copies finished slices while the next range is
    sorted; with and without that overlap the host buffers hold the same bytes (also when refinement has to
    redo the part afterwards)."""
    monkeypatch.setenv("GTB200_PIPELINE", pipeline)
    enc = synth.multi_dna_with_nruns(5_000_000, nseq=8, seed=95)
    want = oracle.build(enc, 4)
    got = sfx.suffixerator(enc, "dna")
    _compare(got, want, f"pipeline{pipeline}/plain")
    rep = synth.genome_like_dna(3_000_000, seed=96)
    want = oracle.build(rep, 4)
    got = sfx.suffixerator(rep, "dna", parts=1)
    _compare(got, want, f"pipeline{pipeline}/repeats")
    assert got["fallback_suffixes"] > 0


@pytest.mark.parametrize("fused", ["0", "1"])
def test_suf_image_from_sorters(fused, monkeypatch):
    """On the direct-insertion path the sorters write the 64-bit .suf image themselves (GTB200_FUSED_WIDEN=1,
    the default) instead of a separate widening pass: same bytes and the same `longest`, with one part, with
    parts, with medium buckets (pl forced down) and when refinement has to redo the image."""
    monkeypatch.setenv("GTB200_FUSED_WIDEN", fused)
    enc = synth.multi_dna_with_nruns(700_000, nseq=5, seed=201)
    want = oracle.build(enc, 4)
    for parts in (1, 3):
        got = sfx.suffixerator(enc, "dna", parts=parts)
        _compare(got, want, f"fused{fused}/parts{parts}")
        assert got["longest"] == int(np.flatnonzero(want["suf_array"] == 0)[0])
    wantm = oracle.build(enc, 4, 4)                       # ~2700 suffixes per bucket: the block sorter finishes them
    gotm = sfx.suffixerator(enc, "dna", prefixlength=4)
    _compare(gotm, wantm, f"fused{fused}/medium")
    rep = synth.repeat_rich_dna(200_000, seed=202)
    wantr = oracle.build(rep, 4)
    gotr = sfx.suffixerator(rep, "dna")
    _compare(gotr, wantr, f"fused{fused}/repeats")
    assert gotr["fallback_suffixes"] > 0


@pytest.mark.parametrize("tile", ["1", "2"])
@pytest.mark.parametrize("gather", ["0", "2"])
def test_gather_table_modes(gather, tile, monkeypatch):
    """Keys fetched through the 16-byte gather table (one load, bitmap only for records that hold a special) or
    through two text words + bitmap (table not built): same index, on the direct and on the tile path, with
    specials at every offset of a record, for DNA, protein and a reversed readmode."""
    monkeypatch.setenv("GTB200_GATHER", gather)
    monkeypatch.setenv("GTB200_TILE", tile)
    rng = np.random.default_rng(77)
    enc = rng.integers(0, 4, size=120_000, dtype=np.uint8)
    enc[rng.integers(0, enc.shape[0], size=3000)] = 254          # single wildcards everywhere
    enc[5000:5070] = 254
    enc[np.arange(20_000, 26_000, 97)] = 255                     # separators at drifting record offsets
    want = oracle.build(enc, 4)
    _compare(sfx.suffixerator(enc, "dna", parts=2), want, f"gather{gather}/tile{tile}/dna")
    wantr = oracle.build(_apply_readmode(enc, 3), 4)
    _compare(sfx.suffixerator(enc, "dna", readmode=3), wantr, f"gather{gather}/tile{tile}/rcl")
    prot = synth.random_protein(800, seed=78)
    wantp = oracle.build(prot, 20, 3)
    _compare(sfx.suffixerator(prot, "protein", prefixlength=3), wantp, f"gather{gather}/tile{tile}/protein")
    rep = synth.repeat_rich_dna(unit=3000, copies=7, seed=79)
    _compare(sfx.suffixerator(rep, "dna"), oracle.build(rep, 4), f"gather{gather}/tile{tile}/repeats")


@pytest.mark.parametrize("overlap", ["0", "1"])
def test_padded_suffixes_beside_the_main_stream(overlap, monkeypatch):
    """The list of special-padded suffixes is built beside the k-mer count and placed beside the insertion on a
    second stream (GTB200_OVERLAP=1, the default): same index with and without, on inputs where those suffixes
    are many (a special every few bases), with parts, repeatedly on one context."""
    monkeypatch.setenv("GTB200_OVERLAP", overlap)
    rng = np.random.default_rng(310)
    enc = rng.integers(0, 4, size=200_000, dtype=np.uint8)
    enc[rng.integers(0, enc.shape[0], size=12_000)] = 254
    enc[rng.integers(0, enc.shape[0], size=300)] = 255
    want = oracle.build(enc, 4)
    for parts in (1, 3):
        _compare(sfx.suffixerator(enc, "dna", parts=parts), want, f"overlap{overlap}/parts{parts}")
    import ctypes as C
    L = sfx.load_library()
    it = sfx.Sfxiterator(4, enc.shape[0], None, 1, enc=enc)           # three passes on the same context
    width = it.parts[0][2]
    suf = np.empty(width, dtype=np.uint64)
    lcp = np.empty(width, dtype=np.uint8)
    for _ in range(3):
        suf[:] = 0
        sfx._check(L.gtb200_count(it.ctx, None, None, None, None))
        pi = sfx.PartInfo()
        sfx._check(L.gtb200_sort_part(it.ctx, it.parts[0][0], it.parts[0][1], suf.ctypes.data, 0, lcp.ctypes.data, None, 0,
                                      C.byref(pi)))
        assert np.array_equal(suf, want["suf_array"][:width]) and lcp.tobytes() == want["lcp"][:width]
    it.delete()


def _apply_readmode(enc, readmode):
    e = enc[::-1].copy() if readmode in (1, 3) else enc.copy()
    return np.where(e < 4, 3 - e, e).astype(np.uint8) if readmode in (2, 3) else e


@pytest.mark.parametrize("readmode", [1, 2, 3])
@pytest.mark.parametrize("use_twobit", [True, False])
def test_readmodes(readmode, use_twobit):
    """-dir rev / cpl / rcl (GtReadmode 1/2/3): the index of the transformed sequence (checked against the
    reference in tests/test_oracle.py); the caller still hands over the forward sequence."""
    enc = synth.multi_dna_with_nruns(300_000, nseq=6, seed=97)
    want = oracle.build(_apply_readmode(enc, readmode), 4)
    got = sfx.suffixerator(enc, "dna", parts=2, readmode=readmode, use_twobit=use_twobit)
    _compare(got, want, f"readmode{readmode}/twobit{use_twobit}")
    if readmode == 1:
        prot = synth.random_protein(300, seed=98)
        wantp = oracle.build(prot[::-1].copy(), 20, 3)
        gotp = sfx.suffixerator(prot, "protein", prefixlength=3, readmode=1)
        _compare(gotp, wantp, "readmode1/protein")
    else:
        with pytest.raises(sfx.Gtb200Error):
            sfx.suffixerator(synth.random_protein(50, seed=99), "protein", prefixlength=2, readmode=readmode)


def test_suftabuint_and_tail_paging():
    enc = synth.multi_dna_with_nruns(100_000, nseq=5, seed=21)
    want = oracle.build(enc, 4)
    it = sfx.Sfxiterator(4, enc.shape[0], None, 2, twobit=encseq.twobit_pack(enc), ranges=encseq.special_ranges(enc),
                         suftabuint=True)
    chunks = []
    while True:
        r = it.next()
        if r is None:
            break
        assert r[0].dtype == np.uint32
        chunks.append(r[0])
    assert it.longest() == want["longest"]
    it.delete()
    assert np.array_equal(np.concatenate(chunks).astype(np.uint64), want["suf_array"])

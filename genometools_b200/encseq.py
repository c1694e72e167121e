"""Host-side view of the *input* of the enhanced-suffix-array path.

The product never re-implements GtEncseq ("input format - unchanged").
Inside libgenometools the two things the device layer needs are obtained from the existing
accessors:

* ``gt_encseq_twobitencoding_export`` (src/core/encseq.c:6697) - the packed 2-bit words, base ``i`` in
  bits ``62-2*(i%32)`` of word ``i/32`` (src/core/intbits.h:78-84), and
* ``gt_specialrangeiterator_new/next`` (src/core/encseq.c:3383,3745) - the maximal runs of
  wildcards/separators as half-open ``GtRange``.

The tests, the benchmark and the synthetic-genome generators have no libgenometools to ask, so
this module produces the same two arrays with numpy from FASTA text or from an array of symbol
codes.  Symbol codes follow src/core/alphabet.c:84-91 (DNA ``acgt`` -> 0..3, ``u`` -> 3, IUPAC
ambiguity letters -> wildcard; protein ``LVIFKREDAGSTNQYWPHMC`` -> 0..19, ``XUBZJO*-`` -> wildcard)
and src/core/chardef_api.h:34-58 (``GT_WILDCARD`` = 254, ``GT_SEPARATOR`` = 255).
"""
from __future__ import annotations

import numpy as np

WILDCARD = 254
SEPARATOR = 255

_DNA_MAP = np.full(256, 253, dtype=np.uint8)  # 253 = undefined character
for _i, _c in enumerate("acgt"):
    _DNA_MAP[ord(_c)] = _i
    _DNA_MAP[ord(_c.upper())] = _i
_DNA_MAP[ord("u")] = 3
_DNA_MAP[ord("U")] = 3
for _c in "nsywrkvbdhm":
    _DNA_MAP[ord(_c)] = WILDCARD
    _DNA_MAP[ord(_c.upper())] = WILDCARD

_PROT_MAP = np.full(256, 253, dtype=np.uint8)
for _i, _c in enumerate("LVIFKREDAGSTNQYWPHMC"):
    _PROT_MAP[ord(_c)] = _i
for _c in "XUBZJO*-":
    _PROT_MAP[ord(_c)] = WILDCARD


def alphabet_size(alphabet: str) -> int:
    return {"dna": 4, "protein": 20}[alphabet]


def read_fasta(path: str) -> list[bytes]:
    """Sequences of a (multi-)FASTA file, description lines dropped."""
    with open(path, "rb") as fh:
        data = fh.read()
    seqs: list[bytes] = []
    for rec in data.split(b">")[1:]:
        nl = rec.find(b"\n")
        body = rec[nl + 1:] if nl >= 0 else b""
        seqs.append(body.replace(b"\n", b"").replace(b"\r", b"").replace(b" ", b""))
    return seqs


def encode_sequences(seqs: list[bytes], alphabet: str = "dna") -> np.ndarray:
    """Concatenate sequences with exactly one separator between them (none after the last) and
    map characters to codes: the text model of SURVEY.md section 9."""
    table = _DNA_MAP if alphabet == "dna" else _PROT_MAP
    parts = []
    for k, s in enumerate(seqs):
        if k:
            parts.append(np.array([SEPARATOR], dtype=np.uint8))
        arr = table[np.frombuffer(s, dtype=np.uint8)]
        if (arr == 253).any():
            bad = s[int(np.argmax(arr == 253))]
            raise ValueError(f"illegal character {chr(bad)!r} for alphabet {alphabet}")
        parts.append(arr)
    if not parts:
        return np.zeros(0, dtype=np.uint8)
    return np.concatenate(parts)


def encode_fasta(path: str, alphabet: str = "dna") -> np.ndarray:
    return encode_sequences(read_fasta(path), alphabet)


def special_ranges(enc: np.ndarray) -> np.ndarray:
    """Maximal runs of special symbols as half-open [start, end) pairs, ascending: what
    gt_specialrangeiterator_next yields in forward direction."""
    pos = np.flatnonzero(enc >= WILDCARD)
    if pos.size == 0:
        return np.zeros((0, 2), dtype=np.uint64)
    brk = np.flatnonzero(np.diff(pos) != 1)
    starts = np.concatenate([pos[:1], pos[brk + 1]])
    ends = np.concatenate([pos[brk], pos[-1:]]) + 1
    return np.stack([starts, ends], axis=1).astype(np.uint64)


def twobit_pack(enc: np.ndarray, chunk: int = 1 << 26) -> np.ndarray:
    """GtTwobitencoding words of a DNA sequence (src/core/encseq.c:73-81).  Special positions get
    the 2-bit filler 0; like in the reference they must never be interpreted without the
    special-range table.  One spare word is appended, as gt_unitsoftwobitencoding does
    (src/core/intbits.h:194-205).  Works in chunks so that a 3.1 Gbp genome needs no large temporaries."""
    n = int(enc.shape[0])
    nwords = (n + 31) // 32 + 1
    out = np.zeros(nwords, dtype=np.uint64)
    for lo in range(0, n, chunk):          # chunk is a multiple of 32
        hi = min(n, lo + chunk)
        c = enc[lo:hi]
        c = np.where(c >= WILDCARD, 0, c).astype(np.uint8)
        pad = (-c.shape[0]) % 32
        if pad:
            c = np.concatenate([c, np.zeros(pad, dtype=np.uint8)])
        q = c.reshape(-1, 4)
        by = (q[:, 0] << 6) | (q[:, 1] << 4) | (q[:, 2] << 2) | q[:, 3]     # 4 bases per byte, first base on top
        words = by.view(">u8").astype(np.uint64)                             # 8 bytes big-endian = 32 bases
        out[lo // 32: lo // 32 + words.shape[0]] = words
    return out


def write_fasta(path: str, seqs: list[bytes], width: int = 80) -> None:
    with open(path, "wb") as fh:
        for k, s in enumerate(seqs):
            fh.write(b">seq%d\n" % k)
            for i in range(0, len(s), width):
                fh.write(s[i:i + width])
                fh.write(b"\n")


def decode_to_fasta_bytes(enc: np.ndarray, alphabet: str = "dna") -> list[bytes]:
    """Inverse of encode_sequences (wildcards become N / X) - used to hand a synthetic input to the
    reference binary, which only reads FASTA."""
    if alphabet == "dna":
        lut = np.zeros(256, dtype=np.uint8)
        lut[:4] = np.frombuffer(b"ACGT", dtype=np.uint8)
        lut[WILDCARD] = ord("N")
    else:
        lut = np.zeros(256, dtype=np.uint8)
        lut[:20] = np.frombuffer(b"LVIFKREDAGSTNQYWPHMC", dtype=np.uint8)
        lut[WILDCARD] = ord("X")
    seps = np.flatnonzero(enc == SEPARATOR)
    bounds = np.concatenate([[-1], seps, [enc.shape[0]]])
    out = []
    for a, b in zip(bounds[:-1], bounds[1:]):
        out.append(lut[enc[a + 1:b]].tobytes())
    return out

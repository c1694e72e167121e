"""Deterministic synthetic inputs of the shapes BASELINE.json names (recipes: SURVEY.md section 8(d)).

All generators return the *encoded* sequence (uint8 symbol codes, 254 wildcard, 255 separator),
i.e. what GtEncseq would hold after ``gt_encseq_encoder_encode``; ``encseq.decode_to_fasta_bytes``
turns it back into FASTA for the reference binary.
"""
from __future__ import annotations

import numpy as np

from .encseq import SEPARATOR, WILDCARD


def random_dna(n: int, seed: int = 42) -> np.ndarray:
    """Config 2 shape: one sequence, i.i.d. uniform ACGT (4.6 Mbp, seed 42 in BASELINE.json)."""
    rng = np.random.default_rng(seed)
    return rng.integers(0, 4, size=n, dtype=np.uint8)


def multi_dna_with_nruns(total: int, nseq: int = 24, seed: int = 250, lead_n: int = 10_000,
                         runs_per_seq: int = 20, max_run: int = 50_000, singles: int = 200) -> np.ndarray:
    """Config 3/4 shape: ``nseq`` sequences (multinomial split of ``total`` bases), uniform ACGT, each
    with a leading N-run, ``runs_per_seq`` N-runs of U[1,max_run] and ``singles`` isolated N.  Run
    sizes are scaled down for small totals so that wildcards stay a ~5 % minority."""
    rng = np.random.default_rng(seed)
    lens = rng.multinomial(total, np.full(nseq, 1.0 / nseq))
    scale = min(1.0, total / 250_000_000)
    lead = max(1, int(lead_n * scale))
    mrun = max(1, int(max_run * scale))
    nsingle = max(1, int(singles * max(scale, 0.01)))
    out = np.empty(int(lens.sum()) + nseq - 1, dtype=np.uint8)
    pos = 0
    for k, ln in enumerate(lens):
        ln = int(ln)
        if k:
            out[pos] = SEPARATOR
            pos += 1
        s = rng.integers(0, 4, size=ln, dtype=np.uint8)
        s[:min(lead, ln)] = WILDCARD
        for _ in range(runs_per_seq):
            if ln < 2:
                break
            a = int(rng.integers(0, ln))
            b = min(ln, a + int(rng.integers(1, mrun + 1)))
            s[a:b] = WILDCARD
        if ln:
            s[rng.integers(0, ln, size=nsingle)] = WILDCARD
        out[pos:pos + ln] = s
        pos += ln
    return out


def random_protein(nseq: int, mean_len: float = 333.0, seed: int = 5, xfrac: float = 1e-4) -> np.ndarray:
    """Config 5 shape: many short sequences (geometric lengths), residues from a fixed non-uniform
    background over the 20 amino-acid codes, a 0.01 % sprinkle of wildcards."""
    rng = np.random.default_rng(seed)
    lens = rng.geometric(1.0 / mean_len, size=nseq).astype(np.int64)
    bg = np.linspace(2.0, 1.0, 20)
    bg /= bg.sum()
    total = int(lens.sum()) + nseq - 1
    body = rng.choice(20, size=total, p=bg).astype(np.uint8)
    body[rng.random(total) < xfrac] = WILDCARD
    ends = np.cumsum(lens + 1)[:-1] - 1
    body[ends] = SEPARATOR
    return body


def repeat_rich_dna(unit: int = 700, copies: int = 6, seed: int = 7, mutate: float = 0.002,
                    with_specials: bool = True) -> np.ndarray:
    """Stress case outside BASELINE.json: tandem and dispersed near-exact repeats, so that many
    LCP values exceed 255 (.llv non-empty) and the prefix-doubling refinement is exercised."""
    rng = np.random.default_rng(seed)
    u = rng.integers(0, 4, size=unit, dtype=np.uint8)
    pieces = []
    for c in range(copies):
        v = u.copy()
        flips = rng.random(unit) < mutate
        v[flips] = rng.integers(0, 4, size=int(flips.sum()), dtype=np.uint8)
        pieces.append(v)
        if with_specials and c % 2 == 1:
            pieces.append(np.array([WILDCARD] * int(rng.integers(1, 4)), dtype=np.uint8))
        if with_specials and c == copies // 2:
            pieces.append(np.array([SEPARATOR], dtype=np.uint8))
    pieces.append(np.full(600, 3, dtype=np.uint8))          # poly-T
    pieces.append(np.tile(np.array([0, 1], dtype=np.uint8), 400))  # (AC)n
    return np.concatenate(pieces)


def genome_like_dna(total: int, seed: int = 77, repeat_fraction: float = 0.35, nseq: int = 6) -> np.ndarray:
    """Stress case closer to a real genome than i.i.d. bases: interspersed repeat families (300 bp
    to 6 kb consensus sequences copied with 0-15 % divergence), a few recent exact segmental
    duplications (10-50 kb), microsatellites, N-runs and several sequences.  Produces many buckets
    with ties beyond pl+28 bases and LCP values in the thousands."""
    rng = np.random.default_rng(seed)
    out = rng.integers(0, 4, size=total, dtype=np.uint8)
    fams = [rng.integers(0, 4, size=int(L), dtype=np.uint8) for L in rng.choice([300, 300, 300, 1200, 6000], size=12)]
    budget = int(total * repeat_fraction)
    while budget > 0:
        f = fams[int(rng.integers(0, len(fams)))]
        div = float(rng.choice([0.0, 0.01, 0.05, 0.15]))
        c = f.copy()
        if div > 0:
            m = rng.random(c.shape[0]) < div
            c[m] = rng.integers(0, 4, size=int(m.sum()), dtype=np.uint8)
        a = int(rng.integers(0, max(1, total - c.shape[0])))
        out[a:a + c.shape[0]] = c[: total - a]
        budget -= c.shape[0]
    for _ in range(4):                                   # segmental duplications (exact)
        L = int(rng.integers(10_000, 50_000))
        if total > 4 * L:
            a, b = int(rng.integers(0, total - L)), int(rng.integers(0, total - L))
            out[b:b + L] = out[a:a + L]
    for _ in range(max(1, total // 200_000)):            # microsatellites
        unit = rng.integers(0, 4, size=int(rng.integers(1, 5)), dtype=np.uint8)
        L = int(rng.integers(40, 400))
        a = int(rng.integers(0, max(1, total - L)))
        out[a:a + L] = np.resize(unit, L)[: total - a]
    for _ in range(max(1, total // 1_000_000)):          # N-runs
        L = int(rng.integers(1, 5000))
        a = int(rng.integers(0, max(1, total - L)))
        out[a:a + L] = WILDCARD
    cuts = np.sort(rng.choice(np.arange(1, total - 1), size=nseq - 1, replace=False))
    out[cuts] = SEPARATOR
    return out

"""TEST INFRASTRUCTURE ONLY: Python face of the oracle.

* ``build(enc, sigma, pl)`` runs the plain-C restatement (oracle/sfx_oracle.c) and lays the result out
  as the bytes of the reference's on-disk files (.suf / .lcp / .llv / .bck).
* ``run_reference(fasta, alphabet, ...)`` runs the UNMODIFIED reference binary oracle/_ref/gt_ref
  (built by oracle/Makefile from /root/reference) and returns the files it wrote.

Nothing under genometools_b200/ imports this module.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import tempfile
import time

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
GT_REF = os.path.join(_HERE, "_ref", "gt_ref")
_LIB = None


class _Stats(ctypes.Structure):
    _fields_ = [("nonspecial", ctypes.c_uint64), ("longest", ctypes.c_uint64),
                ("maxbranchdepth", ctypes.c_uint64), ("numlargelcp", ctypes.c_uint64),
                ("lcpsum", ctypes.c_double)]


def lib() -> ctypes.CDLL:
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "libsfx_oracle.so")
        src = os.path.join(_HERE, "sfx_oracle.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.check_call(["gcc", "-O2", "-Wall", "-shared", "-fPIC", "-o", so, src])
        _LIB = ctypes.CDLL(so)
        _LIB.sfxo_recommendedprefixlength.restype = ctypes.c_uint
        _LIB.sfxo_recommendedprefixlength.argtypes = [ctypes.c_uint, ctypes.c_uint64]
        _LIB.sfxo_numofdistpfxidxcounters.restype = ctypes.c_uint64
        _LIB.sfxo_numofdistpfxidxcounters.argtypes = [ctypes.c_uint, ctypes.c_uint]
        _LIB.sfxo_build.restype = ctypes.c_int
    return _LIB


def recommended_prefixlength(sigma: int, totallength: int) -> int:
    return int(lib().sfxo_recommendedprefixlength(sigma, totallength))


def _pad8(b: bytes) -> bytes:
    return b + b"\0" * (-len(b) % 8)


def build(enc: np.ndarray, sigma: int, pl: int | None = None) -> dict:
    """Returns dict with the byte images of .suf/.lcp/.llv/.bck, the raw arrays and the .prj
    statistics for the encoded sequence ``enc`` (uint8 codes, 254/255 specials)."""
    L = lib()
    enc = np.ascontiguousarray(enc, dtype=np.uint8)
    n = int(enc.shape[0])
    if pl is None:
        pl = recommended_prefixlength(sigma, n)
    numcodes = sigma ** pl
    suf = np.zeros(n + 1, dtype=np.uint64)
    lcp = np.zeros(n + 1, dtype=np.uint32)
    leftborder = np.zeros(numcodes + 1, dtype=np.uint32)
    csc = np.zeros(sigma ** (pl - 1), dtype=np.uint32)
    ndist = int(L.sfxo_numofdistpfxidxcounters(sigma, pl))
    dist = np.zeros(max(ndist, 1), dtype=np.uint32)
    st = _Stats()
    rc = L.sfxo_build(enc.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint64(n), ctypes.c_uint(sigma),
                      ctypes.c_uint(pl), suf.ctypes.data_as(ctypes.c_void_p),
                      lcp.ctypes.data_as(ctypes.c_void_p), leftborder.ctypes.data_as(ctypes.c_void_p),
                      csc.ctypes.data_as(ctypes.c_void_p), dist.ctypes.data_as(ctypes.c_void_p),
                      ctypes.byref(st))
    if rc != 0:
        raise MemoryError("sfxo_build failed")
    dist = dist[:ndist]
    large = np.flatnonzero(lcp >= 255)
    llv = np.stack([large.astype(np.uint64), lcp[large].astype(np.uint64)], axis=1)
    bck = _pad8(leftborder.tobytes()) + _pad8(csc.tobytes())
    if pl > 2:
        bck += _pad8(dist.tobytes())
    return {
        "n": n, "prefixlength": pl,
        "suf": suf.tobytes(), "lcp": np.minimum(lcp, 255).astype(np.uint8).tobytes(),
        "llv": llv.tobytes(), "bck": bck,
        "suf_array": suf, "lcp_array": lcp, "leftborder": leftborder,
        "countspecialcodes": csc, "distpfxidx": dist,
        "longest": int(st.longest), "maxbranchdepth": int(st.maxbranchdepth),
        "largelcpvalues": int(st.numlargelcp), "nonspecial": int(st.nonspecial),
        "averagelcp": "%.2f" % (st.lcpsum / (n + 1)),
    }


def have_reference() -> bool:
    return os.path.isfile(GT_REF) and os.access(GT_REF, os.X_OK)


def parse_prj(text: str) -> dict:
    out = {}
    for line in text.splitlines():
        if "=" in line:
            k, v = line.split("=", 1)
            out[k.strip()] = v.strip()
    return out


def run_reference(fasta: str | list[str], alphabet: str = "dna", parts: int = 1, pl: int | None = None,
                  extra: list[str] | None = None, keep_dir: str | None = None,
                  showtime: bool = False, drop_alphabet_flag: bool = False) -> dict:
    """`gt suffixerator -dna|-protein -suf -lcp -bck -tis` on the unmodified reference."""
    if not have_reference():
        raise RuntimeError("oracle/_ref/gt_ref is not built (run `make -C oracle`)")
    files = [fasta] if isinstance(fasta, str) else list(fasta)
    tmp = keep_dir or tempfile.mkdtemp(prefix="gtref_")
    cmd = [GT_REF, "suffixerator"] + ([] if drop_alphabet_flag else ["-" + alphabet]) + \
          ["-suf", "-lcp", "-bck", "-tis", "-v", "-indexname", os.path.join(tmp, "ref"), "-parts", str(parts)]
    if pl is not None:
        cmd += ["-pl", str(pl)]
    cmd += (extra or []) + ["-db"] + files
    env = dict(os.environ)
    if showtime:
        env["GT_ENV_OPTIONS"] = "-showtime"
    t0 = time.perf_counter()
    proc = subprocess.run(cmd, capture_output=True, text=True, env=env)
    wall = time.perf_counter() - t0
    if proc.returncode != 0:
        raise RuntimeError("reference failed: " + proc.stderr + proc.stdout)
    out = {"log": proc.stdout, "wall_s": wall, "dir": tmp}
    for ext in ("suf", "lcp", "llv", "bck"):
        with open(os.path.join(tmp, "ref." + ext), "rb") as fh:
            out[ext] = fh.read()
    with open(os.path.join(tmp, "ref.prj")) as fh:
        out["prj"] = parse_prj(fh.read())
    times = {}
    for line in proc.stdout.splitlines():
        if line.startswith("# TIME "):
            name, _, val = line[7:].rpartition(" ")
            try:
                times[name] = float(val)
            except ValueError:
                pass
    out["times"] = times
    if keep_dir is None:
        for f in os.listdir(tmp):
            os.unlink(os.path.join(tmp, f))
        os.rmdir(tmp)
    return out

"""Host-side mirror of the reference's Sfxiterator interface over the C ABI of libgtsfx_b200.so.

Inside libgenometools the host side stays C (INTEGRATION.md shows the patched
src/match/sfx-suffixer.c).  This module is the same control flow for the Python test and
benchmark harness: `Sfxiterator` has the reference's entry points with the reference's meaning

    gt_Sfxiterator_new_withadditionalvalues  (src/match/sfx-suffixer.c:1363)  -> Sfxiterator(...)
    gt_Sfxiterator_next                      (:2162)                          -> .next()
    gt_Sfxiterator_longest                   (:2219)                          -> .longest()
    gt_Sfxiterator_bcktab2file               (:2206)                          -> .bcktab2file(fp)
    gt_Sfxiterator_delete                    (:537)                           -> .delete()

and `suffixerator()` is the loop of suffixeratorwithoutput (src/match/sfx-run.c:212-317) that writes
.suf/.lcp/.llv/.bck exactly as gt_suffixsortspace_to_file, outlcpvalues and
gt_bcktab_flush_to_file do.  There is no CPU fallback: if the CUDA library or a GPU is missing the
constructor raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import encseq as _encseq

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgtsfx_b200.so")
_lib = None

UINT64_MAX = (1 << 64) - 1


class Gtb200Error(RuntimeError):
    pass


class Range(C.Structure):
    _fields_ = [("start", C.c_uint64), ("end", C.c_uint64)]


class LargeLcp(C.Structure):
    _fields_ = [("position", C.c_uint64), ("value", C.c_uint64)]


class SeqDesc(C.Structure):
    _fields_ = [("numofchars", C.c_uint), ("totallength", C.c_uint64),
                ("twobitencoding", C.c_void_p), ("numofwords", C.c_uint64),
                ("encoded", C.c_void_p), ("specialranges", C.c_void_p),
                ("numofspecialranges", C.c_uint64), ("memory", C.c_int), ("readmode", C.c_int)]


class CountInfo(C.Structure):
    _fields_ = [("numofallcodes", C.c_uint64), ("numofspecialcodes", C.c_uint64),
                ("numofdistpfxidx", C.c_uint64), ("numofsuffixestosort", C.c_uint64),
                ("specialcharacters", C.c_uint64), ("largestbucket", C.c_uint64)]


class FastaInfo(C.Structure):
    _fields_ = [("totallength", C.c_uint64), ("numofsequences", C.c_uint64), ("specialcharacters", C.c_uint64),
                ("prefixlength", C.c_uint)]


class PartInfo(C.Structure):
    _fields_ = [("partoffset", C.c_uint64), ("width", C.c_uint64),
                ("numoflargelcpvalues", C.c_uint64), ("maxbranchdepth", C.c_uint64),
                ("lcptabsum", C.c_double), ("longest", C.c_uint64),
                ("fallback_suffixes", C.c_uint64), ("fallback_rounds", C.c_uint64)]


EXPORTS = [
    "gtb200_device_count", "gtb200_last_error", "gtb200_open", "gtb200_close", "gtb200_stream",
    "gtb200_count", "gtb200_sort_part_device", "gtb200_part_device", "gtb200_fetch_part",
    "gtb200_sort_part", "gtb200_tail", "gtb200_recommendedprefixlength", "gtb200_suftabparts",
    "gtb200_profile_enable", "gtb200_profile_read", "gtb200_profile_reset",
    "gtb200_leftborder_partialsums_noalloc", "gtb200_release_cached_memory",
    "gtb200_count_shard", "gtb200_count_finish", "gtb200_part_diag",
    "gtb200_multi_open", "gtb200_multi_close", "gtb200_multi_count", "gtb200_multi_sort_part", "gtb200_multi_tail",
    "gtb200_multi_devices", "gtb200_multi_ctx", "gtb200_multi_peer_bytes", "gtb200_host_register",
    "gtb200_host_unregister", "gtb200_multi_free_memory", "gtb200_tail_device", "gtb200_open_fasta", "gtb200_split_range", "gtb200_parse_devices", "gtb200_widen_u32_u64",
]


def load_library() -> C.CDLL:
    """dlopen the in-tree CUDA library; a missing build is a hard error, never a fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise Gtb200Error(f"{LIB_PATH} is not built - run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback for this path)")
    L = C.CDLL(LIB_PATH)
    L.gtb200_last_error.restype = C.c_char_p
    L.gtb200_device_count.restype = C.c_int
    L.gtb200_open.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.POINTER(SeqDesc), C.c_uint]
    L.gtb200_close.argtypes = [C.c_void_p]
    L.gtb200_close.restype = None
    L.gtb200_stream.argtypes = [C.c_void_p]
    L.gtb200_stream.restype = C.c_void_p
    L.gtb200_count.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(CountInfo)]
    L.gtb200_sort_part_device.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(PartInfo)]
    L.gtb200_part_device.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
    L.gtb200_fetch_part.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_uint64]
    L.gtb200_sort_part.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_int, C.c_void_p,
                                   C.c_void_p, C.c_uint64, C.POINTER(PartInfo)]
    L.gtb200_tail.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_int]
    L.gtb200_recommendedprefixlength.argtypes = [C.c_uint, C.c_uint64]
    L.gtb200_recommendedprefixlength.restype = C.c_uint
    L.gtb200_suftabparts.argtypes = [C.c_uint, C.c_void_p, C.c_uint64, C.c_uint, C.c_void_p, C.c_void_p, C.c_void_p]
    L.gtb200_profile_enable.argtypes = [C.c_void_p, C.c_int]
    L.gtb200_profile_read.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.POINTER(C.c_double),
                                      C.POINTER(C.c_uint64), C.c_int]
    L.gtb200_profile_reset.argtypes = [C.c_void_p]
    L.gtb200_profile_reset.restype = None
    L.gtb200_count_shard.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, C.POINTER(C.c_void_p),
                                     C.POINTER(C.c_uint64)]
    L.gtb200_count_finish.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(CountInfo)]
    L.gtb200_leftborder_partialsums_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p,
                                                        C.c_char_p, C.c_size_t]
    L.gtb200_part_diag.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.c_int]
    L.gtb200_split_range.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p]
    L.gtb200_open_fasta.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_void_p, C.c_uint64, C.c_int, C.c_uint,
                                    C.POINTER(FastaInfo)]
    L.gtb200_multi_open.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.c_int, C.POINTER(SeqDesc), C.c_uint]
    L.gtb200_multi_close.argtypes = [C.c_void_p]
    L.gtb200_multi_close.restype = None
    L.gtb200_multi_count.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(CountInfo)]
    L.gtb200_multi_sort_part.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_int, C.c_void_p,
                                         C.c_void_p, C.c_uint64, C.POINTER(PartInfo)]
    L.gtb200_multi_tail.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_int]
    L.gtb200_multi_devices.argtypes = [C.c_void_p]
    L.gtb200_multi_ctx.argtypes = [C.c_void_p, C.c_int]
    L.gtb200_multi_ctx.restype = C.c_void_p
    L.gtb200_multi_peer_bytes.argtypes = [C.c_void_p]
    L.gtb200_multi_peer_bytes.restype = C.c_uint64
    L.gtb200_host_register.argtypes = [C.c_void_p, C.c_size_t]
    L.gtb200_host_unregister.argtypes = [C.c_void_p]
    L.gtb200_host_unregister.restype = None
    _lib = L
    return L


def _check(rc: int) -> None:
    if rc != 0:
        raise Gtb200Error(load_library().gtb200_last_error().decode())


def recommended_prefixlength(sigma: int, totallength: int) -> int:
    return int(load_library().gtb200_recommendedprefixlength(sigma, totallength))


def suftabparts(numofparts: int, leftborder: np.ndarray, prefixlength: int):
    """[(mincode, maxcode, width)] - gt_suftabparts_new for the whole code range."""
    L = load_library()
    numcodes = leftborder.shape[0] - 1
    mn = np.zeros(numofparts, dtype=np.uint64)
    mx = np.zeros(numofparts, dtype=np.uint64)
    wd = np.zeros(numofparts, dtype=np.uint64)
    k = L.gtb200_suftabparts(numofparts, leftborder.ctypes.data, numcodes, prefixlength,
                             mn.ctypes.data, mx.ctypes.data, wd.ctypes.data)
    return [(int(mn[i]), int(mx[i]), int(wd[i])) for i in range(k)]


def split_range(leftborder: np.ndarray, mincode: int, maxcode: int, nshards: int):
    """[(lo, hi)] - the per-GPU cut of one part (gtb200_split_range)."""
    lo = np.zeros(nshards, dtype=np.uint64)
    hi = np.zeros(nshards, dtype=np.uint64)
    k = load_library().gtb200_split_range(leftborder.ctypes.data, mincode, maxcode, nshards, lo.ctypes.data, hi.ctypes.data)
    return [(int(lo[i]), int(hi[i])) for i in range(k)]


class Sfxiterator:
    """Iterates over the suffix array part by part, like gt_Sfxiterator_next.

    sequence: either ``enc`` (uint8 codes incl. 254/255, any alphabet) or, for DNA,
    ``twobit`` + ``ranges`` exactly as GtEncseq exports them; ``device_ptrs=True`` says the arrays are
    integers holding device addresses already resident on ``device`` (multi-GPU broadcast).
    """

    def __init__(self, sigma: int, totallength: int, prefixlength: int | None = None, numofparts: int = 1,
                 enc: np.ndarray | int | None = None, twobit: np.ndarray | int | None = None,
                 ranges: np.ndarray | int | None = None, numofwords: int = 0, numofranges: int = 0,
                 device: int = 0, device_ptrs: bool = False, with_lcp: bool = True, suftabuint: bool = False,
                 code_range: tuple[int, int] | None = None, readmode: int = 0, devices: list[int] | None = None,
                 fasta: bytes | None = None):
        L = load_library()
        self._L = L
        self._ctx = C.c_void_p()
        self._multi = C.c_void_p()          # devices given: gtb200_multi_* (every part is sorted by all listed GPUs)
        self.devices = list(devices) if devices else None
        self.sigma = sigma
        self.totallength = totallength
        self.prefixlength = prefixlength or recommended_prefixlength(sigma, totallength)
        self.suftabuint = suftabuint
        self.with_lcp = with_lcp
        self.fastainfo = None
        if fasta is not None:
            # FASTA text straight to the device (gtb200_open_fasta): description lines, white space, symbol map and
            # separators are handled there; totallength and the automatic prefix length come back
            buf = np.frombuffer(fasta, dtype=np.uint8)
            fi = FastaInfo()
            _check(L.gtb200_open_fasta(C.byref(self._ctx), device, buf.ctypes.data, buf.shape[0], 1 if sigma == 20 else 0,
                                       prefixlength or 0, C.byref(fi)))
            self.fastainfo = fi
            self.totallength = totallength = int(fi.totallength)
            self.prefixlength = int(fi.prefixlength)
        desc = SeqDesc()
        desc.numofchars = sigma
        desc.totallength = totallength
        desc.memory = 1 if device_ptrs else 0
        desc.readmode = readmode
        self._keep = []
        if fasta is not None:
            pass                                    # already open
        elif twobit is not None:
            if device_ptrs:
                desc.twobitencoding, desc.numofwords = int(twobit), numofwords
                desc.specialranges, desc.numofspecialranges = int(ranges or 0), numofranges
            else:
                tw = np.ascontiguousarray(twobit, dtype=np.uint64)
                rg = np.ascontiguousarray(ranges if ranges is not None else np.zeros((0, 2)), dtype=np.uint64)
                self._keep += [tw, rg]
                desc.twobitencoding, desc.numofwords = tw.ctypes.data, tw.shape[0]
                desc.specialranges, desc.numofspecialranges = rg.ctypes.data, rg.shape[0]
        else:
            if device_ptrs:
                desc.encoded = int(enc)
            else:
                e = np.ascontiguousarray(enc, dtype=np.uint8)
                self._keep.append(e)
                desc.encoded = e.ctypes.data
        if fasta is not None:
            pass
        elif self.devices:
            devs = (C.c_int * len(self.devices))(*self.devices)
            _check(L.gtb200_multi_open(C.byref(self._multi), devs, len(self.devices), C.byref(desc), self.prefixlength))
            self._ctx = C.c_void_p(L.gtb200_multi_ctx(self._multi, 0))
        else:
            _check(L.gtb200_open(C.byref(self._ctx), device, C.byref(desc), self.prefixlength))
        # "counting prefix distribution"
        info = CountInfo()
        numcodes = sigma ** self.prefixlength
        self.leftborder = np.zeros(numcodes + 1, dtype=np.uint32)
        self.countspecialcodes = np.zeros(sigma ** (self.prefixlength - 1), dtype=np.uint32)
        ndist = sum(sigma ** i for i in range(1, self.prefixlength - 1))
        self.distpfxidx = np.zeros(ndist, dtype=np.uint32)
        if self.devices:
            _check(L.gtb200_multi_count(self._multi, self.leftborder.ctypes.data, self.countspecialcodes.ctypes.data,
                                        self.distpfxidx.ctypes.data if ndist else None, C.byref(info)))
        else:
            _check(L.gtb200_count(self._ctx, self.leftborder.ctypes.data, self.countspecialcodes.ctypes.data,
                                  self.distpfxidx.ctypes.data if ndist else None, C.byref(info)))
        self.countinfo = info
        self.numofsuffixestosort = int(info.numofsuffixestosort)
        self.specialcharacters = int(info.specialcharacters)
        self.parts = suftabparts(numofparts, self.leftborder, self.prefixlength)
        if code_range is not None:  # one shard of a multi-GPU run: a single, externally chosen part
            lo, hi = code_range
            self.parts = [(lo, hi, int(self.leftborder[hi + 1]) - int(self.leftborder[lo]))]
        self._sufbuf = self._lcpbuf = None
        self._registered = []
        if self.devices and len(self.devices) > 1 and self.parts:
            # like the drop-in's constructor: page-lock the part buffers once (seconds for tens of GB, not per part)
            largest = max(w for _, _, w in self.parts)
            self._sufbuf = np.empty(max(largest, 1), dtype=np.uint32 if self.suftabuint else np.uint64)
            self._lcpbuf = np.empty(max(largest, 1), dtype=np.uint8)
            for b in (self._sufbuf, self._lcpbuf):
                if L.gtb200_host_register(b.ctypes.data, b.nbytes) == 0:
                    self._registered.append(b)
        self._part = 0
        self._tail_done = False
        self._longest = None
        self.lcptabsum = 0.0
        self.maxbranchdepth = 0
        self.numoflargelcpvalues = 0
        self.fallback_suffixes = 0
        self.fallback_rounds = 0
        self.partinfos: list[PartInfo] = []

    # -- raw access for the benchmark ---------------------------------------------------------
    @property
    def ctx(self):
        return self._ctx

    def stream_handle(self) -> int:
        return int(self._L.gtb200_stream(self._ctx))

    def sort_part_device(self, mincode: int, maxcode: int) -> PartInfo:
        info = PartInfo()
        _check(self._L.gtb200_sort_part_device(self._ctx, mincode, maxcode, C.byref(info)))
        return info

    def part_device_pointers(self):
        a, b, c = C.c_void_p(), C.c_void_p(), C.c_void_p()
        _check(self._L.gtb200_part_device(self._ctx, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def part_diag(self) -> dict:
        """What the library did for the last sorted part (gtb200_part_diag)."""
        v = (C.c_uint64 * 8)()
        self._L.gtb200_part_diag(self._ctx, v, 8)
        return {"tile": int(v[0]), "levels": int(v[1]), "keyed": int(v[2]), "medium_suffixes": int(v[3]),
                "padded_suffixes": int(v[4])}

    def _account(self, info: PartInfo) -> None:
        self.partinfos.append(info)
        self.lcptabsum += info.lcptabsum
        self.maxbranchdepth = max(self.maxbranchdepth, int(info.maxbranchdepth))
        self.numoflargelcpvalues += int(info.numoflargelcpvalues)
        self.fallback_suffixes += int(info.fallback_suffixes)
        self.fallback_rounds += int(info.fallback_rounds)
        if info.longest != UINT64_MAX:
            self._longest = int(info.longest)

    # -- the reference's iterator protocol ----------------------------------------------------
    def next(self):
        """Returns (suftab, lcptab, llv, specialsuffixes) or None when exhausted.
        suftab: uint64 (uint32 with suftabuint) array of the part; lcptab: uint8 array (255 =
        overflow) or None for tail pages; llv: structured (position, value) array."""
        if self._part < len(self.parts):
            mincode, maxcode, width = self.parts[self._part]
            self._part += 1
            if self.devices and len(self.devices) > 1:
                # several GPUs copy straight into the part buffers: one page-locked pair for all parts, set up by the
                # constructor (the returned arrays are views, valid until the next call - gt_Sfxiterator_next's contract)
                suf, lcp = self._sufbuf[:width], self._lcpbuf[:width]
            else:
                suf = np.empty(width, dtype=np.uint32 if self.suftabuint else np.uint64)
                lcp = np.empty(width, dtype=np.uint8)
            info = PartInfo()
            if self.devices:
                llvbuf = np.empty((max(width, 1), 2), dtype=np.uint64)      # untouched pages cost nothing
                _check(self._L.gtb200_multi_sort_part(self._multi, mincode, maxcode, suf.ctypes.data,
                                                      1 if self.suftabuint else 0, lcp.ctypes.data, llvbuf.ctypes.data,
                                                      llvbuf.shape[0], C.byref(info)))
                assert info.width == width, (info.width, width)
                llv = llvbuf[:int(info.numoflargelcpvalues)].copy()
                del llvbuf
            else:
                _check(self._L.gtb200_sort_part_device(self._ctx, mincode, maxcode, C.byref(info)))
                assert info.width == width, (info.width, width)
                llv = np.zeros((int(info.numoflargelcpvalues), 2), dtype=np.uint64)
                _check(self._L.gtb200_fetch_part(self._ctx, suf.ctypes.data, 1 if self.suftabuint else 0,
                                                 lcp.ctypes.data, llv.ctypes.data if llv.size else None, llv.shape[0]))
            self._account(info)
            return suf, lcp, llv, False
        if not self._tail_done:
            self._tail_done = True
            count = self.specialcharacters + 1
            suf = np.empty(count, dtype=np.uint32 if self.suftabuint else np.uint64)
            _check(self._L.gtb200_tail(self._ctx, 0, count, suf.ctypes.data, 1 if self.suftabuint else 0))
            if self._longest is None:   # position 0 is a special (or the text is empty)
                hit = np.flatnonzero(suf == 0)
                if hit.size:
                    self._longest = self.numofsuffixestosort + int(hit[0])
            return suf, None, None, True
        return None

    def longest(self) -> int:
        return -1 if self._longest is None else self._longest

    def bcktab_bytes(self) -> bytes:
        """Byte image of the .bck file: the three tables, each zero-padded to 8 bytes
        (src/match/bcktab.c:519-570, src/core/mapspec.c:350-365)."""
        def pad8(b):
            return b + b"\0" * (-len(b) % 8)
        out = pad8(self.leftborder.tobytes()) + pad8(self.countspecialcodes.tobytes())
        if self.prefixlength > 2:
            out += pad8(self.distpfxidx.tobytes())
        return out

    def bcktab2file(self, fp) -> None:
        fp.write(self.bcktab_bytes())

    def profile(self, enable: bool | None = None, reset: bool = False):
        if enable is not None:
            _check(self._L.gtb200_profile_enable(self._ctx, 1 if enable else 0))
        if reset:
            self._L.gtb200_profile_reset(self._ctx)
            return None
        names = (C.c_char_p * 32)()
        ms = (C.c_double * 32)()
        ln = (C.c_uint64 * 32)()
        k = self._L.gtb200_profile_read(self._ctx, names, ms, ln, 32)
        if k < 0:
            _check(-1)
        return {names[i].decode(): (ms[i], int(ln[i])) for i in range(k)}

    def delete(self) -> None:
        for b in getattr(self, "_registered", []):
            self._L.gtb200_host_unregister(b.ctypes.data)
        self._registered = []
        if self._multi:
            self._L.gtb200_multi_close(self._multi)
            self._multi = C.c_void_p()
            self._ctx = C.c_void_p()
        if self._ctx:
            self._L.gtb200_close(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.delete()
        except Exception:
            pass


def suffixerator(enc: np.ndarray | None, alphabet: str = "dna", indexname: str | None = None, parts: int = 1,
                 prefixlength: int | None = None, device: int = 0, use_twobit: bool | None = None,
                 readmode: int = 0, devices: list[int] | None = None, sigma: int | None = None,
                 fasta: bytes | None = None) -> dict:
    """`gt suffixerator -suf -lcp -bck` on an encoded sequence: runs the iterator to exhaustion and
    returns the byte images of .suf/.lcp/.llv/.bck plus the .prj statistics; writes the files when
    ``indexname`` is given."""
    if sigma is None:               # any other alphabet (a -smap symbol map): give its size, the sequence goes in as bytes
        sigma = _encseq.alphabet_size(alphabet)
    n = int(enc.shape[0]) if fasta is None else 0
    if use_twobit is None:
        use_twobit = alphabet == "dna" and sigma == 4
    if fasta is not None:            # FASTA text in, everything from the parse on happens on the device
        it = Sfxiterator(sigma, 0, prefixlength, parts, device=device, fasta=fasta)
        n = it.totallength
    elif use_twobit:
        it = Sfxiterator(sigma, n, prefixlength, parts, twobit=_encseq.twobit_pack(enc),
                         ranges=_encseq.special_ranges(enc), device=device, readmode=readmode, devices=devices)
    else:
        it = Sfxiterator(sigma, n, prefixlength, parts, enc=enc, device=device, readmode=readmode, devices=devices)
    suf_chunks, lcp_chunks, llv_chunks = [], [], []
    try:
        while True:
            r = it.next()
            if r is None:
                break
            suf, lcp, llv, special = r
            suf_chunks.append(suf.tobytes())
            if special:
                lcp_chunks.append(b"\0" * suf.shape[0])   # outmany0lcpvalues (sfx-lcpvalues.c:435-451)
            else:
                lcp_chunks.append(lcp.tobytes())
                llv_chunks.append(llv.tobytes())
        out = {
            "suf": b"".join(suf_chunks), "lcp": b"".join(lcp_chunks), "llv": b"".join(llv_chunks),
            "bck": it.bcktab_bytes(), "prefixlength": it.prefixlength, "longest": it.longest(),
            "maxbranchdepth": it.maxbranchdepth, "largelcpvalues": it.numoflargelcpvalues,
            "averagelcp": "%.2f" % (it.lcptabsum / (n + 1)), "numparts": len(it.parts),
            "fallback_suffixes": it.fallback_suffixes, "fallback_rounds": it.fallback_rounds,
            "specialcharacters": it.specialcharacters,
        }
    finally:
        it.delete()
    if indexname:
        for ext in ("suf", "lcp", "llv", "bck"):
            with open(f"{indexname}.{ext}", "wb") as fh:
                fh.write(out[ext])
    return out

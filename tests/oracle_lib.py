"""ctypes access to the CPU checker under oracle/ (TEST INFRASTRUCTURE: never imported by the product).

``oracle``  = our C restatement (oracle/libxsoracle.so, prefix xso_)
``ref``     = the reference's own sources behind oracle/ref_shim.cpp (oracle/_ref/libxsref.so, prefix xsr_),
              None when it has not been built.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_PU64 = C.POINTER(C.c_uint64)


def build_oracle() -> None:
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "all"], stdout=subprocess.DEVNULL)


class Checker:
    def __init__(self, path: str, prefix: str):
        self.lib = C.CDLL(path)
        self.prefix = prefix
        L, p = self.lib, prefix
        vp, u64, u32, i32, sz = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int, C.c_size_t
        self._f = {}
        for name, res, args in [
            ("find_next", C.c_int64, [C.c_char_p, sz, vp, sz, sz]),
            ("find_next_newline", C.c_int64, [vp, sz, sz]),
            ("count_matches", u64, [C.c_char_p, sz, vp, sz]),
            ("count_matching_lines", u64, [C.c_char_p, sz, vp, sz]),
            ("count", u64, [vp, u64, C.c_char_p, u32, _PU64, u64, i32, i32]),
            ("match_byte_offsets", u64, [vp, u64, C.c_char_p, u32, _PU64, u64, _PU64, u64, i32]),
            ("line_byte_offsets", u64, [vp, u64, C.c_char_p, u32, _PU64, u64, _PU64, u64, i32]),
            ("line_indices", u64, [vp, u64, C.c_char_p, u32, _PU64, u64, _PU64, u64, i32]),
            ("lines", u64, [vp, u64, C.c_char_p, u32, _PU64, u64, _PU64, _PU64, u64, i32]),
        ]:
            fn = getattr(L, p + name)
            fn.restype, fn.argtypes = res, args
            self._f[name] = fn
        fn = getattr(L, p + "to_lower")
        fn.restype, fn.argtypes = None, [vp, u64]
        self._f["to_lower"] = fn
        if prefix == "xso_":
            fn = L.xso_chunk_bounds
            fn.restype, fn.argtypes = u64, [vp, u64, u64, _PU64, u64]
            self._f["chunk_bounds"] = fn

    @staticmethod
    def _buf(buf):
        a = np.frombuffer(buf, dtype=np.uint8) if isinstance(buf, (bytes, bytearray)) else buf
        return a, (a.ctypes.data if a.size else 0)

    def to_lower(self, buf) -> np.ndarray:
        """Copy of buf run through the reference's ignore_case processor (simd::toLower)."""
        a, _ = self._buf(buf)
        out = np.ascontiguousarray(a).copy()
        if out.size:
            self._f["to_lower"](out.ctypes.data, out.size)
        return out

    def find_next(self, pat: bytes, buf, shift: int) -> int:
        a, addr = self._buf(buf)
        if not a.size:
            a = np.zeros(1, dtype=np.uint8); addr = a.ctypes.data
        return self._f["find_next"](pat, len(pat), addr, len(buf), shift)

    def find_next_newline(self, buf, shift: int) -> int:
        a, addr = self._buf(buf)
        return self._f["find_next_newline"](addr, a.size, shift)

    def count_matches(self, pat, buf) -> int:
        a, addr = self._buf(buf)
        return self._f["count_matches"](pat, len(pat), addr, a.size)

    def count_matching_lines(self, pat, buf) -> int:
        a, addr = self._buf(buf)
        return self._f["count_matching_lines"](pat, len(pat), addr, a.size)

    def chunk_bounds(self, buf, min_chunk: int) -> np.ndarray:
        a, addr = self._buf(buf)
        nb = self._f["chunk_bounds"](addr, a.size, min_chunk, None, 0)
        out = np.empty(nb, dtype=np.uint64)
        if nb:
            self._f["chunk_bounds"](addr, a.size, min_chunk, out.ctypes.data_as(_PU64), nb)
        return out

    def count(self, buf, pat, bounds, lines: bool, nthreads: int = 1) -> int:
        a, addr = self._buf(buf)
        b = np.ascontiguousarray(bounds, dtype=np.uint64)
        return self._f["count"](addr, a.size, pat, len(pat), b.ctypes.data_as(_PU64), b.size, int(lines), nthreads)

    def _vec(self, name, buf, pat, bounds, nthreads):
        a, addr = self._buf(buf)
        b = np.ascontiguousarray(bounds, dtype=np.uint64)
        bp = b.ctypes.data_as(_PU64)
        n = self._f[name](addr, a.size, pat, len(pat), bp, b.size, None, 0, nthreads)
        out = np.empty(n, dtype=np.uint64)
        if n:
            self._f[name](addr, a.size, pat, len(pat), bp, b.size, out.ctypes.data_as(_PU64), n, nthreads)
        return out

    def match_byte_offsets(self, buf, pat, bounds, nthreads=1):
        return self._vec("match_byte_offsets", buf, pat, bounds, nthreads)

    def line_byte_offsets(self, buf, pat, bounds, nthreads=1):
        return self._vec("line_byte_offsets", buf, pat, bounds, nthreads)

    def line_indices(self, buf, pat, bounds, nthreads=1):
        return self._vec("line_indices", buf, pat, bounds, nthreads)

    def lines(self, buf, pat, bounds, nthreads=1):
        a, addr = self._buf(buf)
        b = np.ascontiguousarray(bounds, dtype=np.uint64)
        bp = b.ctypes.data_as(_PU64)
        n = self._f["lines"](addr, a.size, pat, len(pat), bp, b.size, None, None, 0, nthreads)
        st, ln = np.empty(n, dtype=np.uint64), np.empty(n, dtype=np.uint64)
        if n:
            self._f["lines"](addr, a.size, pat, len(pat), bp, b.size, st.ctypes.data_as(_PU64),
                             ln.ctypes.data_as(_PU64), n, nthreads)
        return st, ln


def load(build: bool = True):
    so = os.path.join(ORACLE_DIR, "libxsoracle.so")
    if build or not os.path.exists(so):
        build_oracle()
    oracle = Checker(so, "xso_")
    ref_so = os.path.join(ORACLE_DIR, "_ref", "libxsref.so")
    ref = Checker(ref_so, "xsr_") if os.path.exists(ref_so) else None
    return oracle, ref


class RegexRef:
    """The reference's regex wrappers over its vendored re2 (oracle/_ref/libxsref_re.so, oracle/ref_regex_shim.cpp):
    the checker of the regex prefilter.  line_indices / lines are derived from byte_offsets_line the way ref_shim.cpp
    derives them for the literal wrappers (the reference snapshot has no regex twin of either)."""

    BAD = (1 << 64) - 1  # re2 rejected the pattern

    def __init__(self, path: str):
        L = self.lib = C.CDLL(path)
        vp, u64, i32 = C.c_void_p, C.c_uint64, C.c_int
        L.xsr_re_count.restype, L.xsr_re_count.argtypes = u64, [vp, u64, C.c_char_p, i32, _PU64, u64, i32]
        for name in ("xsr_re_match_byte_offsets", "xsr_re_line_byte_offsets"):
            fn = getattr(L, name)
            fn.restype, fn.argtypes = u64, [vp, u64, C.c_char_p, i32, _PU64, u64, _PU64, u64]

    def count(self, buf, pattern: bytes, bounds, lines: bool, ignore_case: bool = False) -> int:
        a, addr = Checker._buf(buf)
        b = np.ascontiguousarray(bounds, dtype=np.uint64)
        n = self.lib.xsr_re_count(addr, a.size, pattern, int(ignore_case), b.ctypes.data_as(_PU64), b.size, int(lines))
        if n == self.BAD:
            raise ValueError("re2 rejected %r" % pattern)
        return n

    def _vec(self, name, buf, pattern, bounds, ignore_case):
        a, addr = Checker._buf(buf)
        b = np.ascontiguousarray(bounds, dtype=np.uint64)
        bp = b.ctypes.data_as(_PU64)
        fn = getattr(self.lib, name)
        n = fn(addr, a.size, pattern, int(ignore_case), bp, b.size, None, 0)
        if n == self.BAD:
            raise ValueError("re2 rejected %r" % pattern)
        out = np.empty(n, dtype=np.uint64)
        if n:
            fn(addr, a.size, pattern, int(ignore_case), bp, b.size, out.ctypes.data_as(_PU64), n)
        return out

    def match_byte_offsets(self, buf, pattern, bounds, ignore_case=False):
        return self._vec("xsr_re_match_byte_offsets", buf, pattern, bounds, ignore_case)

    def line_byte_offsets(self, buf, pattern, bounds, ignore_case=False):
        return self._vec("xsr_re_line_byte_offsets", buf, pattern, bounds, ignore_case)

    def expected_output(self, buf, pattern, bounds, tag, ignore_case=False) -> str:
        """What xsgrep --tag <tag> prints for this search."""
        a, _ = Checker._buf(buf)
        if tag == "count":
            return "%d\n" % self.count(a, pattern, bounds, False, ignore_case)
        if tag == "count_lines":
            return "%d\n" % self.count(a, pattern, bounds, True, ignore_case)
        if tag == "match_byte_offsets":
            return "".join("%d\n" % int(x) for x in self.match_byte_offsets(a, pattern, bounds, ignore_case))
        lo = self.line_byte_offsets(a, pattern, bounds, ignore_case)
        if tag == "line_byte_offsets":
            return "".join("%d\n" % int(x) for x in lo)
        nl = np.flatnonzero(a == 10)
        if tag == "line_indices":  # 0-based number of '\n' before the line
            return "".join("%d\n" % int(i) for i in np.searchsorted(nl, lo, side="left"))
        assert tag == "lines"
        raw, out = a.tobytes(), []
        for s in lo:
            k = int(np.searchsorted(nl, s, side="left"))
            if k < nl.size:  # like the literal xs::lines: an unterminated last line is dropped
                out.append(raw[int(s):int(nl[k])].decode("latin-1") + "\n")
        return "".join(out)


def load_regex_ref():
    so = os.path.join(ORACLE_DIR, "_ref", "libxsref_re.so")
    return RegexRef(so) if os.path.exists(so) else None

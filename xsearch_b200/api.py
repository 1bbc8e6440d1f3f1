"""Host-side mirror of the reference interface for the scan-and-match path, on top of the C ABI.

Reference shapes kept (file:line in /root/reference):
  * ``extern_search<Tag>(pattern, file, meta_file, num_threads, max_num_readers)`` -> Searcher whose
    work has already started (README.md:72-83, checkit.cpp:4)
  * ``Searcher::join()`` idempotent, destructor joins (include/xsearch/Searcher.h:41,52-59)
  * ``getResult()``: ``size()``, ``copyResultSafe()``, blocking iteration; count tags iterate running
    totals whose last value is the total (include/xsearch/ResultTypes.h:48-60,
    test/src/xsearchTest.cpp:346,450,735-739)
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import Iterator, List, Optional, Sequence, Tuple

import numpy as np

from . import (SCAN_IGNORE_CASE, SCAN_PUBLISH, COUNT, COUNT_LINES, DEFAULT_MIN_CHUNK, LINE_BYTE_OFFSETS, LINE_INDICES, LINES,
               MATCH_BYTE_OFFSETS, SYNTH_BLOCK, TAGS, ScanOpts, SynthParams, Timings, XsgError, _check, lib)

_PU64 = C.POINTER(C.c_uint64)
_VECTOR_KINDS = (MATCH_BYTE_OFFSETS, LINE_BYTE_OFFSETS, LINE_INDICES, LINES)


def _u64ptr(a: np.ndarray):
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_PU64)


def synth_host(n_bytes: int, params: SynthParams, first_block: int = 0) -> np.ndarray:
    """Synthetic corpus on the host (same bytes as Corpus.fill_synth on the device)."""
    assert n_bytes % SYNTH_BLOCK == 0
    out = np.empty(n_bytes, dtype=np.uint8)
    lib.xsg_synth_host(out.ctypes.data, first_block, n_bytes // SYNTH_BLOCK, C.byref(params))
    return out


class _PinnedBlock:
    """Owner of one xsg_host_alloc block (freed when the last array viewing it is gone)."""

    def __init__(self, nbytes: int):
        p = C.c_void_p()
        _check(lib.xsg_host_alloc(max(int(nbytes), 1), C.byref(p)))
        self.addr, self.nbytes = p.value, int(nbytes)

    def __del__(self):
        try:
            if self.addr:
                lib.xsg_host_free(C.c_void_p(self.addr))
                self.addr = None
        except Exception:
            pass


def pinned_empty(n: int, dtype=np.uint64) -> np.ndarray:
    """An uninitialised array in pinned host memory (xsg_host_alloc): the destination to hand to
    ``Result.vector(kind, out=...)`` when the result vector is large -- the device->host copy is then a
    single DMA at PCIe speed instead of a staged copy into pageable memory."""
    dt = np.dtype(dtype)
    blk = _PinnedBlock(int(n) * dt.itemsize)
    buf = (C.c_char * max(blk.nbytes, 1)).from_address(blk.addr)
    buf._xsg_block = blk  # the array's base keeps buf alive, buf keeps the pinned block alive
    return np.frombuffer(buf, dtype=dt, count=int(n))


def read_meta_bounds(path: str) -> np.ndarray:
    """Chunk table of an x-search v1 meta file as exclusive chunk ends (format: include/xsearch_b200/metafile.hpp;
    u32 compression type, then per chunk 5 x u64 + n_map x (u64, u64)).  Compressed inputs are rejected like in the
    C++ layer."""
    import struct
    raw = open(path, "rb").read()
    if len(raw) < 4:
        raise ValueError("meta file '%s' is truncated" % path)
    (ctype,) = struct.unpack_from("<I", raw, 0)
    if ctype != 1:
        raise ValueError("compressed inputs (meta file compression type %d) are outside the GPU path" % ctype)
    pos, ends, expect = 4, [], 0
    while pos < len(raw):
        if pos + 40 > len(raw):
            raise ValueError("meta file '%s' is truncated" % path)
        off, _, size, _, n_map = struct.unpack_from("<5Q", raw, pos)
        pos += 40 + 16 * n_map
        if pos > len(raw) or off != expect or size == 0:
            raise ValueError("meta file '%s' is truncated or has a gap in its chunk table" % path)
        expect = off + size
        ends.append(expect)
    return np.array(ends, dtype=np.uint64)


def chunk_bounds_host(buf: np.ndarray, min_chunk: int = DEFAULT_MIN_CHUNK) -> np.ndarray:
    """v1 chunking rule (>= min_chunk bytes, extended through the next newline) on a host buffer."""
    n = int(buf.size)
    nb = lib.xsg_chunk_bounds_host(buf.ctypes.data, n, min_chunk, None, 0)
    out = np.empty(nb, dtype=np.uint64)
    if nb:
        lib.xsg_chunk_bounds_host(buf.ctypes.data, n, min_chunk, _u64ptr(out), nb)
    return out


class Context:
    """One device + the library's streams and scratch (xsg_ctx)."""

    def __init__(self, device: int = 0):
        h = C.c_void_p()
        _check(lib.xsg_init(device, C.byref(h)))
        self._h = h
        self.device = device

    def close(self) -> None:
        if self._h:
            lib.xsg_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def device_info(self) -> dict:
        sm, mem, name = C.c_int(), C.c_uint64(), C.create_string_buffer(128)
        _check(lib.xsg_device_info(self._h, C.byref(sm), C.byref(mem), name, 128))
        return {"sm_count": sm.value, "total_mem": mem.value, "name": name.value.decode()}

    def set_stream(self, cuda_stream: int) -> None:
        """Run the context's device work on a caller-owned CUDA stream (0/None = its own stream)."""
        _check(lib.xsg_set_stream(self._h, C.c_void_p(cuda_stream or None)))

    def summary_tensor(self):
        """Zero-copy torch view (int64[4], on this device) of {newlines, matches, matching lines, lines} of the
        context's last scan -- the input of the multi-GPU all-gather (xsg_scan_summary_device)."""
        import torch
        if getattr(self, "_summary", None) is None:
            ptr = lib.xsg_scan_summary_device(self._h)

            class _View:
                __cuda_array_interface__ = {"shape": (4,), "typestr": "<i8", "data": (int(ptr), False), "version": 2}

            self._summary = torch.as_tensor(_View(), device=torch.device("cuda", self.device))
        return self._summary

    # -- corpora ----------------------------------------------------------------------------
    def corpus(self, n_bytes: int) -> "Corpus":
        h = C.c_void_p()
        _check(lib.xsg_corpus_alloc(self._h, n_bytes, C.byref(h)))
        return Corpus(self, h, n_bytes)

    def corpus_from_host(self, buf: np.ndarray, bounds: Optional[np.ndarray] = None,
                         min_chunk: int = DEFAULT_MIN_CHUNK) -> "Corpus":
        c = self.corpus(int(buf.size))
        c.upload(buf)
        if bounds is None:
            c.chunk(min_chunk)
        else:
            c.set_bounds(bounds)
        return c

    def wrap(self, dev_ptr: int, n_bytes: int) -> "Corpus":
        h = C.c_void_p()
        _check(lib.xsg_corpus_wrap(self._h, C.c_void_p(dev_ptr), n_bytes, C.byref(h)))
        return Corpus(self, h, n_bytes)

    # -- searches ---------------------------------------------------------------------------
    def scan(self, corpus: "Corpus", pattern: bytes, kinds: int, offset_base: int = 0, line_index_base: int = 0,
             occ_capacity: int = 0, ignore_case: bool = False, publish: bool = False) -> "Result":
        """publish: the scan's last kernel also publishes the scan's summary through the Exchange attached to this
        context (XSG_SCAN_PUBLISH) -- the push of that round without a launch of its own."""
        flags = (SCAN_IGNORE_CASE if ignore_case else 0) | (SCAN_PUBLISH if publish else 0)
        opts = ScanOpts(offset_base, line_index_base, occ_capacity, flags, 0)
        h = C.c_void_p()
        _check(lib.xsg_scan(self._h, corpus._h, pattern, len(pattern), kinds, C.byref(opts), C.byref(h)))
        return Result(h, kinds, self)

    def search_host(self, buf, n_bytes: Optional[int], pattern: bytes, kinds: int,
                    bounds: Optional[np.ndarray] = None, min_chunk: int = DEFAULT_MIN_CHUNK,
                    offset_base: int = 0, line_index_base: int = 0, ignore_case: bool = False) -> "Result":
        """buf: numpy uint8 array or raw address (int) of a host buffer (pinned for full speed)."""
        if isinstance(buf, np.ndarray):
            addr, n = buf.ctypes.data, int(buf.size)
        else:
            addr, n = int(buf), int(n_bytes)
        opts = ScanOpts(offset_base, line_index_base, 0, SCAN_IGNORE_CASE if ignore_case else 0, 0)
        h = C.c_void_p()
        bp, nb = (None, 0) if bounds is None else (_u64ptr(bounds), int(bounds.size))
        _check(lib.xsg_search_host(self._h, C.c_void_p(addr), n, bp, nb, min_chunk, pattern, len(pattern), kinds,
                                   C.byref(opts), C.byref(h)))
        return Result(h, kinds, self)


class MultiContext:
    """One host process, several devices (xsg_init_multi): a list of Contexts with peer access between them.
    ``search_host`` deals the segments of ONE host buffer round-robin to the devices and returns ONE result in file
    order (xsg_search_host_multi); ``exchanges()`` connects one Exchange per context without IPC handles."""

    def __init__(self, devices: Sequence[int]):
        n = len(devices)
        devs = (C.c_int * n)(*devices)
        hs = (C.c_void_p * n)()
        _check(lib.xsg_init_multi(devs, n, hs))
        self.contexts = []
        for d, h in zip(devices, hs):
            c = Context.__new__(Context)
            c._h, c.device = C.c_void_p(h), d
            self.contexts.append(c)

    def __len__(self) -> int:
        return len(self.contexts)

    def __getitem__(self, i: int) -> "Context":
        return self.contexts[i]

    def search_host(self, buf, n_bytes: Optional[int], pattern: bytes, kinds: int, bounds: Optional[np.ndarray] = None,
                    min_chunk: int = DEFAULT_MIN_CHUNK, offset_base: int = 0, line_index_base: int = 0,
                    ignore_case: bool = False) -> "Result":
        if isinstance(buf, np.ndarray):
            addr, n = buf.ctypes.data, int(buf.size)
        else:
            addr, n = int(buf), int(n_bytes)
        opts = ScanOpts(offset_base, line_index_base, 0, SCAN_IGNORE_CASE if ignore_case else 0, 0)
        hs = (C.c_void_p * len(self.contexts))(*[c._h for c in self.contexts])
        h = C.c_void_p()
        bp, nb = (None, 0) if bounds is None else (_u64ptr(bounds), int(bounds.size))
        _check(lib.xsg_search_host_multi(hs, len(self.contexts), C.c_void_p(addr), n, bp, nb, min_chunk, pattern,
                                         len(pattern), kinds, C.byref(opts), C.byref(h)))
        return Result(h, kinds, self.contexts[0])

    def exchanges(self) -> List["Exchange"]:
        """One connected Exchange per context (rank = position in the device list), reached through peer access."""
        return Exchange.local(self.contexts)

    def close(self) -> None:
        for c in self.contexts:
            c.close()


class Stream:
    """A streamed search (xsg_stream): push segments of whole reference chunks, read result parts as they
    finish, close() for the complete Result.  Mirrors the reference's reader -> searcher -> result loop."""

    def __init__(self, ctx: Context, pattern: bytes, kinds: int, offset_base: int = 0, line_index_base: int = 0,
                 ignore_case: bool = False):
        opts = ScanOpts(offset_base, line_index_base, 0, SCAN_IGNORE_CASE if ignore_case else 0, 0)
        h = C.c_void_p()
        _check(lib.xsg_stream_open(ctx._h, pattern, len(pattern), kinds, C.byref(opts), C.byref(h)))
        self._h, self.kinds, self.ctx = h, kinds, ctx

    def push(self, buf: np.ndarray, rel_bounds: np.ndarray) -> int:
        """buf must stay alive and unchanged until wait_copied(ticket) returned."""
        b = np.ascontiguousarray(rel_bounds, dtype=np.uint64)
        t = C.c_uint64()
        _check(lib.xsg_stream_push(self._h, C.c_void_p(buf.ctypes.data), int(buf.size), _u64ptr(b), int(b.size),
                                   C.byref(t)))
        return int(t.value)

    def wait_copied(self, ticket: int) -> None:
        _check(lib.xsg_stream_wait_copied(self._h, ticket))

    def num_parts(self) -> int:
        return int(lib.xsg_result_num_parts(lib.xsg_stream_result(self._h)))

    def part_count(self, part: int, kind: int) -> int:
        return int(lib.xsg_result_part_count(lib.xsg_stream_result(self._h), part, kind))

    def part_newlines(self, part: int) -> int:
        """'\\n' in the bytes of one finished part (valid when LINE_INDICES was asked for)."""
        return int(lib.xsg_result_part_newlines(lib.xsg_stream_result(self._h), part))

    def part_vector(self, part: int, kind: int) -> np.ndarray:
        res = lib.xsg_stream_result(self._h)
        n = int(lib.xsg_result_part_count(res, part, kind))
        out = np.empty(n, dtype=np.uint64)
        if n:
            _check(lib.xsg_result_part_copy(res, part, kind, _u64ptr(out), n))
        return out

    def close(self) -> "Result":
        h = C.c_void_p()
        st, self._h = self._h, None
        _check(lib.xsg_stream_close(st, C.byref(h)))
        return Result(h, self.kinds, self.ctx)

    def abort(self) -> None:
        if self._h:
            lib.xsg_stream_abort(self._h)
            self._h = None

    def __del__(self):
        try:
            self.abort()
        except Exception:
            pass


class Corpus:
    """Device-resident bytes + reference-chunk bounds (xsg_corpus)."""

    def __init__(self, ctx: Context, handle, n: int):
        self.ctx, self._h, self.size = ctx, handle, n

    def free(self) -> None:
        if self._h:
            if self.ctx._h:  # see Result.free
                lib.xsg_corpus_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    @property
    def device_ptr(self) -> int:
        return lib.xsg_corpus_device_ptr(self._h) or 0

    def upload(self, buf: np.ndarray, offset: int = 0) -> None:
        _check(lib.xsg_corpus_upload(self._h, buf.ctypes.data, offset, int(buf.size)))

    def download(self, offset: int = 0, n: Optional[int] = None) -> np.ndarray:
        n = self.size - offset if n is None else n
        out = np.empty(n, dtype=np.uint8)
        _check(lib.xsg_corpus_download(self._h, out.ctypes.data, offset, n))
        return out

    def fill_synth(self, params: SynthParams, first_block: int = 0) -> None:
        _check(lib.xsg_corpus_fill_synth(self._h, first_block, C.byref(params)))

    def chunk(self, min_chunk: int = DEFAULT_MIN_CHUNK) -> None:
        _check(lib.xsg_corpus_chunk(self._h, min_chunk))

    def set_bounds(self, bounds: np.ndarray) -> None:
        b = np.ascontiguousarray(bounds, dtype=np.uint64)
        _check(lib.xsg_corpus_set_bounds(self._h, _u64ptr(b), int(b.size)))

    def bounds(self) -> np.ndarray:
        nb = lib.xsg_corpus_num_chunks(self._h)
        out = np.empty(nb, dtype=np.uint64)
        if nb:
            _check(lib.xsg_corpus_get_bounds(self._h, _u64ptr(out), nb))
        return out


class Result:
    """Device-resident result vectors of one search (xsg_result)."""

    def __init__(self, handle, kinds: int, ctx: Optional[Context] = None):
        self._h, self.kinds, self.ctx = handle, kinds, ctx

    def free(self) -> None:
        if self._h:
            # a result outliving its (already destroyed) context must not touch the dangling xsg_ctx
            if self.ctx is None or self.ctx._h:
                lib.xsg_result_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    def count(self, kind: int) -> int:
        return int(lib.xsg_result_count(self._h, kind))

    def vector(self, kind: int, out: Optional[np.ndarray] = None) -> np.ndarray:
        """The result vector on the host.  ``out``: a preallocated uint64 array (e.g. ``pinned_empty``) of at
        least ``count(kind)`` entries; the filled prefix is returned."""
        n = self.count(kind)
        if out is None:
            out = np.empty(n, dtype=np.uint64)
        elif out.dtype != np.uint64 or out.size < n or not out.flags.c_contiguous:
            raise ValueError("out must be a contiguous uint64 array of at least %d entries" % n)
        if n:
            _check(lib.xsg_result_copy(self._h, kind, _u64ptr(out), n))
        return out[:n]

    def line_spans(self) -> Tuple[np.ndarray, np.ndarray]:
        n = self.count(LINES)
        st, ln = np.empty(n, dtype=np.uint64), np.empty(n, dtype=np.uint64)
        if n:
            _check(lib.xsg_result_copy(self._h, LINES, _u64ptr(st), n))
            _check(lib.xsg_result_copy_line_lengths(self._h, _u64ptr(ln), n))
        return st, ln

    def device_ptr(self, kind: int) -> int:
        return lib.xsg_result_device_ptr(self._h, kind) or 0

    @property
    def newlines(self) -> int:
        return int(lib.xsg_result_newlines(self._h))

    @property
    def raw_occurrences(self) -> int:
        return int(lib.xsg_result_raw_occurrences(self._h))

    def rebase(self, offset_delta: int = 0, line_index_delta: int = 0) -> None:
        """Multi-GPU fix-up: add the shard's global byte / line-number base on the device."""
        _check(lib.xsg_result_rebase(self._h, offset_delta, line_index_delta))

    def rebase_from_table(self, table, rank: int) -> None:
        """Device-side fix-up after an all-gather of every rank's summary into `table` (torch int64
        [world, stride] on this device): adds sum(table[r][0] for r < rank) to every line index."""
        _check(lib.xsg_result_rebase_from_table(self._h, C.c_void_p(table.data_ptr()), int(table.stride(0)), rank))

    def timings(self) -> Timings:
        t = Timings()
        _check(lib.xsg_result_timings(self._h, C.byref(t)))
        return t


class _ResultView:
    """What Searcher.get_result() hands out: size(), copy_result_safe(), blocking iteration."""

    def __init__(self, searcher: "Searcher"):
        self._s = searcher

    def size(self) -> int:
        self._s.join()
        return self._s._size

    def copy_result_safe(self):
        self._s.join()
        return list(self._s._values)

    def __iter__(self) -> Iterator:
        self._s.join()  # blocks until the producer closed the result (ResultTypes.h:48-60)
        if self._s.kind in (COUNT, COUNT_LINES):
            yield self._s._size  # running totals; the last value yielded is the total
        else:
            yield from self._s._values

    def __len__(self) -> int:
        return self.size()


class Searcher:
    """Mirror of xs::Searcher for this path: the search runs on a worker thread from construction."""

    def __init__(self, kind: int, pattern: bytes, file_path: str, meta_file: Optional[str] = None,
                 num_threads: int = 1, max_num_readers: int = 1, device: int = 0,
                 min_chunk: int = DEFAULT_MIN_CHUNK, ignore_case: bool = False):
        if isinstance(pattern, str):
            pattern = pattern.encode()
        self.kind, self.pattern, self.file_path = kind, pattern, file_path
        self._size, self._values, self._error = 0, [], None
        self._device, self._min_chunk, self._ignore_case = device, min_chunk, ignore_case
        # num_threads / max_num_readers: accepted like the reference's; this Python mirror reads the file in one piece
        # (the C++ layer, include/xsearch_b200/xsearch.hpp, is the one with reader threads and pinned segments)
        self._meta_file = meta_file or None
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()

    def _run(self) -> None:
        try:
            data = np.fromfile(self.file_path, dtype=np.uint8)
            bounds = None
            if self._meta_file:  # the chunk table of a v1 meta file is the reference-chunk boundary list
                bounds = read_meta_bounds(self._meta_file)
                if (int(bounds[-1]) if bounds.size else 0) != data.size:
                    raise ValueError("the meta file does not describe this file")
            ctx = Context(self._device)
            try:
                res = ctx.search_host(data, None, self.pattern, self.kind, bounds=bounds, min_chunk=self._min_chunk,
                                      ignore_case=self._ignore_case)
                self._size = res.count(self.kind)
                if self.kind == LINES:
                    st, ln = res.line_spans()
                    raw = data.tobytes()
                    self._values = [raw[int(s):int(s) + int(l)].decode("utf-8", "replace") for s, l in zip(st, ln)]
                elif self.kind in _VECTOR_KINDS:
                    self._values = [int(v) for v in res.vector(self.kind)]
                res.free()
            finally:
                ctx.close()
        except Exception as e:  # surfaced by join()
            self._error = e

    def join(self) -> None:
        if self._thread.is_alive():
            self._thread.join()
        if self._error is not None:
            raise self._error

    def running(self) -> bool:
        return self._thread.is_alive()

    def get_result(self) -> _ResultView:
        return _ResultView(self)

    getResult = get_result


def extern_search(tag, pattern, file_path: str, meta_file: Optional[str] = None, num_threads: int = 1,
                  max_num_readers: int = 1, ignore_case: bool = False, **kw) -> Searcher:
    """xs::extern_search<Tag>(pattern, file, meta_file, num_threads, max_num_readers)."""
    kind = TAGS[tag] if isinstance(tag, str) else int(tag)
    return Searcher(kind, pattern, file_path, meta_file, num_threads, max_num_readers, ignore_case=ignore_case, **kw)


def shard_chunks(bounds: np.ndarray, world: int) -> List[Tuple[int, int]]:
    """Contiguous, newline-aligned shards for `world` ranks: reference chunks [c0, c1) per rank, balanced by
    bytes.  Shard boundaries are chunk boundaries, so no line, no match and no tail zone spans two GPUs
    (SURVEY.md 8(e)); ranks beyond the number of chunks get an empty range."""
    nb = int(bounds.size)
    total = int(bounds[-1]) if nb else 0
    out, c0 = [], 0
    for r in range(world):
        if r == world - 1:
            c1 = nb
        else:
            target = total * (r + 1) // world
            c1 = int(np.searchsorted(bounds, target, side="left")) + 1 if target > 0 else c0
            c1 = max(c0, min(c1, nb))
            # take the chunk whose end is nearest to the ideal cut
            if c1 - 1 > c0 and abs(int(bounds[c1 - 2]) - target) <= abs(int(bounds[c1 - 1]) - target):
                c1 -= 1
        out.append((c0, c1))
        c0 = c1
    return out


def shard_exchange(newlines: int, counts: Sequence[int], group=None) -> Tuple[int, List[int], List[int]]:
    """The one exchange step of the multi-GPU path: all-gather every shard's {newline count, result
    counts}; returns (line-index base of this rank, output bases per count, global totals)."""
    import torch
    import torch.distributed as dist

    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return 0, [0 for _ in counts], [int(c) for c in counts]
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    mine = torch.tensor([int(newlines)] + [int(c) for c in counts], dtype=torch.int64, device=dev)
    world = dist.get_world_size(group)
    allv = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allv, mine, group=group)
    table = torch.stack(allv).cpu().numpy()
    rank = dist.get_rank(group)
    excl = table[:rank].sum(axis=0) if rank else np.zeros(table.shape[1], dtype=np.int64)
    tot = table.sum(axis=0)
    return int(excl[0]), [int(x) for x in excl[1:]], [int(x) for x in tot[1:]]


class Exchange:
    """The shard exchange of a multi-GPU search over NVLink peer memory (xsg_exchange, include/xsg.h).  Per round:
    the summary of the rank's shard is published into every rank's table -- by the scan itself
    (``ctx.scan(..., publish=True)``, no extra launch) or by ``push(result)`` -- and ``rebase(result)`` launches the
    fix-up kernel that waits for all ranks on the device and adds the lower shards' newline counts to the line
    indices.  One process per GPU: only the 64-byte IPC handles travel through torch.distributed, once.  Several
    contexts in one process: ``Exchange.local(contexts)``."""

    def __init__(self, ctx: "Context", group=None, _local=None):
        self.ctx = ctx
        if _local is not None:  # one of the exchanges of Exchange.local()
            self.world, self.rank = _local
            h = C.c_void_p()
            _check(lib.xsg_exchange_create(ctx._h, self.world, self.rank, C.byref(h), None))
            self._h = h
            return
        import torch
        import torch.distributed as dist
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        h = C.c_void_p()
        mine = (C.c_ubyte * 64)()
        _check(lib.xsg_exchange_create(ctx._h, self.world, self.rank, C.byref(h), C.cast(mine, C.c_void_p)))
        self._h = h
        dev = torch.device("cuda", ctx.device)
        t = torch.tensor(list(bytes(mine)), dtype=torch.uint8, device=dev)
        allh = torch.empty(self.world * 64, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(allh, t, group=group)
        raw = bytes(allh.cpu().numpy().tobytes())
        _check(lib.xsg_exchange_connect(self._h, C.cast(C.create_string_buffer(raw, len(raw)), C.c_void_p)))
        dist.barrier(group)  # every rank has mapped every table before the first push

    @staticmethod
    def local(contexts: Sequence["Context"]) -> List["Exchange"]:
        """All ranks in this process: contexts[r] is rank r (xsg_exchange_connect_local, peer access)."""
        n = len(contexts)
        exs = [Exchange(c, _local=(n, r)) for r, c in enumerate(contexts)]
        hs = (C.c_void_p * n)(*[e._h for e in exs])
        _check(lib.xsg_exchange_connect_local(hs, n))
        return exs

    def push(self, res: "Result") -> None:
        """Publish the totals of a finished result (any number of parts) as this rank's summary of the next round."""
        _check(lib.xsg_exchange_push(self._h, res._h))

    def rebase(self, res: "Result") -> None:
        _check(lib.xsg_result_rebase_exchanged(res._h, self._h))

    def totals(self) -> np.ndarray:
        """[world, 4] = {newlines, matches, matching lines, lines} of every rank for the last round."""
        out = np.empty(self.world * 4, dtype=np.uint64)
        _check(lib.xsg_exchange_totals(self._h, _u64ptr(out)))
        return out.reshape(self.world, 4)

    def check(self) -> None:
        """Raises if a wait of this exchange timed out (a peer never published)."""
        _check(lib.xsg_exchange_check(self._h))

    def close(self) -> None:
        if self._h:
            if self.ctx._h:
                lib.xsg_exchange_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def sharded_scan_async(ctx: "Context", corpus: "Corpus", pattern: bytes, kinds: int, shard_base: int, table,
                       group=None) -> "Result":
    """Like sharded_scan, but the exchange never returns to the host: the scan's device-resident summary is
    all-gathered into `table` (int64 [world, 4] on this device, caller-owned) on the current torch stream --
    which must be the stream the context runs on (Context.set_stream) -- and the line-index base is added
    by a kernel that reads the table.  Global totals are table[:, 1:].sum(0) whenever the caller wants them."""
    import torch.distributed as dist
    res = ctx.scan(corpus, pattern, kinds, offset_base=shard_base)
    dist.all_gather_into_tensor(table.view(-1), ctx.summary_tensor(), group=group)
    res.rebase_from_table(table, dist.get_rank(group))
    return res


def sharded_scan(ctx: "Context", corpus: "Corpus", pattern: bytes, kinds: int, shard_base: int, group=None):
    """One rank's part of a multi-GPU search: scan the local shard with its global byte base, exchange
    {newlines, counts} with the other ranks, add the global line-number base on the device.
    Returns (result, output bases per requested kind, global totals per requested kind)."""
    res = ctx.scan(corpus, pattern, kinds, offset_base=shard_base)
    ks = [k for k in (COUNT, COUNT_LINES, MATCH_BYTE_OFFSETS, LINE_BYTE_OFFSETS, LINE_INDICES, LINES) if kinds & k]
    line_base, bases, totals = shard_exchange(res.newlines, [res.count(k) for k in ks], group)
    if line_base:
        res.rebase(0, line_base)
    return res, dict(zip(ks, bases)), dict(zip(ks, totals))

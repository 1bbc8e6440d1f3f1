"""xsg_count_static_scratch (include/xsg.h): xs::count with the scan's scratch memory in a fixed __device__ array of
libxsg.so -- no allocation inside the call.  Checked through the C ABI against the existing entry point (xsg_scan,
kinds = XSG_COUNT) on the same device bytes and chunk bounds, bit-exact, on calls in a row (the array is shared and
keeps the last call's contents).  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def xs():
    import xsearch_b200
    return xsearch_b200


@pytest.fixture(scope="module")
def ctx(xs):
    c = xs.Context(0)
    yield c
    c.close()


class DeviceText:
    """text + chunk bounds in device memory owned by the TEST (torch), wrapped for xsg_scan as well"""

    def __init__(self, xs, ctx, buf: np.ndarray, min_chunk: int):
        import torch
        self.torch = torch
        self.n = int(buf.size)
        padded = np.zeros((self.n + 15) // 16 * 16 + 16, np.uint8)
        padded[:self.n] = buf
        self.text = torch.from_numpy(padded).cuda()
        self.h_bounds = xs.chunk_bounds_host(buf, min_chunk)
        self.bounds = torch.from_numpy(self.h_bounds.astype(np.int64)).cuda()
        self.out = torch.full((2,), -1, dtype=torch.int64, device="cuda")
        self.corpus = ctx.wrap(self.text.data_ptr(), self.n)
        self.corpus.set_bounds(self.h_bounds)

    def static_count(self, xs, pat: bytes, flags: int = 0, bounds="own"):
        """-> (rc, count, flag word) of one xsg_count_static_scratch call on torch's current stream"""
        torch = self.torch
        self.out.fill_(-1)
        if bounds == "own":
            bp, nb = self.bounds.data_ptr(), int(self.bounds.numel())
        elif bounds is None:
            bp, nb = None, 0
        else:
            bp, nb = bounds.data_ptr(), int(bounds.numel())
        rc = xs.lib.xsg_count_static_scratch(0, C.c_void_p(torch.cuda.current_stream().cuda_stream),
                                             C.c_void_p(self.text.data_ptr()), self.n, C.c_void_p(bp), nb,
                                             pat, len(pat), flags, C.c_void_p(self.out.data_ptr()))
        if rc != 0:
            return rc, None, None
        got = self.out.cpu().numpy().astype(np.uint64)  # stream-ordered behind the call's kernels
        return rc, int(got[0]), int(got[1])

    def scan_count(self, xs, ctx, pat: bytes, ignore_case: bool = False) -> int:
        res = ctx.scan(self.corpus, pat, xs.COUNT, ignore_case=ignore_case)
        try:
            return res.count(xs.COUNT)
        finally:
            res.free()


@pytest.fixture(scope="module")
def text(xs, ctx):
    buf = xs.synth_host(8192 * 4096, xs.SynthParams.make(21, b"Sherlock", 30.0 / 4096))  # 32 MiB, ~1 hit per 4 KiB of lines
    t = DeviceText(xs, ctx, buf, 1 << 16)
    assert t.h_bounds.size > 100
    return t


def test_two_calls_in_a_row_match_xsg_scan(xs, ctx, text):
    expect = {p: text.scan_count(xs, ctx, p) for p in (b"Sherlock", b"the", b"e", b"that", b"of th")}
    assert expect[b"Sherlock"] > 100 and expect[b"the"] > 10000
    # back to back, no xsg_scan in between: every call finds the previous call's scalars and counters in the array
    for p in (b"Sherlock", b"the", b"Sherlock", b"e", b"of th", b"that", b"Sherlock"):
        rc, count, flag = text.static_count(xs, p)
        assert rc == 0 and flag == 0, (p, rc, flag)
        assert count == expect[p], p


def test_no_allocation_inside_the_call(xs, ctx, text):
    import torch
    text.static_count(xs, b"Sherlock")  # warm: module loaded, symbol address known
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    stats0 = torch.cuda.memory_stats().get("allocation.all.allocated", 0)
    for _ in range(3):
        rc = xs.lib.xsg_count_static_scratch(0, C.c_void_p(torch.cuda.current_stream().cuda_stream),
                                             C.c_void_p(text.text.data_ptr()), text.n, C.c_void_p(text.bounds.data_ptr()),
                                             int(text.bounds.numel()), b"Sherlock", 8, 0, C.c_void_p(text.out.data_ptr()))
        assert rc == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 >= free0
    assert torch.cuda.memory_stats().get("allocation.all.allocated", 0) == stats0


def test_ignore_case_and_whole_text_chunk(xs, ctx, text):
    rc, count, flag = text.static_count(xs, b"sherlock", flags=xs.SCAN_IGNORE_CASE)
    assert rc == 0 and flag == 0
    assert count == text.scan_count(xs, ctx, b"sherlock", ignore_case=True) > 100
    # NULL bounds = one chunk [0, n): the same as xsg_scan with that single bound
    whole = DeviceText.__new__(DeviceText)
    whole.__dict__.update(text.__dict__)
    whole.corpus = ctx.wrap(text.text.data_ptr(), text.n)
    whole.corpus.set_bounds(np.array([text.n], np.uint64))
    for p in (b"the", b"Sherlock"):
        rc, count, flag = text.static_count(xs, p, bounds=None)
        assert rc == 0 and flag == 0
        assert count == whole.scan_count(xs, ctx, p)
    whole.corpus.free()


def test_short_and_empty_texts(xs, ctx):
    buf = np.frombuffer(b"a needle in a haystack\nno hit here\nneedle needle\n", np.uint8)
    t = DeviceText(xs, ctx, buf, 1)
    for p in (b"needle", b"hay", b"n"):
        rc, count, flag = t.static_count(xs, p)
        assert rc == 0 and flag == 0 and count == t.scan_count(xs, ctx, p), p
    t.n = 0
    rc, count, flag = t.static_count(xs, b"needle", bounds=None)
    assert (rc, count, flag) == (0, 0, 0)


def test_flags_overlap_and_bad_bounds(xs, ctx, text):
    import torch
    buf = np.frombuffer(b"xx aaaa yy\naa\n" * 400, np.uint8)
    t = DeviceText(xs, ctx, buf, 64)
    rc, count, flag = t.static_count(xs, b"aa")
    assert rc == 0 and flag == xs.STATIC_OVERLAP         # "aaaa": the greedy selection needs the list -> xsg_scan
    rc, count, flag = t.static_count(xs, b"yy")          # bordered, but no two occurrences overlap here
    assert rc == 0 and flag == 0 and count == t.scan_count(xs, ctx, b"yy") == 400
    bad = torch.tensor([5, t.n], dtype=torch.int64, device="cuda")  # byte 4 is not a newline
    rc, count, flag = t.static_count(xs, b"xx", bounds=bad)
    assert rc == 0 and flag == xs.STATIC_BAD_BOUNDS
    rc, count, flag = t.static_count(xs, b"xx")          # and the flag does not stick
    assert rc == 0 and flag == 0 and count == 400


def test_rejects_what_it_cannot_do(xs, text):
    n_ok = text.n
    try:
        text.n = xs.STATIC_SCRATCH_MAX_TEXT + 1
        rc, _, _ = text.static_count(xs, b"Sherlock")
        assert rc == -2
        msg = xs.lib.xsg_last_error().decode()
        assert "exceeds" in msg and str(xs.STATIC_SCRATCH_MAX_TEXT) in msg
    finally:
        text.n = n_ok
    rc, _, _ = text.static_count(xs, b"abcabcab")        # self-overlapping and longer than 4 bytes
    assert rc == -2 and "occurrence list" in xs.lib.xsg_last_error().decode()
    rc, _, _ = text.static_count(xs, b"a\nb")
    assert rc == -2
    rc, count, flag = text.static_count(xs, b"Sherlock")  # still fine afterwards
    assert rc == 0 and flag == 0 and count > 100

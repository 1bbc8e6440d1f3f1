"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/xsg.h declares, its
host helpers agree with the oracle, and compute entry points fail loudly without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import oracle_lib
import xsearch_b200 as xs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    hdr = open(os.path.join(ROOT, "include", "xsg.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(xsg_[a-z0-9_]+)\s*\(", hdr)))


def test_exports_match_header():
    names = _declared()
    assert len(names) >= 30
    for n in names:
        assert hasattr(xs.lib, n), "libxsg.so does not export %s" % n
    assert sorted(xs.EXPORTS) == names
    assert xs.lib.xsg_version() == 1


def test_no_torch_types_in_signatures():
    hdr = open(os.path.join(ROOT, "include", "xsg.h")).read()
    assert "torch" not in hdr and "at::" not in hdr and "std::" not in hdr


def test_synth_host_deterministic_and_shaped():
    p = xs.SynthParams.make(5, b"Sherlock", 0.01)
    a = xs.synth_host(64 * 4096, p)
    b = xs.synth_host(64 * 4096, p)
    assert np.array_equal(a, b)
    # block-wise pure function: block 7 alone equals block 7 of the big buffer
    c = xs.synth_host(4096, p, first_block=7)
    assert np.array_equal(c, a[7 * 4096:8 * 4096])
    blocks = a.reshape(-1, 4096)
    assert (blocks[:, -1] == 10).all()           # every block ends with a newline
    text = a.tobytes()
    assert text.count(b"Sherlock") > 10
    lines = text.split(b"\n")
    assert 15 < np.mean([len(l) for l in lines]) < 60
    assert xs.synth_host(4096, xs.SynthParams.make(6)).tobytes() != c.tobytes()


def test_synth_blocks_are_independent_streams():
    """round 1 seeded block b with seed*C + b*phi and stepped by phi, so every block was its
    neighbour shifted by one word and needle hits came in bursts (variance 5541 at mean 26 per MiB).  Now every
    block walks its own stream: hit counts per MiB are Poisson-like and neighbours share no text."""
    import bench
    tags, pat, needle, rate, seed, nbytes, scaling = bench.WORKLOADS[bench.DEFAULT_WORKLOAD]
    n = 1 << 30
    buf = bench.host_corpus(n, needle, rate, seed, os.cpu_count() or 4)
    text = buf.tobytes()
    hits = np.array([text.count(needle, i << 20, (i + 1) << 20) for i in range(n >> 20)])
    assert 0.8 * (n / 65536) < hits.sum() < 1.25 * (n / 65536)      # ~1 hit per 64 KiB as configured
    assert hits.var() / hits.mean() < 2.0, (hits.mean(), hits.var())
    # adjacent blocks share no 32-byte substring that starts at a word start of the later block
    for b in range(64):
        prev, cur = text[b * 4096:(b + 1) * 4096], text[(b + 1) * 4096:(b + 2) * 4096]
        body = cur[:cur.rindex(b"\n", 0, 4095) + 1]                 # without the filler line of dots
        starts = [0] + [i + 1 for i, c in enumerate(body[:-33]) if c in b" \n"]
        assert not any(body[s:s + 32] in prev for s in starts), b
    # the library's generator and the CUDA-free one of oracle/ (bench.py --impl reference) give the same bytes
    assert np.array_equal(buf[:1 << 22], xs.synth_host(1 << 22, xs.SynthParams.make(seed, needle, rate)))


def test_chunk_bounds_host_matches_oracle():
    oracle, _ = oracle_lib.load()
    a = xs.synth_host(256 * 4096, xs.SynthParams.make(11))
    for min_chunk in (1, 100, 4096, 50000, 10 ** 7):
        assert np.array_equal(xs.chunk_bounds_host(a, min_chunk), oracle.chunk_bounds(a, min_chunk))
    assert xs.chunk_bounds_host(a[:0], 100).size == 0


def test_init_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(xs.XsgError) as e:
        xs.Context(0)
    assert "no CPU fallback" in str(e.value)


def test_static_scratch_entry_checks_arguments_before_touching_cuda():
    """xsg_count_static_scratch: the size limit and the pattern rules are host-side checks (clear errors, no GPU needed);
    with valid arguments and no GPU it fails loudly like every compute entry point."""
    import torch
    out = (C.c_uint64 * 2)()
    call = lambda n, pat: xs.lib.xsg_count_static_scratch(0, None, C.c_void_p(4096), n, None, 0, pat, len(pat), 0,
                                                          C.cast(out, C.c_void_p))
    assert call(xs.STATIC_SCRATCH_MAX_TEXT + 1, b"Sherlock") == -2
    assert "exceeds the static scratch's maximum of %d bytes" % xs.STATIC_SCRATCH_MAX_TEXT in xs.lib.xsg_last_error().decode()
    assert call(4096, b"abcabcab") == -2 and "occurrence list" in xs.lib.xsg_last_error().decode()
    assert call(4096, b"") == -2
    if not torch.cuda.is_available():
        assert call(4096, b"Sherlock") == -1  # XSG_ERR_CUDA

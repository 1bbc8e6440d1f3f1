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

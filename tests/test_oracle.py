"""Pins the CPU oracle (oracle/xs_oracle.c): against the reference's own known-answer tests, against
golden outputs produced by the reference's code, and -- when oracle/_ref/libxsref.so is present --
differentially against the reference's unmodified sources.  CPU only."""
import json
import os

import numpy as np
import pytest

import gen_cases
import oracle_lib

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def libs():
    return oracle_lib.load()


@pytest.fixture(scope="module")
def kat():
    return json.load(open(os.path.join(GOLD, "kat.json")))


def _checkers(libs):
    oracle, ref = libs
    return [("oracle", oracle)] + ([("ref", ref)] if ref is not None else [])


def test_kat_simd_primitives(libs, kat):
    # reference: test/src/string_search/simd_searchTest.cpp:63-98
    text = open(os.path.join(GOLD, "simd_dummy_text.bin"), "rb").read()
    assert len(text) == 1240
    for name, chk in _checkers(libs):
        for pat, shift, want in kat["findNext"]:
            assert chk.find_next(pat.encode("latin-1"), text, shift) == want, (name, pat, shift)
        for shift, want in kat["findNextNewLine"]:
            assert chk.find_next_newline(text, shift) == want, (name, shift)
        for pat, want in kat["findAllPerLine"]:
            assert chk.count_matching_lines(pat.encode("latin-1"), text) == want, (name, pat)
        for pat, want in kat["findAll"]:
            assert chk.count_matches(pat.encode("latin-1"), text) == want, (name, pat)


def test_kat_wrappers(libs, kat):
    # reference: test/src/string_search/search_wrappersTest.cpp:26,39,52 (+ line contents of :64-69)
    text = np.frombuffer(open(os.path.join(GOLD, "wrappers_dummy_text.bin"), "rb").read(), dtype=np.uint8)
    w = kat["wrappers"]
    pat = w["pattern"].encode()
    bounds = np.array([text.size], dtype=np.uint64)
    for name, chk in _checkers(libs):
        assert list(chk.match_byte_offsets(text, pat, bounds)) == w["byte_offsets_match"], name
        assert list(chk.line_byte_offsets(text, pat, bounds)) == w["byte_offsets_line"], name
        assert chk.count(text, pat, bounds, True) == w["count"], name
        st, ln = chk.lines(text, pat, bounds)
        got = [text[int(s):int(s) + int(l)].tobytes().decode() for s, l in zip(st, ln)]
        assert got == w["lines"], name
        assert list(chk.line_indices(text, pat, bounds)) == [0, 2, 3, 8], name


def test_known_quirks(libs):
    # the scalar tail skips past a partial match
    oracle, _ = libs
    assert oracle.find_next(b"Sherlock", b"SSherlock", 0) == -1
    buf = bytearray(b"." * 100)
    buf[64:72] = b"Sherlock"
    buf[63:64] = b"S"
    assert oracle.find_next(b"Sherlock", bytes(buf), 0) == 64  # start 64 lies in the exact AVX2 region
    buf = bytearray(b"." * 100)
    buf[65:73] = b"Sherlock"
    buf[64:65] = b"S"
    assert oracle.find_next(b"Sherlock", bytes(buf), 0) == -1  # start 65 lies in the scalar tail
    a = np.frombuffer(b"a" * 200, dtype=np.uint8)
    b = np.array([200], dtype=np.uint64)
    assert oracle.count(a, b"aa", b, False) == 100
    assert oracle.count(a, b"aaa", b, False) == 66


def test_golden_reference_vectors(libs):
    """Outputs of the reference's own code, committed by tools/make_golden.py."""
    oracle, _ = libs
    data = json.load(open(os.path.join(GOLD, "ref_vectors.json")))
    assert len(data["vectors"]) > 100
    for v in data["vectors"]:
        buf, pat, bounds = gen_cases.materialize(v["case"])
        assert oracle.count(buf, pat, bounds, False) == v["count"], v["case"]
        assert oracle.count(buf, pat, bounds, True) == v["count_lines"], v["case"]
        assert [int(x) for x in oracle.match_byte_offsets(buf, pat, bounds)] == v["match_byte_offsets"], v["case"]
        assert [int(x) for x in oracle.line_byte_offsets(buf, pat, bounds)] == v["line_byte_offsets"], v["case"]
        assert [int(x) for x in oracle.line_indices(buf, pat, bounds)] == v["line_indices"], v["case"]
        st, ln = oracle.lines(buf, pat, bounds)
        assert [[int(a), int(b)] for a, b in zip(st, ln)] == v["lines"], v["case"]


def test_differential_vs_reference(libs):
    """Restatement vs the reference's unmodified sources on random adversarial inputs."""
    oracle, ref = libs
    if ref is None:
        pytest.skip("oracle/_ref/libxsref.so not built (no reference tree here)")
    n_cases = 0
    for case in gen_cases.random_cases(1500, seed0=7):
        buf, pat, bounds = gen_cases.materialize(case)
        for threads in (1,):
            assert oracle.count(buf, pat, bounds, False, threads) == ref.count(buf, pat, bounds, False, threads), case
            assert oracle.count(buf, pat, bounds, True, threads) == ref.count(buf, pat, bounds, True, threads), case
        assert np.array_equal(oracle.match_byte_offsets(buf, pat, bounds), ref.match_byte_offsets(buf, pat, bounds)), case
        assert np.array_equal(oracle.line_byte_offsets(buf, pat, bounds), ref.line_byte_offsets(buf, pat, bounds)), case
        assert np.array_equal(oracle.line_indices(buf, pat, bounds), ref.line_indices(buf, pat, bounds)), case
        so, lo = oracle.lines(buf, pat, bounds)
        sr, lr = ref.lines(buf, pat, bounds)
        assert np.array_equal(so, sr) and np.array_equal(lo, lr), case
        n_cases += 1
    assert n_cases == 1500


def test_find_next_model_exhaustive_small(libs):
    """findNext model vs the reference for every shift of small buffers."""
    oracle, ref = libs
    if ref is None:
        pytest.skip("oracle/_ref/libxsref.so not built")
    rng = np.random.default_rng(3)
    for _ in range(400):
        n = int(rng.integers(0, 140))
        alphabet = [b"ab", b"abc", b"a", b"ab\n"][int(rng.integers(0, 4))]
        buf = bytes(alphabet[int(i)] for i in rng.integers(0, len(alphabet), size=n))
        m = int(rng.integers(1, 6))
        pat = bytes(alphabet[int(i)] for i in rng.integers(0, len(alphabet), size=m)).replace(b"\n", b"a")
        for shift in range(0, n + 2):
            assert oracle.find_next(pat, buf, shift) == ref.find_next(pat, buf, shift), (buf, pat, shift)


def test_threads_and_chunking(libs):
    oracle, _ = libs
    case = {"kind": "planted", "seed": 99, "n": 60000, "alphabet": "abcdefgh \n", "pattern": "abc", "nl_prob": 0.03,
            "min_chunk": 5000, "plants": 300, "end_nl": True}
    buf, pat, bounds = gen_cases.materialize(case)
    assert np.array_equal(oracle.chunk_bounds(buf, 5000), bounds)
    assert bounds[-1] == buf.size and all(buf[int(b) - 1] == 10 for b in bounds[:-1])
    one = oracle.match_byte_offsets(buf, pat, bounds, 1)
    four = oracle.match_byte_offsets(buf, pat, bounds, 4)
    assert np.array_equal(one, four) and one.size > 100
    assert oracle.count(buf, pat, bounds, True, 4) == oracle.line_indices(buf, pat, bounds, 3).size


def test_to_lower_matches_reference(libs):
    """simd::toLower (string_utils.cpp:11-33): only ASCII A-Z move, at every alignment / tail length."""
    oracle, ref = libs
    allb = np.arange(256, dtype=np.uint8)
    want = allb.copy()
    want[65:91] += 32
    for n in (0, 1, 31, 32, 33, 63, 64, 65, 256, 700):
        buf = np.resize(allb, n)
        assert np.array_equal(oracle.to_lower(buf), np.resize(want, n)), n
        if ref is not None:
            assert np.array_equal(ref.to_lower(buf), np.resize(want, n)), n


def test_golden_ignore_case_vectors(libs):
    """ignore_case = the reference's toLower over chunk and pattern, then its matchers (committed outputs)."""
    oracle, _ = libs
    data = json.load(open(os.path.join(GOLD, "icase_vectors.json")))
    assert len(data["vectors"]) >= 40 and sum(v["count"] for v in data["vectors"]) > 300
    for v in data["vectors"]:
        buf, pat, bounds = gen_cases.materialize(v["case"])
        lbuf, lpat = oracle.to_lower(buf), oracle.to_lower(np.frombuffer(pat, dtype=np.uint8)).tobytes()
        assert oracle.count(lbuf, lpat, bounds, False) == v["count"], v["case"]
        assert oracle.count(lbuf, lpat, bounds, True) == v["count_lines"], v["case"]
        assert [int(x) for x in oracle.match_byte_offsets(lbuf, lpat, bounds)] == v["match_byte_offsets"], v["case"]
        assert [int(x) for x in oracle.line_indices(lbuf, lpat, bounds)] == v["line_indices"], v["case"]
        st, ln = oracle.lines(lbuf, lpat, bounds)
        assert [[int(a), int(b)] for a, b in zip(st, ln)] == v["lines"], v["case"]

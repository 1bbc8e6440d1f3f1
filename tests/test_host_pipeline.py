"""The host pipeline of xs::extern_search (include/xsearch_b200/xsearch.hpp) on the CPU, against a TEST DOUBLE of the C ABI.

The GPU tests of tests/test_cpp_api.py and tests/test_regex.py drive `xsgrep` (the C++20 layer over libxsg.so) on a
B200.  Everything above the C ABI is host logic -- reader threads that pread slices of the same segment, the v1 chunk
rule applied to segments, carries across segment edges (also longer than the headroom), recycling of the pinned
buffers, the live iterator, the regex host stage, the failure path -- and a container without a GPU can check it when
the C ABI underneath answers: tests/cpp/xsg_testdouble.cpp implements the 19 entry points the header calls with the CPU
oracle and reads every pushed segment LATE (inside the next push, like the real library's scan), comparing a checksum
taken at push time, so a buffer recycled too early is an error and not luck.

The double lives under tests/, is built into tests/cpp/_build/ and is linked only into the copy of the driver built
there (xsgrep_double); the product (xsearch_b200/cpp/xsgrep, libxsg.so) never sees it.  The same check functions run
here against the double and, marked gpu, against the real library: whatever passes here and fails there is a
difference between the double and the CUDA path, not a host-pipeline bug.

Also here: the same driver under ThreadSanitizer (readers + pusher + verification threads + a live consumer), and the
failure path (an injected push failure closes the result, reports the error, leaves no thread behind)."""
import os
import subprocess

import numpy as np
import pytest

import oracle_lib
import test_cpp_api as capi
import test_regex as rx

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPP = os.path.join(ROOT, "tests", "cpp")
OUT = os.path.join(CPP, "_build")
SOURCES = [os.path.join(ROOT, "xsearch_b200", "cpp", "xsgrep.cpp"), os.path.join(CPP, "xsg_testdouble.cpp")]
HEADERS = [os.path.join(ROOT, "include", "xsg.h"), os.path.join(ROOT, "oracle", "xs_oracle.c"),
           os.path.join(ROOT, "oracle", "xs_oracle.h")] + \
          [os.path.join(ROOT, "include", "xsearch_b200", f) for f in sorted(os.listdir(os.path.join(ROOT, "include", "xsearch_b200")))]


def _build(name, flags, main_src=None):
    """xsgrep.cpp (or another program on xsearch.hpp) + the test double + the oracle's C restatement in one executable
    (no libxsg.so involved)."""
    os.makedirs(OUT, exist_ok=True)
    exe = os.path.join(OUT, name)
    sources = SOURCES if main_src is None else [main_src, SOURCES[1]]
    newest = max(os.path.getmtime(p) for p in sources + HEADERS)
    if os.path.exists(exe) and os.path.getmtime(exe) >= newest:
        return exe
    obj = os.path.join(OUT, name + ".oracle.o")
    subprocess.check_call(["gcc", "-O2", "-c", *[f for f in flags if f.startswith("-fsanitize") or f == "-g"],
                           "-o", obj, os.path.join(ROOT, "oracle", "xs_oracle.c")])
    subprocess.check_call(["g++", "-std=c++20", "-Wall", "-Wextra", "-pthread", *flags, "-I" + os.path.join(ROOT, "include"),
                           "-o", exe, *sources, obj])
    return exe


VARIANTS = {
    "xsgrep_double": (["-O2"], None),
    "xsgrep_double_tsan": (["-O1", "-g", "-fsanitize=thread"], None),
    "xsgrep_double_asan": (["-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined"], None),
    "reference_usage": (["-O2"], os.path.join(CPP, "reference_usage.cpp")),
}


@pytest.fixture(scope="module")
def built():
    """All variants compiled side by side (each is ~25 s of g++ on the header-only host layer)."""
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(len(VARIANTS)) as pool:
        jobs = {name: pool.submit(_build, name, *spec) for name, spec in VARIANTS.items()}
        out = {}
        for name, job in jobs.items():
            try:
                out[name] = job.result()
            except subprocess.CalledProcessError as e:
                out[name] = e
    return out


def _variant(built, name, what):
    exe = built[name]
    if isinstance(exe, Exception):
        if name == "xsgrep_double" or name == "reference_usage":
            raise exe
        pytest.skip(what + " runtime not available")
    return exe


@pytest.fixture(scope="module")
def double(built):
    return _variant(built, "xsgrep_double", "")


@pytest.fixture(scope="module")
def double_tsan(built):
    exe = _variant(built, "xsgrep_double_tsan", "ThreadSanitizer")
    probe = subprocess.run([exe, "--selftest-host"], capture_output=True, text=True)
    if probe.returncode != 0 and "ThreadSanitizer" not in probe.stderr:
        pytest.skip("ThreadSanitizer binaries do not run here: " + probe.stderr[-200:])
    return exe


@pytest.fixture(scope="module")
def double_asan(built):
    exe = _variant(built, "xsgrep_double_asan", "AddressSanitizer")
    probe = subprocess.run([exe, "--selftest-host"], capture_output=True, text=True)
    if probe.returncode != 0 and "Sanitizer" not in probe.stderr:
        pytest.skip("sanitizer binaries do not run here: " + probe.stderr[-200:])
    return exe


def test_double_is_not_the_product():
    """The double must stay test infrastructure: nothing outside tests/ names it, and the product driver links libxsg.so."""
    for base in ("xsearch_b200", "include", "bench.py", "__graft_entry__.py", "tools"):
        path = os.path.join(ROOT, base)
        files = [path] if os.path.isfile(path) else [os.path.join(d, f) for d, _, fs in os.walk(path) for f in fs
                                                     if f.endswith((".py", ".cpp", ".hpp", ".h", ".cu", ".cuh", ".inc", ".sh")) or f == "Makefile"]
        for f in files:
            text = open(f, errors="replace").read()
            assert "testdouble" not in text and "xsgrep_double" not in text, f
    assert "-lxsg" in open(os.path.join(ROOT, "xsearch_b200", "cpp", "Makefile")).read()


def test_all_tags(double, tmp_path):
    capi.check_all_tags(double, tmp_path)


def test_with_meta_file(double, tmp_path):
    capi.check_with_meta_file(double, tmp_path)


def test_edge_files(double, tmp_path):
    capi.check_edge_files(double, tmp_path)


def test_ignore_case(double, tmp_path):
    capi.check_ignore_case(double, tmp_path)


def test_errors(double, tmp_path):
    capi.check_errors(double, tmp_path)


def test_multi_device_lanes(double, tmp_path):
    """xs::Options::devices: the round-robin lanes, the file-order harvest and the host-side line-index bases, with
    three pretend devices (the double accepts any device list)."""
    capi.check_multi_device(double, tmp_path, [0, 1, 2])


def test_regex(double, tmp_path):
    golden = rx.json.load(open(os.path.join(rx.GOLD, "regex_vectors.json")))
    rx.check_extern_search_regex(golden, rx.corpora(), double, tmp_path)


def test_reference_call_shapes(built, tmp_path):
    """tests/cpp/reference_usage.cpp: every call shape the reference's README, examples and tests use (overloads, a
    char* meta path, join + size, copyResultSafe, live loops, running totals, an ignored searcher) compiles against
    xsearch.hpp and returns what the oracle says."""
    exe = _variant(built, "reference_usage", "")
    oracle, _ = oracle_lib.load(build=False)
    rng = np.random.default_rng(12)
    words = [b"Needle", b"needle", b"NEEDLE", b"nneedle", b"hay", b"stack of", b""]
    buf = np.frombuffer(b"\n".join(b" ".join(words[int(i)] for i in rng.integers(0, len(words), int(rng.integers(0, 9))))
                                   for _ in range(60000)) + b"\n", dtype=np.uint8).copy()
    path, meta = tmp_path / "u.txt", tmp_path / "u.meta"
    buf.tofile(path)
    mb = oracle.chunk_bounds(buf, 65536)
    capi.write_meta(meta, mb)
    one = np.array([buf.size], dtype=np.uint64)
    low = oracle.to_lower(buf)
    pat = b"needle"
    out = subprocess.run([exe, pat.decode(), str(path), str(meta)], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    got = dict(line.split(" ", 1) for line in out.stdout.splitlines())
    st, ln = oracle.lines(buf, pat, one)
    assert got["S1"] == "%d %d" % (len(st), int(ln.sum()))
    assert got["S2"] == "%d" % len(oracle.lines(buf, pat, mb)[0])
    assert got["S3"] == "%d" % oracle.count(buf, pat, mb, False)
    assert got["S4"] == "%d" % oracle.count(buf, pat, one, False)
    assert got["S4i"] == "%d" % oracle.count(low, pat, one, True)
    assert got["S5"] == "%d" % oracle.count(low, pat, mb, False)
    lbo = oracle.line_byte_offsets(buf, pat, one)
    assert got["S6"] == "%d %d 1" % (len(lbo), int(lbo.sum()))
    li = oracle.line_indices(buf, pat, mb)
    assert got["S6i"] == "%d %d" % (len(li), int(li.sum()))
    assert got["S6m"] == "%d" % len(oracle.match_byte_offsets(buf, pat, one))
    assert got["S7"] == "%d 1 1" % oracle.count(buf, pat, one, False)
    assert got["S8"] == "ok"
    assert int(got["S5"]) > int(got["S3"]) > 0  # ignore case finds more: the corpus is not trivial


def _ragged_text(rng, n_lines):
    """Lines of wildly different lengths (0 .. 200 kB) so that chunk ends, segment ends and slice ends fall everywhere."""
    words = [b"needle", b"nneedle", b"need", b"le", b"x", b"hay stack", b""]
    lines = []
    for _ in range(n_lines):
        r = rng.random()
        if r < 0.02:
            body = b"y" * int(rng.integers(50000, 200000)) + (b"needle" if rng.random() < 0.5 else b"")
        else:
            body = b" ".join(words[int(i)] for i in rng.integers(0, len(words), int(rng.integers(0, 12))))
        lines.append(body)
    return np.frombuffer(b"\n".join(lines) + (b"\n" if rng.random() < 0.5 else b""), dtype=np.uint8).copy()


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_random_geometry(double, tmp_path, seed):
    """Random files x random (min_chunk, segment, readers): carries of every length, segments that hold no complete
    chunk, one-off buffers, an unterminated last line -- every tag equals the oracle on the same chunk table."""
    oracle, _ = oracle_lib.load(build=False)
    rng = np.random.default_rng(seed)
    buf = _ragged_text(rng, 4000)
    path = tmp_path / "ragged.txt"
    buf.tofile(path)
    for _ in range(4):
        min_chunk = int(rng.choice([4096, 20000, 65536, 300000]))
        seg = int(min_chunk * rng.choice([2, 3, 7]))
        readers = int(rng.integers(1, 5))
        bounds = oracle.chunk_bounds(buf, min_chunk)
        extra = ["--min-chunk", str(min_chunk), "--segment-bytes", str(seg), "--readers", str(readers),
                 "--reader", str(rng.choice(["pread", "mmap"]))]
        for tag in capi.TAGS:
            live = ["--live"] if rng.random() < 0.5 else []
            exp = capi.expected_output(oracle, buf, b"needle", bounds, tag)
            assert capi.run_xsgrep(double, tag, b"needle", path, extra + live) == exp, (seed, min_chunk, seg, readers, tag)


def test_injected_failure_closes_the_result(double, tmp_path):
    """A failing push (XSG_DOUBLE_FAIL_PUSH) must end the search: exit code 2 with the library's message, results of
    the segments before it still delivered in order, no hang -- joined and live, one and several readers."""
    oracle, _ = oracle_lib.load(build=False)
    buf = _ragged_text(np.random.default_rng(9), 3000)
    path = tmp_path / "f.txt"
    buf.tofile(path)
    full = capi.expected_output(oracle, buf, b"needle", oracle.chunk_bounds(buf, 20000), "match_byte_offsets")
    for fail_at in (0, 1, 5):
        for extra in ([], ["--live", "--readers", "3"]):
            env = dict(os.environ, XSG_DOUBLE_FAIL_PUSH=str(fail_at))
            out = subprocess.run([double, "--tag", "match_byte_offsets", "--min-chunk", "20000", "--segment-bytes", "40000",
                                  *extra, "needle", str(path)], capture_output=True, text=True, env=env, timeout=60)
            assert out.returncode == 2, (fail_at, extra, out.stderr)
            assert "injected failure of push %d" % fail_at in out.stderr
            assert full.startswith(out.stdout) and len(out.stdout) < len(full), (fail_at, extra)
            if fail_at == 0:
                assert out.stdout == ""


def test_thread_sanitizer(double_tsan, tmp_path):
    """Readers, pusher, verification threads and a live consumer under ThreadSanitizer: no report, same output."""
    oracle, _ = oracle_lib.load(build=False)
    buf = _ragged_text(np.random.default_rng(5), 1500)
    path = tmp_path / "t.txt"
    buf.tofile(path)
    bounds = oracle.chunk_bounds(buf, 20000)
    env = dict(os.environ, TSAN_OPTIONS="halt_on_error=1 exitcode=66")
    base = ["--min-chunk", "20000", "--segment-bytes", "60000", "--readers", "4", "-j", "4"]
    for tag in ("count", "lines", "line_indices"):
        for live in ([], ["--live"], ["--live", "--reader", "mmap"]):
            out = subprocess.run([double_tsan, "--tag", tag, *base, *live, "needle", str(path)], capture_output=True,
                                 env=env, timeout=300)
            assert out.returncode == 0 and b"ThreadSanitizer" not in out.stderr, out.stderr.decode()[-2000:]
            assert out.stdout.decode("latin-1") == capi.expected_output(oracle, buf, b"needle", bounds, tag), (tag, live)
    # parts of more than 16 k matching lines: their strings are built on -j threads
    dense = np.frombuffer(b"".join(b"needle %d\n" % i for i in range(120000)), dtype=np.uint8).copy()
    dpath = tmp_path / "dense.txt"
    dense.tofile(dpath)
    out = subprocess.run([double_tsan, "--tag", "lines", "--min-chunk", "400000", "--segment-bytes", "800000", "--readers", "2",
                          "-j", "4", "--live", "needle", str(dpath)], capture_output=True, env=env, timeout=300)
    assert out.returncode == 0 and b"ThreadSanitizer" not in out.stderr, out.stderr.decode()[-2000:]
    assert out.stdout == dense.tobytes()
    # regex mode: candidate verification on four threads with per-thread DFA caches
    out = subprocess.run([double_tsan, "-E", "--tag", "count_lines", *base, "--live", "ne+d(le|y)", str(path)],
                         capture_output=True, env=env, timeout=300)
    assert out.returncode == 0 and b"ThreadSanitizer" not in out.stderr, out.stderr.decode()[-2000:]
    import re
    lines = buf.tobytes().split(b"\n")
    assert int(out.stdout) == sum(1 for line in lines if re.search(rb"ne+d(le|y)", line))
    # match-level kinds run the Pike VM (thread-local scratch) instead of the per-thread DFA
    out = subprocess.run([double_tsan, "-E", "--tag", "count", *base, "ne+d(le|y)", str(path)], capture_output=True, env=env,
                         timeout=300)
    assert out.returncode == 0 and b"ThreadSanitizer" not in out.stderr, out.stderr.decode()[-2000:]
    assert int(out.stdout) == sum(len(re.findall(rb"ne+d(le|y)", line)) for line in lines)


def test_address_sanitizer(double_asan, tmp_path, monkeypatch):
    """The carry arithmetic (headroom copies, one-off buffers, string building from recycled buffers) under
    AddressSanitizer + UBSan with leak checking: the edge files, a meta-file run, a regular expression and the failure
    path.  A report makes the binary exit with 67 (UBSan aborts), which the check functions see as a failed run."""
    monkeypatch.setenv("ASAN_OPTIONS", "detect_leaks=1:exitcode=67")
    capi.check_edge_files(double_asan, tmp_path)
    capi.check_with_meta_file(double_asan, tmp_path)
    capi.check_errors(double_asan, tmp_path)
    for name in ("empty", "no_trailing_newline", "long_line"):  # files written by check_edge_files
        out = subprocess.run([double_asan, "--tag", "lines", "--reader", "mmap", "--min-chunk", "65536", "--segment-bytes",
                              "262144", "--readers", "3", "needle", str(tmp_path / (name + ".txt"))], capture_output=True, timeout=300)
        assert out.returncode == 0 and b"Sanitizer" not in out.stderr, out.stderr.decode()[-3000:]
    buf = _ragged_text(np.random.default_rng(6), 1500)
    path = tmp_path / "a.txt"
    buf.tofile(path)
    for tag in ("lines", "match_byte_offsets"):
        out = subprocess.run([double_asan, "-E", "--tag", tag, "--min-chunk", "20000", "--segment-bytes", "60000", "--readers", "3",
                              "-j", "3", "--live", "n+e{2}dle|stack", str(path)], capture_output=True, timeout=300)
        assert out.returncode == 0 and b"Sanitizer" not in out.stderr, out.stderr.decode()[-3000:]
    monkeypatch.setenv("XSG_DOUBLE_FAIL_PUSH", "2")
    out = subprocess.run([double_asan, "--tag", "lines", "--min-chunk", "20000", "--segment-bytes", "40000", "--readers", "3",
                          "needle", str(path)], capture_output=True, timeout=300)
    assert out.returncode == 2 and b"Sanitizer" not in out.stderr, out.stderr.decode()[-3000:]

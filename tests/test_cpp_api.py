"""The C++20 host layer (include/xsearch_b200/xsearch.hpp) through its grep-like driver xsearch_b200/cpp/xsgrep.

CPU part: it builds, the result collector behaves like the reference's (blocking iterator, running totals,
close), argument errors are exceptions, the meta-file reader parses the v1 format.
GPU part (-m gpu): every tag of xs::extern_search against the CPU oracle on files, joined and live, with one
and several readers, with and without a meta file, with carries longer than the headroom."""
import json
import os
import struct
import subprocess

import numpy as np
import pytest

import oracle_lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPP_DIR = os.path.join(ROOT, "xsearch_b200", "cpp")
XSGREP = os.path.join(CPP_DIR, "xsgrep")
GOLD = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def xsgrep():
    try:  # rebuild when the header changed; a prebuilt binary is good enough where make/g++ are missing
        subprocess.check_call(["make", "-s", "-C", CPP_DIR], stdout=subprocess.DEVNULL)
    except (OSError, subprocess.CalledProcessError):
        if not os.path.exists(XSGREP):
            raise
    return XSGREP


def write_meta(path, bounds, compression=1, mappings=True):
    """v1 meta file: u32 type, then per chunk 5 x u64 + n_map x (u64, u64)."""
    with open(path, "wb") as f:
        f.write(struct.pack("<I", compression))
        prev = 0
        for e in bounds:
            e = int(e)
            maps = [(prev, 0)] if mappings else []
            f.write(struct.pack("<5Q", prev, prev, e - prev, e - prev, len(maps)))
            for m in maps:
                f.write(struct.pack("<2Q", *m))
            prev = e


def test_host_selftest(xsgrep):
    out = subprocess.run([xsgrep, "--selftest-host"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert "selftest-host ok" in out.stdout


def test_metafile_reader(xsgrep, tmp_path):
    p = tmp_path / "x.meta"
    write_meta(p, [1000, 2500, 2600])
    out = subprocess.run([xsgrep, "--metafile-cat", str(p)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    lines = out.stdout.splitlines()
    assert lines[0] == "compression_type: 1" and lines[1] == "chunks: 3" and lines[2] == "original_size: 2600"
    assert lines[4].startswith("original_offset=1000 actual_offset=1000 original_size=1500 ")
    # truncated file -> error, not garbage
    raw = p.read_bytes()
    (tmp_path / "bad.meta").write_bytes(raw[:-5])
    out = subprocess.run([xsgrep, "--metafile-cat", str(tmp_path / "bad.meta")], capture_output=True, text=True)
    assert out.returncode != 0


def test_metafile_reader_on_reference_fixture(xsgrep):
    """The chunk table of the reference's own fixture test/files/sample.meta (committed as a small golden)."""
    gold = json.load(open(os.path.join(GOLD, "sample_meta_chunks.json")))
    fixture = "/root/reference/test/files/sample.meta"
    if not os.path.exists(fixture):
        pytest.skip("reference tree not mounted (GPU box)")
    out = subprocess.run([xsgrep, "--metafile-cat", fixture], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    lines = out.stdout.splitlines()
    assert lines[1] == "chunks: %d" % len(gold["chunks"]) and lines[2] == "original_size: %d" % gold["original_size"]
    for line, c in zip(lines[3:], gold["chunks"]):
        assert line.startswith("original_offset=%d actual_offset=%d original_size=%d actual_size=%d mappings=%d " %
                               (c["original_offset"], c["actual_offset"], c["original_size"], c["actual_size"], c["n_map"]))


# ---------------------------------------------------------------------------------------------------------
TAGS = ["count", "count_lines", "match_byte_offsets", "line_byte_offsets", "line_indices", "lines"]


def expected_output(oracle, buf, pat, bounds, tag):
    if tag == "count":
        return "%d\n" % oracle.count(buf, pat, bounds, False)
    if tag == "count_lines":
        return "%d\n" % oracle.count(buf, pat, bounds, True)
    if tag == "lines":
        st, ln = oracle.lines(buf, pat, bounds)
        raw = buf.tobytes()
        return "".join(raw[int(s):int(s) + int(l)].decode("latin-1") + "\n" for s, l in zip(st, ln))
    v = {"match_byte_offsets": oracle.match_byte_offsets, "line_byte_offsets": oracle.line_byte_offsets,
         "line_indices": oracle.line_indices}[tag](buf, pat, bounds)
    return "".join("%d\n" % int(x) for x in v)


def run_xsgrep(xsgrep, tag, pat, path, extra=()):
    out = subprocess.run([xsgrep, "--tag", tag, *extra, pat.decode("latin-1"), str(path)], capture_output=True)
    assert out.returncode == 0, out.stderr.decode()
    return out.stdout.decode("latin-1")


def check_all_tags(xsgrep, tmp_path):
    import xsearch_b200 as xs
    oracle, _ = oracle_lib.load(build=False)
    p = xs.SynthParams.make(9, b"Sherlock", 30.0 / 8192)
    buf = xs.synth_host(4096 * 6000, p)  # ~24.6 MB
    path = tmp_path / "corpus.txt"
    buf.tofile(path)
    min_chunk, seg = 1 << 20, 4 << 20
    bounds = oracle.chunk_bounds(buf, min_chunk)
    base = ["--min-chunk", str(min_chunk), "--segment-bytes", str(seg)]
    for pat in (b"Sherlock", b"the"):
        for tag in TAGS:
            exp = expected_output(oracle, buf, pat, bounds, tag)
            assert run_xsgrep(xsgrep, tag, pat, path, base) == exp, (tag, pat)
            assert run_xsgrep(xsgrep, tag, pat, path, base + ["--live", "--readers", "3", "-j", "4"]) == exp, (tag, pat, "live")


def check_with_meta_file(xsgrep, tmp_path):
    import xsearch_b200 as xs
    oracle, _ = oracle_lib.load(build=False)
    import gen_cases
    # hazard corpus: chunks whose last line holds prefix hazards ("SSherlock") inside the scalar-tail zone,
    # so the result depends on the chunk table -- which here comes from the meta file
    case = {"kind": "hazard", "pattern": "Sherlock", "seed": 5, "alphabet": "Sherlock abc", "chunks": 40,
            "body_min": 20000, "body_max": 60000}
    buf, pat, bounds = gen_cases.materialize(case)
    path, meta = tmp_path / "hazard.txt", tmp_path / "hazard.meta"
    buf.tofile(path)
    write_meta(meta, bounds)
    differs = False
    for tag in TAGS:
        exp = expected_output(oracle, buf, pat, bounds, tag)
        got = run_xsgrep(xsgrep, tag, pat, path, ["--meta", str(meta), "--segment-bytes", "300000", "--min-chunk", "65536",
                                                  "--readers", "2"])
        assert got == exp, tag
        other = expected_output(oracle, buf, pat, oracle.chunk_bounds(buf, 65536), tag)
        differs = differs or other != exp
    assert differs, "the hazard corpus should make the chunk table observable"
    # compressed meta files are rejected with a clear error
    write_meta(meta, bounds, compression=2)
    out = subprocess.run([xsgrep, "--meta", str(meta), "Sherlock", str(path)], capture_output=True, text=True)
    assert out.returncode != 0 and "compressed" in out.stderr


def check_edge_files(xsgrep, tmp_path):
    oracle, _ = oracle_lib.load(build=False)
    rng = np.random.default_rng(3)
    import gen_cases
    min_chunk = 65536
    base = ["--min-chunk", str(min_chunk), "--segment-bytes", str(4 * min_chunk), "--readers", "2"]
    cases = {
        "empty": np.zeros(0, np.uint8),
        "no_trailing_newline": np.frombuffer(b"one needle\ntwo\nneedle needle", dtype=np.uint8).copy(),
        # one line far longer than the carry headroom (min_chunk + 1 MiB) in the middle of normal text
        "long_line": np.concatenate([gen_cases._random_text(rng, 300000, b"abc needle\n", 0.0),
                                     np.frombuffer(b"\n" + b"x" * 2500000 + b"needle" + b"y" * 40 + b"\n", np.uint8),
                                     gen_cases._random_text(rng, 400000, b"abc needle\n", 0.0),
                                     np.frombuffer(b"\n", np.uint8)]),
    }
    for name, buf in cases.items():
        path = tmp_path / (name + ".txt")
        buf.tofile(path)
        bounds = oracle.chunk_bounds(buf, min_chunk) if buf.size else np.zeros(0, np.uint64)
        for tag in TAGS:
            exp = expected_output(oracle, buf, b"needle", bounds, tag) if buf.size else ("0\n" if tag.startswith("count") else "")
            assert run_xsgrep(xsgrep, tag, b"needle", path, base + ["--live"]) == exp, (name, tag)


def check_ignore_case(xsgrep, tmp_path):
    """The ignore_case overloads (README.md:37, example/grep.cpp:69-77): lines come back with their ORIGINAL bytes."""
    oracle, _ = oracle_lib.load(build=False)
    text = (b"Sherlock holmes\nno match here\nSHERLOCK again, sherlock twice\nshErLoCk\n" * 3000)
    buf = np.frombuffer(text, dtype=np.uint8).copy()
    path = tmp_path / "mixed.txt"
    buf.tofile(path)
    bounds = oracle.chunk_bounds(buf, 65536)
    lbuf = oracle.to_lower(buf)
    base = ["--min-chunk", "65536", "--segment-bytes", "262144", "-i"]
    for tag in TAGS:
        exp = expected_output(oracle, lbuf, b"sherlock", bounds, tag)
        if tag == "lines":  # same lines, original spelling
            st, ln = oracle.lines(lbuf, b"sherlock", bounds)
            exp = "".join(text[int(s):int(s) + int(l)].decode() + "\n" for s, l in zip(st, ln))
        assert run_xsgrep(xsgrep, tag, b"SherLOCK", path, base) == exp, tag
    assert run_xsgrep(xsgrep, "count", b"SherLOCK", path, base[:-1]) == "0\n"


def check_multi_device(xsgrep, tmp_path, devices):
    """ONE Searcher over several devices (xs::Options::devices, xsgrep --devices): segments dealt round-robin, output
    byte-identical to the oracle (= to the single-device run) for all six tags, also live, with a meta file's chunk
    table, and when there are fewer segments than devices."""
    import xsearch_b200 as xs
    oracle, _ = oracle_lib.load(build=False)
    p = xs.SynthParams.make(12, b"Sherlock", 30.0 / 8192)
    buf = xs.synth_host(4096 * 5000, p)  # ~20 MB -> ~10 segments of 2 MiB
    path = tmp_path / "multi.txt"
    buf.tofile(path)
    min_chunk, seg = 1 << 19, 2 << 20
    bounds = oracle.chunk_bounds(buf, min_chunk)
    dev = ["--devices", ",".join(str(d) for d in devices)]
    base = ["--min-chunk", str(min_chunk), "--segment-bytes", str(seg)]
    for pat in (b"Sherlock", b"the"):
        for tag in TAGS:
            exp = expected_output(oracle, buf, pat, bounds, tag)
            assert run_xsgrep(xsgrep, tag, pat, path, base + dev) == exp, (tag, pat)
            assert run_xsgrep(xsgrep, tag, pat, path, base + dev + ["--live", "--readers", "3", "-j", "2"]) == exp, (tag, pat, "live")
    # one segment only: the other devices stay idle
    big = ["--min-chunk", str(min_chunk), "--segment-bytes", str(1 << 30)]
    for tag in ("line_indices", "lines", "count"):
        assert run_xsgrep(xsgrep, tag, b"Sherlock", path, big + dev) == expected_output(oracle, buf, b"Sherlock", bounds, tag)
    # chunk table from a meta file
    meta = tmp_path / "multi.meta"
    mb = oracle.chunk_bounds(buf, 300000)
    write_meta(meta, mb)
    for tag in ("line_indices", "match_byte_offsets", "count_lines"):
        got = run_xsgrep(xsgrep, tag, b"Sherlock", path, ["--meta", str(meta), "--segment-bytes", "1000000", "--min-chunk", "65536"] + dev)
        assert got == expected_output(oracle, buf, b"Sherlock", mb, tag), tag
    err = subprocess.run([xsgrep, "--tag", "count", *dev, "Sherlock", str(path)], capture_output=True, text=True).stderr
    assert "%d device(s)" % len(devices) in err


def check_errors(xsgrep, tmp_path):
    out = subprocess.run([xsgrep, "abc", str(tmp_path / "missing.txt")], capture_output=True, text=True)
    assert out.returncode == 1 and "cannot open" in out.stderr


# the same checks run on the CPU against the test double of the C ABI in tests/test_host_pipeline.py
@pytest.mark.gpu
def test_extern_search_all_tags(xsgrep, tmp_path):
    check_all_tags(xsgrep, tmp_path)


@pytest.mark.gpu
def test_extern_search_with_meta_file(xsgrep, tmp_path):
    check_with_meta_file(xsgrep, tmp_path)


@pytest.mark.gpu
def test_extern_search_edge_files(xsgrep, tmp_path):
    check_edge_files(xsgrep, tmp_path)


@pytest.mark.gpu
def test_extern_search_ignore_case(xsgrep, tmp_path):
    check_ignore_case(xsgrep, tmp_path)


@pytest.mark.gpu
def test_extern_search_multi_device(xsgrep, tmp_path):
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    check_multi_device(xsgrep, tmp_path, list(range(min(n, 4))))


@pytest.mark.gpu
def test_extern_search_errors(xsgrep, tmp_path):
    check_errors(xsgrep, tmp_path)

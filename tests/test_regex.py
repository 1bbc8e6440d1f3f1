"""Regular expressions over the GPU literal scan (include/xsearch_b200/regex_prefilter.hpp).

Checker: the reference's own regex wrappers (search_wrappers.h:63-88,209-271) over its vendored re2, compiled where they
lie into oracle/_ref/libxsref_re.so (oracle/ref_regex_shim.cpp), and the fixture tests/golden/regex_vectors.json made
from it (tools/make_golden.py --regex) for boxes without that library.

CPU part: the checker reproduces the reference's own known answers (search_wrappersTest.cpp:74-105); the fixture agrees
with the checker; literal extraction; the host stage (required literals + our regex VM / lazy DFA over candidate lines, driven by
tests/cpp/regex_host_check.cpp, which replaces the GPU scan by a host loop) agrees with the reference on every case.
GPU part (-m gpu): `xsgrep -E` = xs::extern_search with a regular expression, every tag, against the same expectations."""
import hashlib
import json
import os
import subprocess

import numpy as np
import pytest

import oracle_lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPP_DIR = os.path.join(ROOT, "xsearch_b200", "cpp")
XSGREP = os.path.join(CPP_DIR, "xsgrep")
GOLD = os.path.join(ROOT, "tests", "golden")
TAGS = ["count", "count_lines", "match_byte_offsets", "line_byte_offsets", "line_indices", "lines"]


def corpora():
    """name -> bytes (np.uint8).  Deterministic: the library's host-side generator and hand-made hazards."""
    import xsearch_b200 as xs
    out = {}
    p = xs.SynthParams.make(9, b"Sherlock", 30.0 / 2048)
    out["synth"] = xs.synth_host(4096 * 400, p)  # 1.6 MB of English-like lines
    out["dummy"] = np.fromfile(os.path.join(GOLD, "wrappers_dummy_text.bin"), dtype=np.uint8)
    # hazards of the literal scan's chunk-tail emulation: "SSherlock" / "llock" in the last bytes of the
    # file and of every 64 KiB segment, mixed case, an unterminated last line
    rng = np.random.default_rng(11)
    words = [b"Sherlock", b"SSherlock", b"She lock", b"sherlock", b"SHERLOCK", b"llock", b"lolock", b"She\tlock", b"abc",
             b"Shelock", b"Sheerlock", b"holmes", b"x"]
    lines = []
    for _ in range(30000):
        k = int(rng.integers(1, 6))
        lines.append(b" ".join(words[int(i)] for i in rng.integers(0, len(words), k)))
    out["hazard"] = np.frombuffer(b"\n".join(lines) + b"\nlast line SSherlock", dtype=np.uint8).copy()
    return out


# (corpus, pattern, ignore_case)
CASES = [
    ("dummy", b"(a[n|m]t)", False),            # the reference's own regex test pattern (search_wrappersTest.cpp:74-105)
    ("synth", b"She[r ]lock", False),          # the pattern of the v1 golden vectors (xsearchTest.cpp:17-335)
    ("synth", b"(a[n|l]d)", False),
    ("synth", b"th(e|is|at)\\b", False),
    ("synth", b"(?:Sherlock|which) [a-z]+", False),
    ("synth", b"wh(ich|ere)|Sherlock", False),  # three required literals -> three GPU streams
    ("synth", b"[A-Z][a-z]+lock", False),
    ("synth", b"o{2}k", False),
    ("synth", b"sHE[r ]LOCK", True),
    ("hazard", b"She[r ]lock", False),
    ("hazard", b"S+her?lock", False),
    ("hazard", b"she[r\\t ]lock", True),
    ("hazard", b"(SS|l)lock", False),
]


def case_key(corpus, pattern, icase):
    return "%s|%s|%d" % (corpus, pattern.decode(), int(icase))


def digest(text: str):
    raw = text.encode("latin-1")
    return {"n": raw.count(b"\n"), "sha256": hashlib.sha256(raw).hexdigest(), "head": text[:120]}


def expected_from_ref(ref, buf, pattern, icase):
    bounds = [buf.size]  # re2 has no chunk-dependent behaviour for line-local matches
    return {tag: digest(ref.expected_output(buf, pattern, bounds, tag, icase)) for tag in TAGS}


@pytest.fixture(scope="module")
def golden():
    return json.load(open(os.path.join(GOLD, "regex_vectors.json")))


@pytest.fixture(scope="module")
def data():
    return corpora()


@pytest.fixture(scope="module")
def xsgrep():
    try:
        subprocess.check_call(["make", "-s", "-C", CPP_DIR], stdout=subprocess.DEVNULL)
    except (OSError, subprocess.CalledProcessError):
        if not os.path.exists(XSGREP):
            raise
    return XSGREP


@pytest.fixture(scope="module")
def host_check(tmp_path_factory):
    exe = tmp_path_factory.mktemp("bin") / "regex_host_check"
    subprocess.check_call(["g++", "-std=c++20", "-O2", "-I" + os.path.join(ROOT, "include"), "-o", str(exe),
                           os.path.join(ROOT, "tests", "cpp", "regex_host_check.cpp")])
    return str(exe)


def test_checker_reproduces_reference_kat():
    """search_wrappersTest.cpp:74-105: offsets {2,151,197,507}, line offsets {0,113,176,460}, count 4."""
    ref = oracle_lib.load_regex_ref()
    if ref is None:
        pytest.skip("oracle/_ref/libxsref_re.so not built (reference tree absent)")
    buf = np.fromfile(os.path.join(GOLD, "wrappers_dummy_text.bin"), dtype=np.uint8)
    b = [buf.size]
    assert ref.match_byte_offsets(buf, b"(a[n|m]t)", b).tolist() == [2, 151, 197, 507]
    assert ref.line_byte_offsets(buf, b"(a[n|m]t)", b).tolist() == [0, 113, 176, 460]
    assert ref.count(buf, b"(a[n|m]t)", b, True) == 4
    # chunking at line starts does not change a line-local result
    cut = int(np.flatnonzero(buf == 10)[3]) + 1
    assert ref.match_byte_offsets(buf, b"(a[n|m]t)", [cut, buf.size]).tolist() == [2, 151, 197, 507]


def test_fixture_matches_checker(golden, data):
    ref = oracle_lib.load_regex_ref()
    if ref is None:
        pytest.skip("oracle/_ref/libxsref_re.so not built (reference tree absent)")
    for corpus, pattern, icase in CASES:
        assert golden["cases"][case_key(corpus, pattern, icase)] == expected_from_ref(ref, data[corpus], pattern, icase)
    for name, buf in data.items():
        assert golden["corpora"][name] == hashlib.sha256(buf.tobytes()).hexdigest()


def test_cases_are_not_trivial(golden):
    for key, exp in golden["cases"].items():
        assert exp["count_lines"]["head"] != "0\n", key
        assert exp["count"]["n"] == 1 and exp["line_byte_offsets"]["n"] == exp["line_indices"]["n"]


def test_required_literals(xsgrep):
    def lits(p):
        out = subprocess.run([xsgrep, "--required-literals", p], capture_output=True, text=True)
        if out.returncode != 0:
            return None
        rows = out.stdout.splitlines()
        return int(rows[0].split()[1]), [r[len("literal: "):] for r in rows[1:]]
    assert lits("She[r ]lock") == (1, ["lock"])
    assert lits("Sherlock") == (0, ["Sherlock"])           # plain text by the reference's rule (utils.h:17-24)
    assert lits("wh(ich|ere)|Sherlock") == (1, ["ich", "ere", "Sherlock"])
    assert lits("(?:Sherlock|which) [a-z]+") == (1, ["Sherlock", "which"])
    assert lits("th(e|is|at)\\b") == (1, ["th"])
    assert lits("colou?r") == (1, ["colo"])
    assert lits("x" * 80 + "+") == (1, ["x" * 64])         # longer than the GPU's pattern limit: a prefix is still required
    assert lits("a\\.b\\\\c") == (1, ["a.b\\c"])
    assert lits("(abc|abcd|zabc)+") == (1, ["abc"])        # literals containing another one are redundant
    for bad in ("a*", "[ab]+", "ab|cd|", "x|.", "(?i)abc", "(?=a)b", "(a)\\1", "a\\nb", "\\pL+a"):
        assert lits(bad) is None, bad


def test_host_stage_matches_reference(golden, data, host_check, tmp_path):
    """Literal extraction + xs::LineVerifier (regex_vm.hpp) over the candidate lines == the reference's re2 wrappers, every tag."""
    paths = {}
    for name, buf in data.items():
        paths[name] = tmp_path / (name + ".txt")
        buf.tofile(paths[name])
    for corpus, pattern, icase in CASES:
        exp = golden["cases"][case_key(corpus, pattern, icase)]
        for tag in TAGS:
            out = subprocess.run([host_check, tag, pattern.decode(), str(paths[corpus])] + (["-i"] if icase else []),
                                 capture_output=True)
            assert out.returncode == 0, (pattern, out.stderr.decode())
            assert digest(out.stdout.decode("latin-1")) == exp[tag], (corpus, pattern, icase, tag)


def check_extern_search_regex(golden, data, xsgrep, tmp_path):
    """xs::extern_search with a regular expression: GPU scan for the required literals, host verification."""
    paths = {}
    for name, buf in data.items():
        paths[name] = tmp_path / (name + ".txt")
        buf.tofile(paths[name])
    # small segments: many segment ends (where the unconditional tail candidates come from), carries across them
    base = ["--min-chunk", "65536", "--segment-bytes", "262144"]
    for corpus, pattern, icase in CASES:
        exp = golden["cases"][case_key(corpus, pattern, icase)]
        for tag in TAGS:
            for extra in ([], ["--live", "--readers", "3", "-j", "4"]):
                if extra and (tag in ("count_lines", "match_byte_offsets", "line_byte_offsets") or corpus == "synth"):
                    continue  # live + several readers / threads: the hazard corpus, half of the tags (keeps the test short)
                out = subprocess.run([xsgrep, "-E", "--tag", tag, *base, *extra, *(["-i"] if icase else []),
                                      pattern.decode(), str(paths[corpus])], capture_output=True)
                assert out.returncode == 0, (pattern, out.stderr.decode())
                assert digest(out.stdout.decode("latin-1")) == exp[tag], (corpus, pattern, icase, tag, extra)
    # the reference's rule picks the regex reading by itself; -F forces plain text
    exp = golden["cases"][case_key("synth", b"She[r ]lock", False)]
    out = subprocess.run([xsgrep, "--tag", "count", *base, "She[r ]lock", str(paths["synth"])], capture_output=True)
    assert out.returncode == 0 and digest(out.stdout.decode()) == exp["count"]
    out = subprocess.run([xsgrep, "-F", "--tag", "count", *base, "She[r ]lock", str(paths["synth"])], capture_output=True)
    assert out.returncode == 0 and out.stdout == b"0\n"
    # nothing to scan for -> a clear error before any work starts
    out = subprocess.run([xsgrep, "-E", "[a-z]+", str(paths["synth"])], capture_output=True, text=True)
    assert out.returncode == 1 and "no literal" in out.stderr


@pytest.mark.gpu
def test_extern_search_regex(golden, data, xsgrep, tmp_path):
    check_extern_search_regex(golden, data, xsgrep, tmp_path)


@pytest.mark.gpu
def test_extern_search_regex_against_checker_large(xsgrep, tmp_path):
    """A 40 MB file with the default 16 MiB chunks, against the reference library itself where it travelled."""
    ref = oracle_lib.load_regex_ref()
    if ref is None:
        pytest.skip("oracle/_ref/libxsref_re.so did not travel")
    import xsearch_b200 as xs
    p = xs.SynthParams.make(21, b"Sherlock", 30.0 / 8192)
    buf = xs.synth_host(4096 * 10000, p)
    path = tmp_path / "big.txt"
    buf.tofile(path)
    for pattern in (b"She[r ]lock", b"mo(rn|n)[a-z]+ "):
        for tag in TAGS:
            out = subprocess.run([xsgrep, "-E", "--tag", tag, "--segment-bytes", str(32 << 20), "--readers", "2", "-j", "4",
                                  pattern.decode(), str(path)], capture_output=True)
            assert out.returncode == 0, out.stderr.decode()
            assert out.stdout.decode("latin-1") == ref.expected_output(buf, pattern, [buf.size], tag), (pattern, tag)


def _random_expression(rng, depth=0):
    """A random expression of the supported grammar over a three-letter alphabet (dense matches, many overlaps)."""
    def atom():
        r = rng.random()
        if r < 0.45 or depth > 2:
            return "".join(rng.choice(list("abc")) for _ in range(int(rng.integers(1, 4))))
        if r < 0.60:
            return "[" + "".join(sorted(set(rng.choice(list("abc ")) for _ in range(2)))) + "]"
        if r < 0.65:
            return "."
        inner = "|".join(_random_expression(rng, depth + 1) for _ in range(int(rng.integers(1, 4))))
        return ("(?:" if rng.random() < 0.5 else "(") + inner + ")"
    parts = []
    for _ in range(int(rng.integers(1, 4))):
        a = atom()
        q = rng.random()
        if q < 0.12:
            a = (a if len(a) == 1 or a[0] in "([." else "(?:" + a + ")") + rng.choice(["*", "+", "?", "{2}", "{1,2}", "{0,1}", "+?"])
        parts.append(a)
    return "".join(parts)


def test_prefilter_never_loses_a_line(host_check, tmp_path):
    """Soundness of required_literals on random expressions: regex_host_check fails (exit 2) when a line matches the
    expression but holds none of the literals; its line count must also equal Python's re on the same lines."""
    import re
    rng = np.random.default_rng(2024)
    text = "\n".join("".join(rng.choice(list("abc "), p=[0.35, 0.3, 0.25, 0.1]) for _ in range(int(rng.integers(0, 40))))
                     for _ in range(600)) + "\n"
    path = tmp_path / "abc.txt"
    path.write_text(text)
    lines = text.split("\n")[:-1]
    usable = 0
    for _ in range(150):
        expr = _random_expression(rng)
        out = subprocess.run([host_check, "count_lines", expr, str(path)], capture_output=True, text=True)
        if out.returncode == 1:  # no required literal: rejected, nothing to check
            assert "no literal" in out.stderr or "not usable" in out.stderr, (expr, out.stderr)
            continue
        assert out.returncode == 0, (expr, out.stderr)
        usable += 1
        assert int(out.stdout) == sum(1 for ln in lines if re.search(expr, ln)), expr
    assert usable > 60


def _random_rich_expression(rng, depth=0):
    """Random expression using everything the VM supports: classes, ranges, negation, perl classes, '.', word
    boundaries, anchors, greedy / lazy / counted repetition, nested groups and alternation."""
    def atom():
        r = rng.random()
        if r < 0.40 or depth > 2:
            return "".join(rng.choice(list("abcAB1")) for _ in range(int(rng.integers(1, 4))))
        if r < 0.50:
            return rng.choice(["[a-c]", "[^ab]", "[ab1]", "[A-Ca]", "[^a-c ]", "[\\w]", "[^\\d]", "[a\\-c]", "[[:digit:]b]"])
        if r < 0.58:
            return rng.choice(["\\w", "\\d", "\\s", "\\W", "\\D", "\\S", "."])
        if r < 0.64:
            return rng.choice(["\\b", "\\B", "^", "$"])
        inner = "|".join(_random_rich_expression(rng, depth + 1) for _ in range(int(rng.integers(1, 4))))
        return ("(?:" if rng.random() < 0.5 else "(") + inner + ")"
    parts = []
    for _ in range(int(rng.integers(1, 5))):
        a = atom()
        if rng.random() < 0.25 and a not in ("\\b", "\\B", "^", "$"):
            if not (len(a) == 1 or a[0] in "([.\\"):
                a = "(?:" + a + ")"
            a += rng.choice(["*", "+", "?", "{2}", "{1,2}", "{0,1}", "{2,}", "+?", "*?", "??", "{1,3}?"])
        parts.append(a)
    return "".join(parts)


def test_vm_matches_re2_on_random_expressions(host_check, tmp_path):
    """The host matcher (regex_vm.hpp) against the reference's re2 wrappers: start offsets of all non-overlapping
    matches, line by line (a single line has no chunk- or newline-dependent behaviour in the reference either)."""
    ref = oracle_lib.load_regex_ref()
    if ref is None:
        pytest.skip("oracle/_ref/libxsref_re.so not built (reference tree absent)")
    rng = np.random.default_rng(int(os.environ.get("XS_REGEX_SEED", "77")))
    alphabet, weights = list("abcAB1 _") + ["é", "€"], [0.25, 0.2, 0.15, 0.08, 0.07, 0.08, 0.1, 0.03, 0.02, 0.02]
    lines = ["".join(rng.choice(alphabet, p=weights) for _ in range(int(rng.integers(1, 30)))) for _ in range(250)]
    raw_lines = [ln.encode("utf-8") for ln in lines]
    path = tmp_path / "rich.txt"
    path.write_bytes(b"\n".join(raw_lines) + b"\n")
    starts = np.cumsum([0] + [len(r) + 1 for r in raw_lines])
    checked = 0
    for k in range(220):
        expr = _random_rich_expression(rng)
        icase = k % 5 == 0
        out = subprocess.run([host_check, "match_byte_offsets", expr, str(path)] + (["-i"] if icase else []),
                             capture_output=True, text=True)
        if out.returncode == 1:  # no required literal (or syntax the VM rejects): nothing to compare
            continue
        assert out.returncode == 0, (expr, out.stderr)
        try:
            exp = [int(starts[i]) + int(x) for i, r in enumerate(raw_lines)
                   for x in ref.match_byte_offsets(r, expr.encode(), [len(r)], icase)]
        except ValueError:  # re2 rejects it
            continue
        got = [int(x) for x in out.stdout.split()]
        assert got == exp, (expr, icase)
        checked += 1
    assert checked > 60, checked


def test_dfa_resolves_assertions(host_check, tmp_path):
    """Programs with ^ $ \\b \\B run on the lazy DFA too (pending assertions + position flags in the states,
    regex_vm::Dfa::expand).  Assertion-heavy random expressions over short words: regex_host_check exits 3 when the
    DFA and the Pike VM disagree on any line, and its line count must equal Python's re (ASCII \\b on bytes) as well."""
    import re
    rng = np.random.default_rng(int(os.environ.get("XS_REGEX_SEED", "5")))
    alphabet = list("ab_1 -")
    lines = ["".join(rng.choice(alphabet, p=[0.3, 0.25, 0.05, 0.1, 0.25, 0.05]) for _ in range(int(rng.integers(0, 14))))
             for _ in range(800)]
    path = tmp_path / "words.txt"
    path.write_text("\n".join(lines) + "\n")
    raw = [ln.encode() for ln in lines]

    def piece():
        r = rng.random()
        if r < 0.35:
            return rng.choice(["\\b", "\\B", "^", "$"])
        if r < 0.75:
            return "".join(rng.choice(list("ab1")) for _ in range(int(rng.integers(1, 3))))
        if r < 0.85:
            return rng.choice(["[ab]", "[^a ]", "\\w", "\\W", ".", " "])
        if r < 0.93:
            return rng.choice(["a*", "b+", "[ab]?", "\\w{1,2}", " *"])
        return "(" + "|".join("".join(piece2() for _ in range(int(rng.integers(1, 4)))) for _ in range(2)) + ")"

    def piece2():
        return rng.choice(["\\b", "\\B", "$", "^", "a", "b", "ab", "1", "\\w", " "])

    checked = with_assertion = 0
    for _ in range(260):
        expr = "".join(piece() for _ in range(int(rng.integers(2, 6))))
        out = subprocess.run([host_check, "count_lines", expr, str(path)], capture_output=True, text=True)
        if out.returncode == 1:  # no required literal
            continue
        assert out.returncode == 0, (expr, out.stderr)
        try:
            cre = re.compile(expr.encode())
        except re.error:
            continue
        assert int(out.stdout) == sum(1 for r in raw if cre.search(r)), expr
        checked += 1
        with_assertion += any(t in expr for t in ("\\b", "\\B", "^", "$"))
    assert checked > 80 and with_assertion > 50, (checked, with_assertion)

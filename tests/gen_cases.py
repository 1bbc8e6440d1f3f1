"""Seeded test corpora shared by the oracle tests, the GPU parity tests and tools/make_golden.py.

A case is a small JSON-able dict; ``materialize(case)`` -> (uint8 array, pattern bytes, chunk bounds).
Patterns are latin-1 strings so bytes >= 0x80 can be expressed.
"""
import numpy as np


def _rng(seed):
    return np.random.default_rng(seed)


def _chunk_bounds(buf: np.ndarray, min_chunk: int) -> np.ndarray:
    """v1 rule: >= min_chunk bytes then through the next newline (same rule as the library/oracle)."""
    n = buf.size
    out, start = [], 0
    nl = np.flatnonzero(buf == 10)
    while start < n:
        if n - start <= min_chunk:
            end = n
        else:
            i = np.searchsorted(nl, start + min_chunk - 1)
            end = int(nl[i]) + 1 if i < nl.size else n
        out.append(end)
        start = end
    return np.array(out, dtype=np.uint64)


def _random_text(rng, n, alphabet: bytes, nl_prob: float) -> np.ndarray:
    a = np.frombuffer(alphabet, dtype=np.uint8)
    buf = a[rng.integers(0, a.size, size=n)].copy()
    if nl_prob > 0:
        buf[rng.random(n) < nl_prob] = 10
    return buf


def materialize(case):
    kind = case["kind"]
    pat = case["pattern"].encode("latin-1")
    rng = _rng(case["seed"])
    if kind == "random":
        buf = _random_text(rng, case["n"], case["alphabet"].encode("latin-1"), case.get("nl_prob", 0.0))
        if case.get("end_nl", False) and buf.size:
            buf[-1] = 10
        bounds = _chunk_bounds(buf, case["min_chunk"])
    elif kind == "planted":
        # background text with the pattern (and prefix-hazard variants of it) planted at random places
        buf = _random_text(rng, case["n"], case["alphabet"].encode("latin-1"), case.get("nl_prob", 0.03))
        p = np.frombuffer(pat, dtype=np.uint8)
        for _ in range(case["plants"]):
            k = int(rng.integers(0, 3))
            piece = np.concatenate([p[:k], p]) if k else p  # e.g. "SSherlock": partial match, then the match
            if piece.size >= buf.size:
                continue
            pos = int(rng.integers(0, buf.size - piece.size))
            buf[pos:pos + piece.size] = piece
        if case.get("end_nl", True) and buf.size:
            buf[-1] = 10
        bounds = _chunk_bounds(buf, case["min_chunk"])
    elif kind == "hazard":
        # explicit chunks; the last line of every chunk is filler + hazard + d filler bytes + '\n' so the
        # hazard sweeps through the last 40+m bytes (the scalar-tail zone of simd::strstr)
        p = np.frombuffer(pat, dtype=np.uint8)
        alphabet = case["alphabet"].encode("latin-1")
        parts, bounds, total = [], [], 0
        for c in range(case["chunks"]):
            body = _random_text(rng, int(rng.integers(case["body_min"], case["body_max"])), alphabet, 0.04)
            d = (c * case.get("d_step", 1) + case.get("d0", 0)) % (48 + p.size)
            k = 1 + c % max(1, min(3, p.size - 1)) if p.size > 1 else 0
            haz = np.concatenate([p[:k], p])
            tail = _random_text(rng, d, alphabet, 0.0)
            chunk = np.concatenate([body, np.array([10], np.uint8), _random_text(rng, 7, alphabet, 0.0), haz, tail,
                                    np.array([10], np.uint8)])
            parts.append(chunk)
            total += chunk.size
            bounds.append(total)
        buf = np.concatenate(parts) if parts else np.zeros(0, np.uint8)
        if not case.get("end_nl", True) and buf.size:
            buf = buf[:-1].copy()
            bounds[-1] -= 1
        bounds = np.array(bounds, dtype=np.uint64)
    elif kind == "repeat":
        unit = case["unit"].encode("latin-1")
        reps = case["n"] // len(unit) + 1
        buf = np.frombuffer((unit * reps)[:case["n"]], dtype=np.uint8).copy()
        if case.get("nl_every"):
            buf[case["nl_every"] - 1::case["nl_every"]] = 10
        bounds = _chunk_bounds(buf, case["min_chunk"])
    else:
        raise ValueError(kind)
    return np.ascontiguousarray(buf), pat, bounds


def golden_cases():
    """Small deterministic cases whose reference outputs are committed under tests/golden/."""
    cases = []
    seed = 1000
    for alphabet, pats in [("ab\n", ["a", "ab", "aab", "abab", "aaaa"]), ("abc\n", ["abc", "cab", "bcabc"]),
                           ("a\n", ["a", "aa", "aaa"]), ("ab", ["ab", "ba", "aab", "abba"])]:
        for pat in pats:
            for n in (0, 1, 31, 32, 33, 40, 64, 65, 100, 257, 1500):
                seed += 1
                cases.append({"kind": "random", "seed": seed, "n": n, "alphabet": alphabet, "pattern": pat,
                              "nl_prob": 0.0, "min_chunk": 300})
    for pat in ["Sherlock", "aab", "the", "xyxyz", "needle-needle-needle-needle-0032", "\xe9\xfc"]:
        for d0 in (0, 11, 23):
            seed += 1
            cases.append({"kind": "hazard", "seed": seed, "pattern": pat, "alphabet": "abcdefghij SSheSherl tthh xy",
                          "chunks": 12, "body_min": 40, "body_max": 400, "d0": d0, "d_step": 5})
        seed += 1
        cases.append({"kind": "hazard", "seed": seed, "pattern": pat, "alphabet": "abc SSh", "chunks": 5,
                      "body_min": 10, "body_max": 90, "d0": 3, "d_step": 9, "end_nl": False})
    for unit, pat in [("a", "aa"), ("a", "aaa"), ("ab", "abab"), ("abc", "abcabc"), ("aab", "aabaab")]:
        seed += 1
        cases.append({"kind": "repeat", "seed": seed, "n": 700, "unit": unit, "pattern": pat, "nl_every": 0,
                      "min_chunk": 256})
        seed += 1
        cases.append({"kind": "repeat", "seed": seed, "n": 700, "unit": unit, "pattern": pat, "nl_every": 97,
                      "min_chunk": 256})
    return cases


def random_cases(count, seed0, max_n=4000):
    """Differential cases (oracle vs reference vs GPU); not stored, regenerated from the seed."""
    rng = _rng(seed0)
    alphabets = ["ab\n", "abc\n", "a\n", "ab", "abcd \n", "SherlockS \n", "the t\n", "\xe9\xfca\n"]
    cases = []
    for i in range(count):
        alphabet = alphabets[int(rng.integers(0, len(alphabets)))]
        letters = [c for c in alphabet if c != "\n"]
        m = int(rng.integers(1, 9))
        pat = "".join(letters[int(rng.integers(0, len(letters)))] for _ in range(m))
        kind = ["random", "planted"][int(rng.integers(0, 2))]
        case = {"kind": kind, "seed": seed0 * 100003 + i, "n": int(rng.integers(0, max_n)), "alphabet": alphabet,
                "pattern": pat, "nl_prob": float(rng.choice([0.0, 0.02, 0.2])), "min_chunk": int(rng.integers(40, 600)),
                "plants": int(rng.integers(0, 30)), "end_nl": bool(rng.integers(0, 2))}
        cases.append(case)
    return cases


def icase_cases():
    """Mixed-case corpora for ignore_case (golden outputs: the reference's toLower + its matchers)."""
    cases = []
    seed = 5000
    for pat in ["Sherlock", "sHeRlOcK", "the", "AB", "a", "Z", "needle-NEEDLE-needle-Needle-0032", "\xc9\xfc"]:
        for n in (0, 33, 500, 3000):
            seed += 1
            cases.append({"kind": "planted", "seed": seed, "n": n, "alphabet": "sSHhEeRrLlOoCcKk tT aAbBzZ@[`{\xc9\xe9\xfc\xdc",
                          "pattern": pat, "nl_prob": 0.04, "min_chunk": 400, "plants": 12, "end_nl": bool(seed & 1)})
        seed += 1
        cases.append({"kind": "hazard", "seed": seed, "pattern": pat, "alphabet": "sShHeE rRlL oOcCkK", "chunks": 8,
                      "body_min": 40, "body_max": 300, "d0": 2, "d_step": 7})
    return cases

#!/usr/bin/env python3
"""Builds tests/golden/ from the reference tree (run in the container that has /root/reference).

  * simd_dummy_text.bin / wrappers_dummy_text.bin : the two fixture texts of the reference's unit tests,
    extracted from the C string literals at test/src/string_search/simd_searchTest.cpp:12-34 and
    test/src/string_search/search_wrappersTest.cpp:12-21
  * kat.json      : the expected values those tests assert (file:line noted per entry)
  * ref_vectors.json : outputs of the reference's own code (oracle/_ref/libxsref.so) on small seeded
    adversarial inputs -- inputs are regenerated from the seed by tests/gen_cases.py
  * regex_vectors.json (--regex): digests of what the reference's regex wrappers over its vendored re2
    (oracle/_ref/libxsref_re.so) return for the cases of tests/test_regex.py, per result kind
"""
import json
import os
import re
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
REF = "/root/reference"
GOLD = os.path.join(ROOT, "tests", "golden")


def c_literal(path, start_marker):
    src = open(path, encoding="utf-8").read()
    i = src.index(start_marker)
    j = src.index(";", i)
    body = src[i:j]
    out = bytearray()
    for lit in re.findall(r'"((?:[^"\\]|\\.)*)"', body):
        k = 0
        while k < len(lit):
            ch = lit[k]
            if ch == "\\":
                nxt = lit[k + 1]
                if nxt == "n":
                    out.append(10)
                    k += 2
                elif nxt in "01234567":
                    m = re.match(r"[0-7]{1,3}", lit[k + 1:])
                    out.append(int(m.group(0), 8))
                    k += 1 + len(m.group(0))
                else:
                    out.append(ord(nxt))
                    k += 2
            else:
                out += ch.encode()
                k += 1
    return bytes(out)


def main():
    os.makedirs(GOLD, exist_ok=True)
    simd = c_literal(os.path.join(REF, "test/src/string_search/simd_searchTest.cpp"), "static char dummy_text[1241]")
    assert len(simd) == 1240, len(simd)
    wrap = c_literal(os.path.join(REF, "test/src/string_search/search_wrappersTest.cpp"), "static char dummy_text[]")
    open(os.path.join(GOLD, "simd_dummy_text.bin"), "wb").write(simd)
    open(os.path.join(GOLD, "wrappers_dummy_text.bin"), "wb").write(wrap)
    long_pat = ("smooth-bellied chirognostic inkos BVM antigraphy pagne "
                "bicorne complementizer commorant ever-endingly sheikhly")
    soh = "\\u0001"  # the tests use the byte \1 as an end marker; kept escaped in JSON
    kat = {
        "source": "lfreist/x-search test/src/string_search/{simd_searchTest,search_wrappersTest}.cpp",
        "findNext": [  # simd_searchTest.cpp:63-73
            ["Liane", 0, 0], [long_pat, 0, 113], ["Helladic", 0, 346], ["ly\x01", 0, 1237],
            ["jkahgsf", 0, -1], ["ia", 0, 1], ["ia", 3, 455]],
        "findNextNewLine": [[0, 223], [224, 282], [283, 728], [729, -1]],  # :77-80
        "findAllPerLine": [["is", 2], ["th", 3], ["Van", 1], ["y\x01", 1], ["Liane", 1], ["Vansdf", 0]],  # :84-89
        "findAll": [["is", 8], ["th", 5], ["Van", 1], ["y\x01", 1], ["Liane", 1], ["Vansdf", 0]],  # :93-98
        "wrappers": {  # search_wrappersTest.cpp:26,39,52 and the line contents of :64-69 without '\n'
            "pattern": "ant", "byte_offsets_match": [2, 151, 197, 507], "byte_offsets_line": [0, 113, 176, 460],
            "count": 4,
            "lines": ["Liant reindorsing two-time zippering chromolithography rainbowweed",
                      "smooth-bellied chirognostic inkos BVM antigraphy pagne bicorne",
                      "complementizer commorant ever-endingly sheikhly",
                      "refrangible terebras autobiographal mid-breast ant"]},
    }
    del soh
    json.dump(kat, open(os.path.join(GOLD, "kat.json"), "w"), indent=1)

    import gen_cases
    import oracle_lib
    _, ref = oracle_lib.load()
    assert ref is not None, "build oracle/_ref first (make -C oracle)"
    vecs = []
    for case in gen_cases.golden_cases():
        buf, pat, bounds = gen_cases.materialize(case)
        st, ln = ref.lines(buf, pat, bounds)
        vecs.append({
            "case": case,
            "count": int(ref.count(buf, pat, bounds, False)),
            "count_lines": int(ref.count(buf, pat, bounds, True)),
            "match_byte_offsets": [int(x) for x in ref.match_byte_offsets(buf, pat, bounds)],
            "line_byte_offsets": [int(x) for x in ref.line_byte_offsets(buf, pat, bounds)],
            "line_indices": [int(x) for x in ref.line_indices(buf, pat, bounds)],
            "lines": [[int(a), int(b)] for a, b in zip(st, ln)],
        })
    json.dump({"generator": "tools/make_golden.py via oracle/_ref/libxsref.so (reference @ 86641bb)", "vectors": vecs},
              open(os.path.join(GOLD, "ref_vectors.json"), "w"))
    # ignore_case: the reference's toLower processor (src/utils/string_utils.cpp:11-33) over buffer and pattern,
    # then its unmodified matchers
    ivecs = []
    for case in gen_cases.icase_cases():
        buf, pat, bounds = gen_cases.materialize(case)
        lbuf, lpat = ref.to_lower(buf), ref.to_lower(np.frombuffer(pat, dtype=np.uint8)).tobytes()
        st, ln = ref.lines(lbuf, lpat, bounds)
        ivecs.append({
            "case": case,
            "count": int(ref.count(lbuf, lpat, bounds, False)),
            "count_lines": int(ref.count(lbuf, lpat, bounds, True)),
            "match_byte_offsets": [int(x) for x in ref.match_byte_offsets(lbuf, lpat, bounds)],
            "line_byte_offsets": [int(x) for x in ref.line_byte_offsets(lbuf, lpat, bounds)],
            "line_indices": [int(x) for x in ref.line_indices(lbuf, lpat, bounds)],
            "lines": [[int(a), int(b)] for a, b in zip(st, ln)],
        })
    json.dump({"generator": "tools/make_golden.py: xs::utils::str::simd::toLower + matchers of oracle/_ref/libxsref.so",
               "vectors": ivecs}, open(os.path.join(GOLD, "icase_vectors.json"), "w"))
    # chunk table of the reference's meta-file fixture (format: SURVEY.md 3.5)
    import struct
    raw = open(os.path.join(REF, "test/files/sample.meta"), "rb").read()
    pos, chunks = 4, []
    while pos < len(raw):
        oo, ao, osz, asz, nm = struct.unpack_from("<5Q", raw, pos)
        pos += 40
        first = struct.unpack_from("<2Q", raw, pos) if nm else (0, 0)
        last = struct.unpack_from("<2Q", raw, pos + 16 * (nm - 1)) if nm else (0, 0)
        pos += 16 * nm
        chunks.append(dict(original_offset=oo, actual_offset=ao, original_size=osz, actual_size=asz, n_map=nm,
                           first_mapping=list(first), last_mapping=list(last)))
    json.dump({"source": "test/files/sample.meta of the reference (3 060 484 bytes): chunk table and first/last line "
                         "mapping of each chunk", "compression_type": struct.unpack_from("<I", raw, 0)[0],
               "original_size": chunks[-1]["original_offset"] + chunks[-1]["original_size"], "chunks": chunks},
              open(os.path.join(GOLD, "sample_meta_chunks.json"), "w"), indent=1)
    print("golden: %d KAT groups, %d reference vectors, %d ignore-case vectors" % (len(kat) - 1, len(vecs), len(ivecs)))


def main_regex():
    import hashlib

    import oracle_lib
    import test_regex as T
    ref = oracle_lib.load_regex_ref()
    assert ref is not None, "build oracle/_ref/libxsref_re.so first (make -C oracle)"
    data = T.corpora()
    cases = {T.case_key(c, p, i): T.expected_from_ref(ref, data[c], p, i) for c, p, i in T.CASES}
    json.dump({"generator": "tools/make_golden.py --regex: xs::search::regex::{count, byte_offsets_match, byte_offsets_line} "
                            "(search_wrappers.h:63-88,209-271) over re2 @954656f, oracle/_ref/libxsref_re.so; per kind the number "
                            "of output rows of `xsgrep --tag KIND`, their sha256 and first bytes",
               "corpora": {k: hashlib.sha256(v.tobytes()).hexdigest() for k, v in data.items()}, "cases": cases},
              open(os.path.join(GOLD, "regex_vectors.json"), "w"), indent=1)
    print("golden: %d regex cases" % len(cases))


if __name__ == "__main__":
    if "--regex" in sys.argv:
        main_regex()
    else:
        main()

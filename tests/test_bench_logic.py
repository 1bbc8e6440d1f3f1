"""CPU checks of bench.py's own logic: the parity comparison must see every kind of difference, the reference arm
must run without the product library, and both arms must describe the workload identically."""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def _exp(vals, lens=None):
    a = np.array(vals, dtype=np.uint64)
    if lens is None:
        return (a, a.size)
    return (a, np.array(lens, dtype=np.uint64), a.size)


def test_compare_outputs_sees_every_difference():
    tags = ("count_lines", "line_indices", "line_byte_offsets", "lines")
    exp = {"count_lines": 3, "line_indices": _exp([1, 5, 9]), "line_byte_offsets": _exp([10, 50, 90]),
           "lines": _exp([10, 50, 90], [4, 4, 7])}
    u = lambda v: np.array(v, dtype=np.uint64)  # noqa: E731
    good = {"count_lines": 3, "line_indices": u([101, 105, 109]), "line_byte_offsets": u([1010, 1050, 1090]),
            "lines": (u([1010, 1050, 1090]), u([4, 4, 7]))}
    assert bench.compare_outputs(tags, good, exp, 1000, 100) == (0, 1 + 3 + 3 + 6)
    for tag, bad_value in (("count_lines", 4), ("line_indices", u([101, 105, 108])), ("line_byte_offsets", u([1010, 1050])),
                           ("lines", (u([1010, 1050, 1090]), u([4, 5, 7])))):
        got = dict(good)
        got[tag] = bad_value
        assert bench.compare_outputs(tags, got, exp, 1000, 100)[0] > 0, tag
    # a truncated reference vector (capacity too small) is a mismatch, not a pass
    short = dict(exp)
    short["line_indices"] = (exp["line_indices"][0][:2], 3)
    assert bench.compare_outputs(tags, good, short, 1000, 100)[0] > 0


def test_workload_specs_and_config_identity():
    class A:
        workload, pattern_len, gb = bench.DEFAULT_WORKLOAD, 8, None
    tags, pat, needle, rate, seed, nbytes, scaling = bench.workload_spec(A, 1)
    assert nbytes % bench.SYNTH_BLOCK == 0 and scaling == "weak" and pat == b"Sherlock"
    A.workload = "line_byte_offsets+lines/needle32/100GB-streamed"
    per8 = bench.workload_spec(A, 8)[5]
    assert bench.workload_spec(A, 8)[6] == "strong" and abs(per8 * 8 - 100 * bench.GB) < 8 * bench.SYNTH_BLOCK
    c1 = bench.config_of(A, b"x", ("count",), 4096, "strong", 2)
    assert c1 == bench.config_of(A, b"x", ("count",), 4096, "strong", 2) and c1["total_bytes"] == 8192
    assert set(bench.SWEEP_PATTERNS) == {1, 2, 3, 4, 8, 16, 32, 64}
    assert all(len(p) == m for m, p in bench.SWEEP_PATTERNS.items())


def test_reference_arm_runs_without_the_product_library(tmp_path):
    """bench.py --impl reference: one JSON line, impl = reference, and the process never maps libxsg.so."""
    code = ("import sys, runpy; sys.argv = ['bench.py', '--impl', 'reference', '--workload', 'count/Sherlock/100MB', "
            "'--steps', '2', '--warmup', '1']; runpy.run_path(%r, run_name='__main__'); "
            "maps = open('/proc/self/maps').read(); assert 'libxsg.so' not in maps, 'product library mapped'; "
            "assert 'libxscorpus.so' in maps" % os.path.join(ROOT, "bench.py"))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=str(tmp_path))
    assert out.returncode == 0, out.stderr[-500:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["gpu_launches"] == 0 and line["cpu_baseline"]["kind"] in ("reference", "port")
    assert line["config"]["workload"] == "count/Sherlock/100MB" and line["results"]["per_tag"]["count"] > 0
    assert line["e2e"]["h2d_bytes_per_step"] == 0

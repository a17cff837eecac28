"""The generated sorting networks (psana-ray_b200/csrc/sortnet.cuh) sort: re-derive them and check them here,
and check that the committed header is what the generator produces."""
import importlib.util
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gen():
    spec = importlib.util.spec_from_file_location("gen_sortnet", os.path.join(ROOT, "tools", "gen_sortnet.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_networks_sort_and_match_the_committed_header():
    g = _gen()
    header = open(os.path.join(ROOT, "psana-ray_b200", "csrc", "sortnet.cuh")).read()
    for n in (48, 64):
        ces = g.network(n)
        g.check(n, ces)                                   # random permutations + 0/1 vectors
        body = header[header.index(f"psr_sort{n}("):]
        body = body[:body.index("}\n")]
        got = [(int(a), int(b)) for a, b in re.findall(r"PSR_CE\((\d+),(\d+)\)", body)]
        assert got == ces
        assert all(0 <= i < j < n for i, j in ces)

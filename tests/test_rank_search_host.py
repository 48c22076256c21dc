"""The search-based rank bookkeeping of the Mann-Whitney kernel (foremast_b200/csrc/fm_rank_search.h)
is plain C++: compile it for the host and check it against scipy's rankdata on the pooled sample,
including heavy ties, all-equal samples, disjoint ranges and full tiles."""
import os
import subprocess

import numpy as np
import pytest
from scipy import stats

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("rs") / "rank_search_host")
    subprocess.run(["/usr/bin/g++", "-O1", "-std=c++17", "-o", exe,
                    os.path.join(ROOT, "tests", "host", "rank_search_host.cpp")], check=True)
    return exe


def fkey(x):
    """order-preserving u32 key of finite fp32 values (as the kernel's float_key)."""
    u = (x + np.float32(0)).view(np.uint32)
    return np.where(u & 0x80000000, ~u, u | 0x80000000).astype(np.uint32)


def tie_term(v):
    _, cnt = np.unique(v, return_counts=True)
    cnt = cnt.astype(np.int64)
    return int((cnt ** 3 - cnt).sum())


def test_search_statistics_match_rankdata(harness):
    rng = np.random.default_rng(11)
    cases, expect = [], []
    shapes = [(1, 1), (5, 3), (31, 33), (32, 32), (64, 64), (96, 96), (720, 720), (768, 768), (1000, 17), (17, 1000),
              (1024, 1024), (2048, 2048), (2048, 1)]
    for na, nb in shapes:
        for kind in range(7):
            if kind == 0:
                x, y = rng.normal(0, 1, na), rng.normal(0, 1, nb)
            elif kind == 1:   # heavy ties
                x, y = rng.integers(0, 3, na), rng.integers(0, 3, nb)
            elif kind == 2:   # all equal
                x, y = np.ones(na), np.ones(nb)
            elif kind == 3:   # disjoint, current above baseline
                x, y = rng.normal(0, 1, na), rng.normal(100, 1, nb)
            elif kind == 4:   # current entirely inside one gap of the baseline
                x = np.concatenate([rng.normal(-50, 1, na // 2), rng.normal(50, 1, na - na // 2)])
                y = rng.integers(0, 2, nb) * 0.5
            elif kind == 5:   # moderate ties
                x, y = np.round(rng.normal(0, 2, na)), np.round(rng.normal(0.5, 2, nb))
            else:             # ties only inside the current sample; negative values and zeros
                x, y = rng.normal(0, 1, na), np.round(rng.normal(0, 1, nb), 1) + 0.05
            x = np.sort(x.astype(np.float32))
            y = y.astype(np.float32)  # the current sample stays unsorted
            cases.append("%d %d %s %s" % (na, nb, " ".join(str(int(v)) for v in fkey(x)),
                                          " ".join(str(int(v)) for v in fkey(y))))
            pooled = np.concatenate([x, y]).astype(np.float64)
            r = stats.rankdata(pooled)
            r2 = int(round(2 * r[:na].sum()))
            expect.append((r2, r2 - na * (na + 1), tie_term(pooled)))
    inp = "%d\n%s\n" % (len(cases), "\n".join(cases))
    out = subprocess.run([harness], input=inp, capture_output=True, text=True, check=True).stdout.split("\n")
    got = [tuple(int(v) for v in line.split()) for line in out if line.strip()]
    assert got == expect

"""Regenerates tests/golden/*.json.  Run in the dev container only (needs /root/reference + scipy).

demo_fixture.json  -- the reference's only numeric fixture
                      (examples/spring-boot-demo/src/main/resources/data1.txt, data2.txt; used by
                      examples/demo/continuous/demo_v2.yaml:33 to provoke an Unhealthy verdict)
                      with scipy 1.18.1 cross-check values.  NOT reference-pinned scores: the
                      reference holds no known-answer test for its scorer.
rank_cases.json    -- seeded random baseline/current pairs (with ties, unequal lengths, shifts)
                      and scipy's statistics for them.
"""
import json
import os

import numpy as np
import scipy
from scipy import stats

REF = "/root/reference/examples/spring-boot-demo/src/main/resources"
HERE = os.path.dirname(os.path.abspath(__file__))


def read(name):
    return [float(l.split(",")[1]) for l in open(os.path.join(REF, name)) if l.strip()]


def rank_stats(x, y):
    x = np.asarray(x, np.float32)
    y = np.asarray(y, np.float32)
    n1, n2 = len(x), len(y)
    pooled = np.concatenate([x, y]).astype(np.float64)
    r = stats.rankdata(pooled)
    r1 = r[:n1].sum()
    _, cnt = np.unique(pooled, return_counts=True)
    ties = int((cnt.astype(np.int64) ** 3 - cnt).sum())
    out = {"n1": n1, "n2": n2, "mw_2r1": int(round(2 * r1)),
           "mw_2u1": int(round(2 * r1)) - n1 * (n1 + 1), "mw_ties": ties}
    mw = stats.mannwhitneyu(x.astype(np.float64), y.astype(np.float64), alternative="two-sided",
                            method="asymptotic", use_continuity=True)
    out["p_mw"] = float(mw.pvalue)
    kw = stats.kruskal(x.astype(np.float64), y.astype(np.float64))
    out["kw_h"] = float(kw.statistic)
    out["p_k"] = float(kw.pvalue)
    m = min(n1, n2)
    d = (y[:m] - x[:m]).astype(np.float32)   # fp32 differences, DECISIONS D8
    nz = d[d != 0]
    out["w_n"] = int(nz.size)
    if nz.size:
        rr = stats.rankdata(np.abs(nz).astype(np.float64))
        out["w_2wplus"] = int(round(2 * rr[nz > 0].sum()))
        out["w_2wminus"] = int(round(2 * rr[nz < 0].sum()))
        _, c2 = np.unique(np.abs(nz), return_counts=True)
        out["w_ties"] = int((c2.astype(np.int64) ** 3 - c2).sum())
        try:
            w = stats.wilcoxon(nz.astype(np.float64), zero_method="wilcox", correction=False,
                               method="approx")
            out["p_w"] = float(w.pvalue)
        except Exception:  # degenerate
            out["p_w"] = None
    return out


def main():
    d1, d2 = read("data1.txt"), read("data2.txt")
    fx = {"source": "examples/spring-boot-demo/src/main/resources/data1.txt,data2.txt",
          "scipy": scipy.__version__, "data1": d1, "data2": d2}
    fx["pooled"] = rank_stats(d1, d2)
    fx["paired36"] = rank_stats(d1, d2[:36])
    a = np.asarray(d1, np.float32).astype(np.float64)
    fx["ma"] = {"mean": float(a.mean()), "std": float(a.std()), "upper_k2": float(a.mean() + 2 * a.std()),
                "anomalies_idx": [i for i, v in enumerate(np.asarray(d2, np.float32)) if v > a.mean() + 2 * a.std()]}
    json.dump(fx, open(os.path.join(HERE, "demo_fixture.json"), "w"), indent=1)

    rng = np.random.default_rng(20261106)
    cases = []
    for i in range(40):
        n1 = int(rng.integers(5, 200))
        n2 = int(rng.integers(5, 200)) if i % 3 else n1
        kind = i % 4
        if kind == 0:    # continuous, same distribution
            x, y = rng.normal(10, 2, n1), rng.normal(10, 2, n2)
        elif kind == 1:  # heavy ties (quantised error counts)
            x, y = rng.poisson(2, n1) * 0.001, rng.poisson(2, n2) * 0.001
        elif kind == 2:  # shifted
            x, y = rng.normal(10, 2, n1), rng.normal(12, 2, n2)
        else:            # mixed sign, some exact equal pairs
            x = np.round(rng.normal(0, 1, n1), 1)
            y = np.round(rng.normal(0, 1, n2), 1)
        x = x.astype(np.float32)
        y = y.astype(np.float32)
        c = rank_stats(x, y)
        c["x"] = [float(v) for v in x]
        c["y"] = [float(v) for v in y]
        cases.append(c)
    # degenerate: all equal
    c = {"x": [1.5] * 30, "y": [1.5] * 25}
    cases.append({**c, "n1": 30, "n2": 25, "degenerate": True})
    json.dump({"scipy": scipy.__version__, "cases": cases}, open(os.path.join(HERE, "rank_cases.json"), "w"))


if __name__ == "__main__":
    main()

"""Rank kernels on a batch of small jobs with a few long ones mixed in (the sort size of a launch is chosen from the
longest segment it has to rank).  usage: python scripts/mixed_sizes_time.py [units]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foremast_b200 import _abi as abi  # noqa: E402
from foremast_b200.api import Context, default_params  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 400_000
rng = np.random.default_rng(9)


def make(nb, nc):
    tot = nb + nc
    offs = np.zeros(n + 1, np.int64)
    np.cumsum(tot + (-tot) % 4, out=offs[1:])
    vals = rng.normal(100.0, 10.0, int(offs[-1])).astype(np.float32)
    seg = np.stack([np.zeros(n, np.uint32), nb.astype(np.uint32), nc.astype(np.uint32)], axis=1)
    return abi.HostCsr(vals, offs, seg, np.arange(n) // 5, rng.integers(0, 8, n))


with Context(0) as ctx:
    p = default_params()
    p.algorithms = abi.ALG_MA_ALL
    p.pairwise = abi.PW_ALL
    small_b, small_c = np.full(n, 60), np.full(n, 30)
    cases = {"all 60 + 30": (small_b, small_c)}
    for k, name in ((1, "one unit of 720 + 720 among them"), (n // 100, "1 % of 720 + 720 among them"),
                    (n // 10, "10 % of 720 + 720, 10 % of 200 + 200 among them")):
        nb, nc = small_b.copy(), small_c.copy()
        idx = rng.choice(n, k, replace=False)
        nb[idx], nc[idx] = 720, 720
        if k == n // 10:
            idx2 = rng.choice(n, k, replace=False)
            nb[idx2], nc[idx2] = 200, 200
        cases[name] = (nb, nc)
    for name, (nb, nc) in cases.items():
        b = ctx.batch_from_csr(make(nb, nc))
        for _ in range(3):
            r = ctx.score(b, p)
            r.sync()
            t = r.timings_ms()
            r.free()
        print(f"{name:52s} samples {float((nb + nc).sum()) / 1e6:6.1f} M  wilcoxon {t['rank_wilcoxon']:.3f}  mw {t['rank_mw']:.3f}  band {t['band']:.3f} ms")
        b.free()

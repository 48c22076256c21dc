"""Rank and band kernels on tie-heavy data (integer-valued metrics) against continuous data of the same shape.
usage: python scripts/ties_time.py [units]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foremast_b200 import _abi as abi  # noqa: E402
from foremast_b200.api import Context, default_params  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
rng = np.random.default_rng(3)
L = 720
offs = np.arange(n + 1, dtype=np.int64) * (2 * L)
seg = np.tile(np.array([0, L, L], np.uint32), (n, 1))
with Context(0) as ctx:
    p = default_params()
    p.algorithms = abi.ALG_MA_ALL | abi.ALG_MA_ROLLING | abi.ALG_EWMA | abi.ALG_DES
    p.pairwise = abi.PW_ALL
    base = rng.normal(100.0, 3.0, n * 2 * L).astype(np.float32)
    cases = {"continuous": base,
             "rounded to integers (~20 distinct values)": np.round(base),
             "rounded to 0.1 (~200 distinct values)": np.round(base * 10) / 10,
             "counts 0/1/2 (Poisson 0.7)": rng.poisson(0.7, n * 2 * L).astype(np.float32),
             "constant": np.full(n * 2 * L, 5.0, np.float32)}
    for name, vals in cases.items():
        b = ctx.batch_from_csr(abi.HostCsr(vals.astype(np.float32), offs, seg, np.arange(n) // 80, rng.integers(0, 8, n)))
        for _ in range(3):
            r = ctx.score(b, p)
            r.sync()
            t = r.timings_ms()
            r.free()
        print(f"{name:44s} wilcoxon {t['rank_wilcoxon']:.3f}  mw {t['rank_mw']:.3f}  band {t['band']:.3f} ms")
        b.free()

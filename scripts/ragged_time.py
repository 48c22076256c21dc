"""How the scoring kernels take a batch whose units all have different lengths (jobs with different windows batched
together), against a uniform batch of the same total size.  usage: python scripts/ragged_time.py [units]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foremast_b200 import _abi as abi  # noqa: E402
from foremast_b200.api import Context, default_params  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
rng = np.random.default_rng(5)


def make(nb, nc):
    tot = nb + nc
    pad = (-tot) % 4
    offs = np.zeros(n + 1, np.int64)
    np.cumsum(tot + pad, out=offs[1:])
    vals = rng.normal(100.0, 10.0, int(offs[-1])).astype(np.float32)
    seg = np.stack([np.zeros(n, np.uint32), nb.astype(np.uint32), nc.astype(np.uint32)], axis=1)
    return abi.HostCsr(vals, offs, seg, np.arange(n) // 80, rng.integers(0, 8, n))


with Context(0) as ctx:
    p = default_params()
    p.algorithms = abi.ALG_MA_ALL | abi.ALG_MA_ROLLING | abi.ALG_EWMA | abi.ALG_DES
    p.pairwise = abi.PW_ALL
    cases = {"uniform 560 + 560": (np.full(n, 560), np.full(n, 560)),
             "base, cur ~ U[400, 720] independently": (rng.integers(400, 721, n), rng.integers(400, 721, n)),
             "base = cur ~ U[400, 720]": ((lambda x: (x, x))(rng.integers(400, 721, n))),
             "sorted by length (base = cur ~ U[400, 720])": ((lambda x: (x, x))(np.sort(rng.integers(400, 721, n))))}
    for name, (nb, nc) in cases.items():
        b = ctx.batch_from_csr(make(nb, nc))
        for _ in range(3):
            r = ctx.score(b, p)
            r.sync()
            t = r.timings_ms()
            r.free()
        tot = float((nb + nc).sum())
        print(f"{name:48s} samples {tot / 1e6:7.1f} M  wilcoxon {t['rank_wilcoxon']:.3f}  mw {t['rank_mw']:.3f}  band {t['band']:.3f} ms"
              f"  -> {tot / 1e6 / (t['rank'] + t['band']):.1f} M samples / ms")
        b.free()

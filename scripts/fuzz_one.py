"""One seed of tests/test_gpu_fuzz.py with the differences printed field by field (debugging aid).
usage: python scripts/fuzz_one.py <seed>"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foremast_b200 import _abi as abi  # noqa: E402
from foremast_b200.api import Context  # noqa: E402
from oracle import oracle  # noqa: E402
from tests import test_gpu_fuzz as f  # noqa: E402

seed = int(sys.argv[1])
rng = np.random.default_rng(1000 + seed)
max_seg = f.MAX_SEGS[seed % 7]
n = f.N_UNITS[seed % 7]
units, mts = f.make_units(rng, n, max_seg)
sid = np.sort(rng.integers(0, max(1, n // 7), n))
p = oracle.params_default()
p.algorithms = int(rng.integers(0, 32))
p.pairwise = int(rng.integers(0, 6))
p.pairwise_threshold = float(rng.choice([0.05, 0.01, 0.5]))
p.differ_scale = float(rng.choice([0.5, 0.25, 1.0]))
p.min_mann_white, p.min_wilcoxon, p.min_kruskal = (int(rng.choice([1, 5, 20])) for _ in range(3))
p.min_historical = int(rng.choice([1, 10, 50]))
p.min_anomaly_points = int(rng.choice([1, 1, 3]))
p.ma_window = int(rng.choice([1, 5, 30, 33, 200, 5000]))
p.period = int(rng.choice([32, 48, 100, 1440]))
p.alpha, p.beta, p.gamma = float(rng.uniform(0.05, 0.9)), float(rng.uniform(0.01, 0.5)), float(rng.uniform(0.01, 0.5))
p.rules[abi.MT_TPS].threshold, p.rules[abi.MT_TPS].bound, p.rules[abi.MT_TPS].min_lower_bound = 1.5, abi.BOUND_LOWER, 0.1
p.rules[abi.MT_GC].threshold, p.rules[abi.MT_GC].bound, p.rules[abi.MT_GC].min_lower_bound = 3.0, abi.BOUND_BOTH, 1.0
print("algs", p.algorithms, "pairwise", p.pairwise, "period", p.period, "abg", p.alpha, p.beta, p.gamma, "min_hist", p.min_historical)
with Context(0) as ctx:
    b = ctx.batch_from_units(units, service_id=sid, metric_type=mts)
    host = b.download()
    ref, _, _ = oracle.score(host, p)
    got = ctx.score(b, p).series()
    for name in got.dtype.names:
        a, r = got[name], ref[name]
        if a.dtype.kind == "f":
            bad = ~(np.isclose(a, r, rtol=1e-5, atol=0) | (a == r))
        else:
            bad = a != r
        if bad.any():
            idx = np.argwhere(bad)[:6]
            print(name, "differs at", idx.tolist())
            for i in idx:
                u = int(i[0])
                print("  unit", u, "lens", host.seg_lens[u].tolist(), "mt", int(host.metric_type[u]), "got", a[tuple(i)], "ref", r[tuple(i)])
    print("done")
if len(sys.argv) > 2:
    u = int(sys.argv[2])
    h, bs, c = host.segments(u)
    print("unit", u, "hist", h[:8], "... min/max", h.min() if h.size else None, h.max() if h.size else None, "cur", c[:24])
    print("got", {k: got[k][u].tolist() for k in ("n_anom", "first_idx", "sigma", "status")})
    print("ref", {k: ref[k][u].tolist() for k in ("n_anom", "first_idx", "sigma", "status")})

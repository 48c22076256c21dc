"""cfg0 (BASELINE.json configs[0]): one canary job -- 5 series, 60-point baseline vs 30-point current, MA band +
Mann-Whitney U -- through the one-shot host call fm_score_host.  Prints microseconds per call."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foremast_b200 import _abi as abi  # noqa: E402
from foremast_b200.api import Context, default_params  # noqa: E402

rng = np.random.default_rng(0)
units0 = [([], rng.normal(100, 5, 60).astype(np.float32), rng.normal(100, 5, 30).astype(np.float32)) for _ in range(5)]
h0 = abi.HostCsr.from_units(units0, service_id=[0] * 5, metric_type=[1, 3, 6, 4, 5])
with Context(0) as ctx:
    q = default_params()
    q.algorithms = abi.ALG_MA_ALL
    q.pairwise = abi.PW_MANN_WHITE
    for _ in range(50):
        ctx.score_host(h0, q, 1)
    t = time.perf_counter()
    for _ in range(2000):
        v = ctx.score_host(h0, q, 1)
    print("cfg0 us/call", (time.perf_counter() - t) / 2000 * 1e6, "verdict", int(v[0]["verdict"]))

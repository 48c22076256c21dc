"""Phases of the Prometheus-bodies -> sealed batch path at cfg2's shape (see bench.py other_configs)."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foremast_b200 import _abi as abi  # noqa: E402
from foremast_b200.api import Context  # noqa: E402


def prom_body(n_series, n_points, seed):
    r = np.random.default_rng(seed)
    return json.dumps({"status": "success", "data": {"resultType": "matrix", "result": [
        {"metric": {"__name__": "namespace_app_pod_cpu_usage_seconds_total", "pod": f"pod-{i}", "namespace": "dev"},
         "values": [[1550000000 + 60 * k, f"{v:.4f}"] for k, v in enumerate(r.normal(0.45, 0.08, n_points))]}
        for i in range(n_series)]}}).encode()


hb, cb = prom_body(1, 1380, 1), prom_body(8, 60, 2)
n_jobs = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000
threads = int(sys.argv[2]) if len(sys.argv) > 2 else 0
with Context(0) as ctx:
    t0 = time.perf_counter()
    arr, n, keep = ctx.prom_jobs([(j // 10, abi.MT_CPU, 60.0, hb, None, cb) for j in range(n_jobs)])
    t1 = time.perf_counter()
    mb = n_jobs * (len(hb) + len(cb)) / 1e6
    for rep in range(3):
        h = C.c_void_p()
        ctx._check(ctx._L.fm_batch_begin(ctx._h, n, 0, C.byref(h)))
        per = np.zeros(n, np.uint32)
        t2 = time.perf_counter()
        ctx._check(ctx._L.fm_batch_add_prometheus_many(h, arr, n, abi.PROM_GRID | abi.PROM_FILL, threads, per.ctypes.data))
        t3 = time.perf_counter()
        ctx._check(ctx._L.fm_batch_seal(h))
        t4 = time.perf_counter()
        ctx._L.fm_batch_free(h)
        print(f"{mb:.0f} MB JSON, {int(per.sum())} units: job array {1e3 * (t1 - t0):.0f} ms (python, once), "
              f"add_many {1e3 * (t3 - t2):.0f} ms = {mb / (t3 - t2):.0f} MB/s, seal {1e3 * (t4 - t3):.0f} ms")

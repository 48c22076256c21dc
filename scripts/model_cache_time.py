"""Continuous mode (fm_model_fit / fm_model_score): fit Holt-Winters on the history once, then score successive
windows of new samples from the cached state; against re-scoring history + window every cycle.
usage: python scripts/model_cache_time.py [units] [hist] [cur] [period]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foremast_b200 import _abi as abi  # noqa: E402
from foremast_b200.api import Context, default_params  # noqa: E402

units = int(sys.argv[1]) if len(sys.argv) > 1 else 500_000
hist = int(sys.argv[2]) if len(sys.argv) > 2 else 10_080
cur = int(sys.argv[3]) if len(sys.argv) > 3 else 60
period = int(sys.argv[4]) if len(sys.argv) > 4 else 1440
with Context(0) as ctx:
    p = default_params()
    p.algorithms, p.pairwise, p.period = abi.ALG_HW, abi.PW_NONE, period
    full = ctx.batch_synth(units, hist, 0, cur, flags=abi.SYNTH_ANOMALIES)
    hb = ctx.batch_synth(units, hist, 0, 0, flags=abi.SYNTH_ANOMALIES)
    cb = ctx.batch_synth(units, 0, 0, cur, flags=abi.SYNTH_ANOMALIES)
    for _ in range(2):
        r = ctx.score(full, p); r.verdicts(); t_full = r.timings_ms()["band"]; r.free()
    t0 = time.perf_counter()
    model = ctx.model_fit(hb, p)
    r = model.score(cb); r.verdicts(); r.free()
    t_first = (time.perf_counter() - t0) * 1e3
    ts = []
    for _ in range(5):
        t0 = time.perf_counter()
        r = model.score(cb); r.verdicts()
        ts.append((time.perf_counter() - t0) * 1e3)
        t_band = r.timings_ms()["band"]
        r.free()
    print(f"{units} units, {hist} + {cur} samples, period {period}: full re-score {t_full:.2f} ms (kernel); "
          f"fit + first window {t_first:.2f} ms (wall); later windows {min(ts):.2f} ms (wall), {t_band:.2f} ms (kernel); "
          f"state {units * (period + 4) * 8 / 1e9:.2f} GB")

"""Times k_band on the bench shapes: fused MA+EWMA+DES and single algorithms on 1 M x (720+720),
and cfg2 (80 k x 1380+60, EWMA+DES).  usage: python scripts/band_configs_time.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foremast_b200 import _abi as abi  # noqa: E402
from foremast_b200.api import Context, default_params  # noqa: E402


def timed(ctx, b, algs, n=6):
    p = default_params()
    p.algorithms, p.pairwise = algs, abi.PW_NONE
    ts = []
    for i in range(n + 2):
        r = ctx.score(b, p)
        t = r.timings_ms()["band"]
        r.free()
        if i >= 2:
            ts.append(t)
    return sum(ts) / len(ts)


with Context(0) as ctx:
    b = ctx.batch_synth(1_000_000, 0, 720, 720, flags=3)
    for name, algs in (("fused", abi.ALG_MA_ALL | abi.ALG_EWMA | abi.ALG_DES), ("ma_all", abi.ALG_MA_ALL),
                       ("ewma", abi.ALG_EWMA), ("des", abi.ALG_DES), ("rolling", abi.ALG_MA_ROLLING)):
        print(f"1M x 1440 {name:8s} {timed(ctx, b, algs):.3f} ms")
    b.free()
    b = ctx.batch_synth(80_000, 1380, 0, 60, units_per_service=8, flags=3)
    print(f"cfg2 80k ewma+des   {timed(ctx, b, abi.ALG_EWMA | abi.ALG_DES, 10):.4f} ms")
    b.free()

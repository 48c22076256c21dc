"""Small driver for ncu: one synthetic batch, a few scoring passes, nothing else.
usage: python scripts/profile_step.py [units] [passes] [n_hist n_base n_cur] [algs] [pairwise]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foremast_b200 import _abi as abi  # noqa: E402
from foremast_b200.api import Context, default_params  # noqa: E402

units = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
passes = int(sys.argv[2]) if len(sys.argv) > 2 else 2
shape = tuple(int(x) for x in sys.argv[3:6]) if len(sys.argv) > 5 else (0, 720, 720)
algs = int(sys.argv[6]) if len(sys.argv) > 6 else abi.ALG_MA_ALL | abi.ALG_EWMA | abi.ALG_DES
pairwise = int(sys.argv[7]) if len(sys.argv) > 7 else abi.PW_ALL
with Context(0) as ctx:
    b = ctx.batch_synth(units, *shape, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES)
    p = default_params()
    if os.environ.get("FM_BENCH_RULES", "1") != "0":  # the rule table of bench.py (rows for tps / gc / threads)
        for mt in (abi.MT_TPS, abi.MT_GC, abi.MT_THREADS):
            p.rules[mt].threshold, p.rules[mt].min_lower_bound, p.rules[mt].bound = 5.0, 0.0, abi.BOUND_UPPER
    p.algorithms = algs
    p.pairwise = pairwise
    if os.environ.get("FM_PERIOD"):
        p.period = int(os.environ["FM_PERIOD"])
    for _ in range(passes):
        r = ctx.score(b, p)
        v = r.verdicts()
        t = r.timings_ms()
        r.free()
    print(units, shape, {k: round(x, 3) for k, x in t.items()},
          "unhealth", int((v["verdict"] == abi.V_UNHEALTH).sum()), "of", v.size)

import sys; sys.path.insert(0,'/root/repo')
from foremast_b200 import _abi as abi
from foremast_b200.api import Context, default_params
with Context(0) as ctx:
    b = ctx.batch_synth(1000000, 0, 720, 720, flags=3)
    for alg in (abi.ALG_EWMA, abi.ALG_MA_ALL):
        p = default_params(); p.algorithms = alg; p.pairwise = 0
        for band in (0, alg):
            p.band_algorithm = band
            for _ in range(3):
                r = ctx.score(b, p); t = r.timings_ms(); r.free()
            print(alg, "bands" if band else "no bands", round(t["band"], 3), "ms")

"""Small all-kernels run for compute-sanitizer (one tool per gpurun call, see B200_PROFILING.md):
every kernel family once on ragged batches: rank (R = 1..64), band with all four algorithms,
Holt-Winters, bands materialised, anomaly list (scan + emit), verdict fold, fm_score_host chunks."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("FOREMAST_GPU_CHUNK_BYTES", "65536")
from foremast_b200 import _abi as abi  # noqa: E402
from foremast_b200.api import Context, default_params  # noqa: E402

with Context(0) as ctx:
    for shape, n in (((0, 20, 20), 300), ((200, 60, 45), 300), ((150, 300, 280), 200), ((0, 720, 720), 150),
                     ((0, 1500, 1100), 40), ((400, 0, 90), 120)):
        b = ctx.batch_synth(n, *shape, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES, units_per_service=7)
        p = default_params()
        p.algorithms = abi.ALG_ALL
        p.pairwise = abi.PW_ALL
        p.period = 96
        p.ma_window = 50
        for band in (0, abi.ALG_EWMA, abi.ALG_HW):
            p.band_algorithm = band
            r = ctx.score(b, p)
            r.verdicts()
            if band:
                r.bands(3)
            for alg in (abi.ALG_MA_ROLLING, abi.ALG_HW):
                r.anomalies(alg)
            r.free()
        p.band_algorithm = 0
        ctx.score_host(b.download(), p)
        b.free()
print("sanitize case done")

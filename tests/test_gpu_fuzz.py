"""Randomised differential test: ragged hand-built batches with awkward segment lengths (0, 1, around
the 32-sample staging chunk, around the sort capacities of the rank kernels), random parameters, heavy
ties -- CUDA path (through the C ABI) against the oracle, same bar as test_gpu_parity.py."""
import numpy as np
import pytest

from foremast_b200 import _abi as abi
from tests.test_gpu_parity import assert_scores_equal

pytestmark = pytest.mark.gpu

import os

# segment lengths around the staging chunk (32) and around every sort size of the rank kernels
# (32*R slots, R = 2^q or 3*2^q)
LENS = [0, 1, 2, 5, 9, 10, 11, 19, 20, 21, 31, 32, 33, 63, 64, 65, 95, 96, 97, 100, 127, 128, 129, 191, 192, 193,
        255, 256, 257, 383, 384, 385, 500, 511, 512, 513, 720, 767, 768, 769, 1000, 1023, 1024, 1025, 1440, 1535,
        1536, 1537, 2047, 2048]
MAX_SEGS = [64, 300, 1024, 2048, 192, 768, 1536]
N_UNITS = [400, 300, 200, 120, 300, 200, 150]
SEEDS = int(os.environ.get("FOREMAST_FUZZ_SEEDS", "140"))  # 420 were run clean while developing


def make_units(rng, n, max_seg):
    units, mts = [], []
    lens = [x for x in LENS if x <= max_seg]
    for _ in range(n):
        nh = int(rng.choice(LENS + [3000, 4500]))
        nb, nc = int(rng.choice(lens)), int(rng.choice(lens))
        kind = rng.integers(0, 5)
        tot = nh + nb + nc
        t = np.arange(tot)
        if kind == 0:
            z = rng.normal(100, 10, tot)
        elif kind == 1:   # quantised counters: many ties, many exact zeros
            z = np.maximum(rng.poisson(0.7, tot) - 0.0, 0) * 0.25
        elif kind == 2:   # constant
            z = np.full(tot, float(rng.integers(0, 3)))
        elif kind == 3:   # seasonal + trend + shift in the current window
            z = 50 + 10 * np.sin(2 * np.pi * t / 48) + 0.01 * t + rng.normal(0, 1, tot)
            z[nh + nb:] += rng.choice([0, 0, 8])
        else:             # huge dynamic range
            z = np.abs(rng.normal(0, 1, tot)) * 10.0 ** rng.integers(-3, 9)
        if tot and rng.random() < 0.3:
            z[rng.integers(0, tot)] *= 30
        z = z.astype(np.float32)
        units.append((z[:nh], z[nh:nh + nb], z[nh + nb:]))
        mts.append(int(rng.integers(0, 10)))
    return units, mts


@pytest.mark.parametrize("seed", range(SEEDS))
def test_random_batches_against_oracle(seed):
    from foremast_b200.api import Context
    from oracle import oracle
    rng = np.random.default_rng(1000 + seed)
    max_seg = MAX_SEGS[seed % 7]
    n = N_UNITS[seed % 7]
    units, mts = make_units(rng, n, max_seg)
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
    with Context(0) as ctx:
        b = ctx.batch_from_units(units, service_id=sid, metric_type=mts)
        host = b.download()
        ref, _, _ = oracle.score(host, p)
        res = ctx.score(b, p)
        got = res.series()
        assert_scores_equal(got, ref)
        v = oracle.verdicts(host, ref, b.info()["n_services"])
        assert res.verdicts().tobytes() == v.tobytes()
        res.free()
        b.free()

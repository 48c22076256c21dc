"""Size-independent properties of the scoring spec, checked on the oracle (CPU) -- the same
properties are asserted on the CUDA path at full size in test_gpu_parity.py::test_full_size_properties."""
import numpy as np
import pytest

from foremast_b200 import _abi as abi
from oracle import oracle


def score(units, p, mt=abi.MT_LATENCY):
    csr = abi.HostCsr.from_units(units, metric_type=[mt] * len(units))
    return oracle.score(csr, p)[0]


@pytest.fixture(scope="module")
def pairs():
    rng = np.random.default_rng(42)
    out = []
    for i in range(60):
        n1, n2 = int(rng.integers(20, 300)), int(rng.integers(20, 300))
        q = [None, 0.5, 0.01][i % 3]
        x, y = rng.normal(5, 2, n1), rng.normal(5.3, 2, n2)
        if q:
            x, y = np.round(x / q) * q, np.round(y / q) * q
        out.append((x.astype(np.float32), y.astype(np.float32)))
    return out


def test_rank_statistics_invariant_under_exact_monotone_maps(pairs):
    p = oracle.params_default()
    a = score([([], x, y) for x, y in pairs], p)
    # x -> 4x and x -> -x are exact in fp32: same order (resp. reversed), same ties
    b = score([([], 4 * x, 4 * y) for x, y in pairs], p)
    for f in ("mw_2u1", "mw_2r1", "mw_ties", "w_2wplus", "w_2wminus", "w_ties", "w_n"):
        assert np.array_equal(a[f], b[f]), f
    assert np.array_equal(a["p_mw"], b["p_mw"]) and np.array_equal(a["p_w"], b["p_w"])
    c = score([([], -x, -y) for x, y in pairs], p)
    n1 = np.array([len(x) for x, _ in pairs], np.int64)
    n2 = np.array([len(y) for _, y in pairs], np.int64)
    assert np.array_equal(c["mw_2u1"], 2 * n1 * n2 - a["mw_2u1"])          # U1 + U2 = n1*n2
    assert np.array_equal(c["mw_ties"], a["mw_ties"])
    assert np.array_equal(c["w_2wplus"], a["w_2wminus"]) and np.array_equal(c["w_2wminus"], a["w_2wplus"])
    assert np.array_equal(c["p_mw"], a["p_mw"]) and np.array_equal(c["p_w"], a["p_w"])
    np.testing.assert_allclose(c["kw_h"], a["kw_h"], rtol=1e-9, atol=1e-12)


def test_swapping_samples_mirrors_u(pairs):
    p = oracle.params_default()
    a = score([([], x, y) for x, y in pairs], p)
    b = score([([], y, x) for x, y in pairs], p)
    n1 = np.array([len(x) for x, _ in pairs], np.int64)
    n2 = np.array([len(y) for _, y in pairs], np.int64)
    assert np.array_equal(a["mw_2u1"] + b["mw_2u1"], 2 * n1 * n2)
    assert np.array_equal(a["mw_2r1"] + b["mw_2r1"], (n1 + n2) * (n1 + n2 + 1))   # R1 + R2 = N(N+1)/2
    assert np.array_equal(a["mw_ties"], b["mw_ties"])
    assert np.array_equal(a["p_mw"], b["p_mw"]) and np.array_equal(a["p_k"], b["p_k"])
    # two-group Kruskal-Wallis is Mann-Whitney without the continuity correction: p_k <= p_mw always
    assert np.all(a["p_k"] <= a["p_mw"] + 1e-15)


def test_bands_scale_with_the_data():
    rng = np.random.default_rng(7)
    units = []
    for _ in range(20):
        z = np.maximum(rng.normal(50, 5, 260), 0).astype(np.float32)
        z[230] += 60
        units.append((z[:200], [], z[200:]))
    p = oracle.params_default()
    p.algorithms = abi.ALG_MA_ALL | abi.ALG_MA_ROLLING | abi.ALG_EWMA | abi.ALG_DES
    p.pairwise = abi.PW_NONE
    a = score(units, p, abi.MT_CPU)
    b = score([(2 * h, [], 2 * c) for h, _, c in units], p, abi.MT_CPU)   # doubling is exact in fp32 and fp64
    assert np.array_equal(a["n_anom"], b["n_anom"]) and np.array_equal(a["first_idx"], b["first_idx"])
    assert np.array_equal(2 * a["sigma"], b["sigma"]) and np.array_equal(2 * a["ma_mean"], b["ma_mean"])
    assert (a["n_anom"][:, [0, 1, 2, 3]] > 0).all()


def test_verdict_fold_is_order_independent():
    csr = oracle.synth(800, 60, 30, 30, units_per_service=16, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES)
    p = oracle.params_default()
    sc = oracle.score(csr, p)[0]
    v = oracle.verdicts(csr, sc, csr.n_services)
    assert int(v["n_units"].sum()) == 800
    assert np.array_equal(v["verdict"] == abi.V_UNHEALTH, v["n_anom_units"] > 0)
    perm = np.random.default_rng(1).permutation(800)
    shuffled = abi.HostCsr(csr.values, csr.offsets, csr.seg_lens, csr.service_id[perm], csr.metric_type)
    # same scores attributed to permuted services: totals per verdict class are preserved
    v2 = oracle.verdicts(shuffled, sc, csr.n_services)
    assert int(v2["n_anom_units"].sum()) == int(v["n_anom_units"].sum())
    assert int(v2["n_anom_points"].sum()) == int(v["n_anom_points"].sum())

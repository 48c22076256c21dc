"""CPU tests of the oracle (oracle/fm_oracle.c) against the golden fixtures and against
independent numpy/pandas/scipy restatements.  The oracle is the checker for the CUDA path, so it
is pinned first (parity itself stays "unpinned": the reference holds no scorer tests)."""
import json
import os

import numpy as np
import pytest

from foremast_b200 import _abi as abi
from oracle import oracle

REL = 1e-9


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


@pytest.fixture(scope="module")
def demo(golden_dir):
    return json.load(open(os.path.join(golden_dir, "demo_fixture.json")))


def test_demo_fixture_rank_statistics(demo):
    """data1.txt vs data2.txt of the reference's spring-boot demo: exact rank statistics and the
    scipy p-values."""
    p = oracle.params_default()
    csr = abi.HostCsr.from_units([([], demo["data1"], demo["data2"])], metric_type=[abi.MT_ERROR5XX])
    s = oracle.score(csr, p)[0][0]
    g = demo["pooled"]
    assert int(s["mw_2r1"]) == g["mw_2r1"] == 2946
    assert int(s["mw_2u1"]) == g["mw_2u1"] == 1614
    assert int(s["mw_ties"]) == g["mw_ties"] == 76068
    assert rel(s["p_mw"], g["p_mw"]) < REL
    assert rel(s["kw_h"], g["kw_h"]) < REL
    assert rel(s["p_k"], g["p_k"]) < REL
    # Wilcoxon on the first 36 pairs
    assert int(s["w_n"]) == g["w_n"] and int(s["w_2wplus"]) == g["w_2wplus"] == 185
    assert int(s["w_2wminus"]) == g["w_2wminus"] and int(s["w_ties"]) == g["w_ties"]
    assert rel(s["p_w"], g["p_w"]) < REL
    # all three say "same distribution", yet the spikes make the unit anomalous (design.md:35)
    assert not s["status"] & abi.S_DIFFERS
    assert s["status"] & abi.S_ANOMALY and s["status"] & abi.S_MODEL_BASE
    assert int(s["n_anom"][0]) == 2 and int(s["first_idx"][0]) == 1
    assert rel(s["ma_mean"], demo["ma"]["mean"]) < REL and rel(s["ma_std"], demo["ma"]["std"]) < 1e-7


def test_rank_cases_against_scipy(golden_dir):
    cases = json.load(open(os.path.join(golden_dir, "rank_cases.json")))["cases"]
    p = oracle.params_default()
    p.min_mann_white = p.min_wilcoxon = p.min_kruskal = 1
    csr = abi.HostCsr.from_units([([], c["x"], c["y"]) for c in cases])
    sc = oracle.score(csr, p)[0]
    for c, s in zip(cases, sc):
        n1, n2 = c["n1"], c["n2"]
        if c.get("degenerate"):
            assert s["p_mw"] == 1.0 and s["p_k"] == 1.0 and s["p_w"] == 1.0 and s["w_n"] == 0
            assert int(s["mw_ties"]) == (n1 + n2) ** 3 - (n1 + n2)
            continue
        for k in ("mw_2r1", "mw_2u1", "mw_ties", "w_n"):
            assert int(s[k]) == c[k], k
        # identities: U1 + U2 = n1*n2, R1 + R2 = N(N+1)/2
        assert 0 <= int(s["mw_2u1"]) <= 2 * n1 * n2
        assert rel(s["p_mw"], c["p_mw"]) < 1e-9
        assert rel(s["kw_h"], c["kw_h"]) < 1e-9 or abs(s["kw_h"] - c["kw_h"]) < 1e-12
        assert rel(s["p_k"], c["p_k"]) < 1e-9
        if c["w_n"]:
            for k in ("w_2wplus", "w_2wminus", "w_ties"):
                assert int(s[k]) == c[k], k
            assert int(s["w_2wplus"]) + int(s["w_2wminus"]) == c["w_n"] * (c["w_n"] + 1)
            if c["p_w"] is not None and np.isfinite(c["p_w"]):
                assert rel(s["p_w"], c["p_w"]) < 1e-9


def _ref_bands(src, cur, alg, p, k, bound, mlb, period):
    """Independent float64 numpy restatement of DECISIONS.md (plain loops, no fma)."""
    src = np.asarray(src, np.float64)
    cur = np.asarray(cur, np.float64)
    ups, los = [], []
    if alg == abi.ALG_MA_ALL:
        mu, sd = src.mean(), src.std()
        ups, los = [mu + k * sd] * len(cur), [max(mlb, mu - k * sd)] * len(cur)
        sigma = sd
    elif alg == abi.ALG_MA_ROLLING:
        z = np.concatenate([src, cur])
        for j in range(len(cur)):
            q = len(src) + j
            w = z[max(0, q - p.ma_window):q]
            ups.append(w.mean() + k * w.std())
            los.append(max(mlb, w.mean() - k * w.std()))
        sigma = w.std()
    elif alg == abi.ALG_EWMA:
        s, sse = src[0], 0.0
        for x in src[1:]:
            e = x - s
            sse += e * e
            s += p.alpha * e
        sigma = np.sqrt(sse / (len(src) - 1))
        for x in cur:
            ups.append(s + k * sigma)
            los.append(max(mlb, s - k * sigma))
            s += p.alpha * (x - s)
    elif alg == abi.ALG_DES:
        l, b, sse = src[0], src[1] - src[0], 0.0
        for x in src[1:]:
            l0 = l
            f = l + b
            sse += (x - f) ** 2
            l = p.alpha * x + (1 - p.alpha) * f            # textbook form, not error-correction
            b = p.beta * (l - l0) + (1 - p.beta) * b
        sigma = np.sqrt(sse / (len(src) - 1))
        for x in cur:
            l0 = l
            f = l + b
            ups.append(f + k * sigma)
            los.append(max(mlb, f - k * sigma))
            l = p.alpha * x + (1 - p.alpha) * f
            b = p.beta * (l - l0) + (1 - p.beta) * b
    elif alg == abi.ALG_HW:
        m = period
        L0, M2 = src[:m].mean(), src[m:2 * m].mean()
        l, b = L0, (M2 - L0) / m
        S = list(src[:m] - L0)
        sse = 0.0
        z = np.concatenate([src, cur])
        for t in range(m, len(z)):
            i = t % m
            f = l + b + S[i]
            if t >= len(src):
                ups.append(f + k * sigma)
                los.append(max(mlb, f - k * sigma))
            else:
                sse += (z[t] - f) ** 2
            l0 = l
            l = p.alpha * (z[t] - S[i]) + (1 - p.alpha) * (l + b)   # textbook additive HW
            b = p.beta * (l - l0) + (1 - p.beta) * b
            S[i] = S[i] + p.gamma * (1 - p.alpha) * (z[t] - f)
            if t == len(src) - 1:
                sigma = np.sqrt(sse / (len(src) - m))
    flags = [((bound & 1) and x > u) or ((bound & 2) and x < lo) for x, u, lo in zip(cur, ups, los)]
    return np.array(ups), np.array(los), np.array(flags, bool), sigma


@pytest.mark.parametrize("alg", [abi.ALG_MA_ALL, abi.ALG_MA_ROLLING, abi.ALG_EWMA, abi.ALG_DES, abi.ALG_HW])
@pytest.mark.parametrize("mt", [abi.MT_ERROR5XX, abi.MT_LATENCY])
def test_band_models_against_numpy(alg, mt):
    rng = np.random.default_rng(alg * 10 + mt)
    p = oracle.params_default()
    p.algorithms = alg
    p.band_algorithm = alg
    p.pairwise = abi.PW_NONE
    p.period = 24
    p.ma_window = 17
    p.rules[abi.MT_LATENCY].threshold = 2.5   # make both-bound flags actually fire
    units = []
    for u in range(12):
        n = int(rng.integers(60, 200))
        t = np.arange(n + 40)
        z = 50 + 10 * np.sin(2 * np.pi * t / 24) + rng.normal(0, 2, n + 40) + 0.01 * t
        z[n + 5] += 40
        z[n + 9] -= 45
        z = np.maximum(z, 0).astype(np.float32)
        units.append((z[:n], [], z[n:]))
    csr = abi.HostCsr.from_units(units, metric_type=[mt] * len(units))
    sc, up, lo = oracle.score(csr, p, want_bands=True)
    rule = p.rules[mt]
    a = {1: 0, 2: 1, 4: 2, 8: 3, 16: 4}[alg]
    for u, (h, _, c) in enumerate(units):
        ups, los, flags, sigma = _ref_bands(h, c, alg, p, rule.threshold, rule.bound, rule.min_lower_bound, 24)
        o = int(csr.offsets[u]) + len(h)
        np.testing.assert_allclose(up[o:o + len(c)], ups, rtol=2e-6)
        np.testing.assert_allclose(lo[o:o + len(c)], los, rtol=2e-6, atol=1e-6)
        assert int(sc[u]["n_anom"][a]) == int(flags.sum())
        assert int(sc[u]["first_idx"][a]) == (int(np.argmax(flags)) if flags.any() else abi.NONE_IDX)
        assert rel(sc[u]["sigma"][a], sigma) < 1e-9
        assert bool(sc[u]["status"] & abi.S_ANOMALY) == bool(flags.any())
        assert sc[u]["status"] >> (abi.S_ALG_SHIFT + a) & 1


def test_differs_lowers_threshold():
    """design.md:35: different distributions -> lower threshold."""
    rng = np.random.default_rng(5)
    base = rng.normal(100, 5, 60).astype(np.float32)
    cur_same = rng.normal(100, 5, 60).astype(np.float32)
    cur_shift = rng.normal(112, 5, 60).astype(np.float32)
    p = oracle.params_default()
    csr = abi.HostCsr.from_units([([], base, cur_same), ([], base, cur_shift)],
                                 metric_type=[abi.MT_CPU, abi.MT_CPU])
    s = oracle.score(csr, p)[0]
    assert not s[0]["status"] & abi.S_DIFFERS and s[0]["k_eff"] == 5.0
    assert s[1]["status"] & abi.S_DIFFERS and s[1]["k_eff"] == 2.5
    for f in (abi.S_MW_RUN, abi.S_WILCOXON_RUN, abi.S_KRUSKAL_RUN):
        assert s[1]["status"] & f
    # ANY vs ALL and the MIN_*_DATA_POINTS gates (brain.yaml:74-79)
    p.min_mann_white = 1000
    s = oracle.score(csr, p)[0]
    assert not s[1]["status"] & abi.S_MW_RUN and s[1]["status"] & abi.S_DIFFERS
    p.pairwise = abi.PW_MANN_WHITE
    s = oracle.score(csr, p)[0]
    assert not s[1]["status"] & abi.S_DIFFERS and s[1]["p_mw"] == 1.0


def test_unknown_and_verdict_fold():
    """converter.go:16-21 statuses; unknown when the current metric is missing (design.md:25)."""
    rng = np.random.default_rng(9)
    ok = rng.normal(10, 1, 80).astype(np.float32)
    spike = ok.copy()
    spike[70] = 100
    units = [
        (ok[:50], [], ok[50:]),        # svc0 healthy
        (ok[:50], [], spike[50:]),     # svc0 anomalous -> svc0 unhealth
        (ok[:50], [], []),             # svc1 no current -> unknown
        (ok[:5], ok[:6], ok[50:]),     # svc1 no model window -> unknown
        (ok[:50], [], ok[50:]),        # svc2 healthy
        (ok[:50], [], []),             # svc2 unknown but not all -> health
    ]
    csr = abi.HostCsr.from_units(units, service_id=[0, 0, 1, 1, 2, 2], metric_type=[abi.MT_CPU] * 6)
    p = oracle.params_default()
    sc = oracle.score(csr, p)[0]
    assert [bool(s["status"] & abi.S_UNKNOWN) for s in sc] == [False, False, True, True, False, True]
    v = oracle.verdicts(csr, sc, 4, first_unit=100)
    assert list(v["verdict"]) == [abi.V_UNHEALTH, abi.V_UNKNOWN, abi.V_HEALTH, abi.V_UNKNOWN]
    assert list(v["n_units"]) == [2, 2, 2, 0]
    assert int(v[0]["worst_unit"]) == 101 and int(v[0]["worst_unit_points"]) == 1
    assert int(v[0]["n_anom_units"]) == 1 and int(v[0]["n_anom_points"]) == 1
    assert int(v[2]["worst_unit"]) == abi.NONE_IDX


def test_cfg1_single_canary_job():
    """BASELINE.json configs[0]: 1 app, 5 metrics, 60-pt baseline vs 30-pt current, MA band + MWU."""
    p = oracle.params_default()
    p.pairwise = abi.PW_MANN_WHITE
    csr = oracle.synth(5, 0, 60, 30, units_per_service=5, n_metric_types=5, flags=abi.SYNTH_ANOMALIES)
    sc = oracle.score(csr, p)[0]
    assert all(s["status"] & abi.S_MW_RUN and s["status"] & abi.S_MODEL_BASE for s in sc)
    assert all(0 <= int(s["mw_2u1"]) <= 2 * 60 * 30 for s in sc)
    v = oracle.verdicts(csr, sc, 1)
    assert int(v[0]["n_units"]) == 5


def test_synth_deterministic_and_shaped():
    a = oracle.synth(64, 100, 30, 30, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES, threads=1)
    b = oracle.synth(64, 100, 30, 30, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES, threads=4)
    assert np.array_equal(a.values, b.values) and np.array_equal(a.seg_lens, b.seg_lens)
    # sharding: units 32.. generated with first_unit=32 equal the tail of the full batch
    c = oracle.synth(32, 100, 30, 30, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES, first_unit=32)
    stride = int(a.offsets[1])
    assert np.array_equal(c.values[:32 * stride], a.values[32 * stride:64 * stride])
    assert np.array_equal(c.service_id, a.service_id[32:])
    assert (a.offsets % 4 == 0).all() and np.isfinite(a.values).all() and (a.values >= 0).all()
    err = a.metric_type == abi.MT_ERROR5XX
    u = int(np.argmax(err))
    h = a.segments(u)[0]
    assert len(np.unique(h)) < len(h)  # quantised -> ties

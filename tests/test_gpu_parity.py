"""GPU parity tests: the CUDA path through the C ABI versus the CPU oracle on the same inputs.

Bar (BASELINE.json north_star): rank statistics, anomaly counts/indices, statuses and verdicts
bit-identical; smoothed values, sigmas, bands, H and p-values within 1e-5 relative."""
import json
import os

import numpy as np
import pytest

from foremast_b200 import _abi as abi

pytestmark = pytest.mark.gpu

REL_TOL = 1e-5  # north_star tolerance for floating-point outputs
INT_FIELDS = ["status", "n_anom", "first_idx", "w_n", "mw_2u1", "mw_2r1", "mw_ties", "w_2wplus", "w_2wminus", "w_ties"]
FLT_FIELDS = ["p_mw", "p_w", "p_k", "kw_h", "k_eff", "ma_mean", "ma_std", "sigma"]


@pytest.fixture(scope="module")
def ctx():
    from foremast_b200.api import Context
    c = Context()
    yield c
    c.close()


@pytest.fixture(scope="module")
def oracle():
    from oracle import oracle as o
    return o


def assert_scores_equal(got, ref):
    for f in INT_FIELDS:
        assert np.array_equal(got[f], ref[f]), (f, np.argwhere(got[f] != ref[f])[:5], got[f][got[f] != ref[f]][:5], ref[f][got[f] != ref[f]][:5])
    worst = 0.0
    for f in FLT_FIELDS:
        a, b = got[f].astype(np.float64), ref[f].astype(np.float64)
        err = np.abs(a - b) / np.maximum(np.abs(b), 1e-300)
        err[(a == b)] = 0.0
        if f == "sigma":
            # Holt-Winters residual sigma of a constant series: exactly 0 in the sequential oracle, the
            # re-associated scan leaves rounding noise that a marginally stable parameter set amplifies
            # (seen: 3.5e-11 on level 2.0 after 4500 samples) -- compare on the scale of the guard that
            # keeps such noise out of the flags (1e-9 of the level, DECISIONS "HW")
            scale = np.maximum(np.abs(ref["ma_mean"]), 1.0)
            err[:, 4] = np.where(np.abs(a[:, 4] - b[:, 4]) <= 1e-9 * scale, 0.0, err[:, 4])
        assert np.all(err <= REL_TOL), (f, float(err.max()))
        worst = max(worst, float(err.max()) if err.size else 0.0)
    return worst


def full_params(oracle, algs=abi.ALG_MA_ALL | abi.ALG_MA_ROLLING | abi.ALG_EWMA | abi.ALG_DES, pairwise=abi.PW_ALL):
    p = oracle.params_default()
    p.algorithms = algs
    p.pairwise = pairwise
    return p


def check_batch(ctx, oracle, batch, p, n_services=None):
    host = batch.download()
    res = ctx.score(batch, p)
    got = res.series()
    ref, _, _ = oracle.score(host, p)
    worst = assert_scores_equal(got, ref)
    ns = batch.info()["n_services"]
    v_ref = oracle.verdicts(host, ref, ns)
    v_got = res.verdicts()
    assert v_got.tobytes() == v_ref.tobytes()
    res.free()
    return got, worst


def test_synth_generator_bit_identical(ctx, oracle):
    for shape in [(0, 60, 30), (100, 30, 30), (1440, 0, 60), (33, 7, 1)]:
        flags = abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES
        b = ctx.batch_synth(517, *shape, units_per_service=40, n_metric_types=8, flags=flags, first_unit=1000)
        g = b.download()
        h = oracle.synth(517, *shape, units_per_service=40, n_metric_types=8, flags=flags, first_unit=1000)
        assert np.array_equal(g.offsets, h.offsets) and np.array_equal(g.seg_lens, h.seg_lens)
        assert np.array_equal(g.service_id, h.service_id) and np.array_equal(g.metric_type, h.metric_type)
        assert g.values.view(np.uint32).tobytes() == h.values.view(np.uint32).tobytes()
        b.free()


def test_demo_fixture_through_incremental_ingest(ctx, oracle, golden_dir):
    """The reference's demo profile data1.txt / data2.txt (spikes at idx 1 and 21) -> unhealthy."""
    fx = json.load(open(os.path.join(golden_dir, "demo_fixture.json")))
    p = oracle.params_default()
    b = ctx.batch_from_units([([], fx["data1"], fx["data2"])], metric_type=[abi.MT_ERROR5XX])
    res = ctx.score(b, p)
    s = res.series()[0]
    g = fx["pooled"]
    assert (int(s["mw_2r1"]), int(s["mw_2u1"]), int(s["mw_ties"])) == (g["mw_2r1"], g["mw_2u1"], g["mw_ties"]) == (2946, 1614, 76068)
    assert (int(s["w_n"]), int(s["w_2wplus"]), int(s["w_2wminus"]), int(s["w_ties"])) == (g["w_n"], g["w_2wplus"], g["w_2wminus"], g["w_ties"])
    for k in ("p_mw", "p_k", "p_w", "kw_h"):
        assert abs(s[k] - g[k]) / g[k] < REL_TOL
    assert int(s["n_anom"][0]) == 2 and int(s["first_idx"][0]) == 1
    v = res.verdicts()
    assert int(v[0]["verdict"]) == abi.V_UNHEALTH and abi.V_STATUS[int(v[0]["verdict"])] == "completed_unhealth"
    res.free()
    b.free()


def test_rank_cases_fixture(ctx, oracle, golden_dir):
    cases = json.load(open(os.path.join(golden_dir, "rank_cases.json")))["cases"]
    p = oracle.params_default()
    p.min_mann_white = p.min_wilcoxon = p.min_kruskal = 1
    b = ctx.batch_from_units([([], c["x"], c["y"]) for c in cases])
    res = ctx.score(b, p)
    sc = res.series()
    for c, s in zip(cases, sc):
        if c.get("degenerate"):
            assert s["p_mw"] == 1.0 and s["p_k"] == 1.0 and s["p_w"] == 1.0
            continue
        for k in ("mw_2r1", "mw_2u1", "mw_ties", "w_n"):
            assert int(s[k]) == c[k], k
        if c["w_n"]:
            for k in ("w_2wplus", "w_2wminus", "w_ties"):
                assert int(s[k]) == c[k], k
        assert abs(s["p_mw"] - c["p_mw"]) <= REL_TOL * c["p_mw"]
        assert abs(s["p_k"] - c["p_k"]) <= REL_TOL * c["p_k"]
    res.free()
    b.free()


@pytest.mark.parametrize("shape,n", [((0, 60, 30), 5), ((0, 720, 720), 3000), ((1440, 0, 60), 2000),
                                      ((300, 100, 90), 4097), ((10080, 30, 30), 300), ((0, 1024, 1024), 257),
                                      ((0, 2048, 1500), 131), ((40, 1025, 33), 65)])
def test_score_parity_on_synthetic_batches(ctx, oracle, shape, n):
    flags = abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES
    b = ctx.batch_synth(n, *shape, units_per_service=80 if n > 100 else 5, n_metric_types=8 if n > 5 else 5, flags=flags)
    p = full_params(oracle)
    got, worst = check_batch(ctx, oracle, b, p)
    if n > 1000:
        assert (got["status"] & abi.S_ANOMALY).any() and not (got["status"] & abi.S_ANOMALY).all()
    print(f"shape={shape} n={n} worst float rel err {worst:.3e}")
    b.free()


@pytest.mark.parametrize("longest", [3, 32, 33, 64, 65, 96, 97, 128, 150, 192, 193, 256, 300, 384, 385, 512, 600, 720,
                                     768, 769, 1024, 1025, 1536, 1537, 2048])
def test_rank_sort_size_classes(ctx, oracle, longest):
    """The rank kernels sort in 32*R register slots, R = 2^q or 3*2^q chosen from the longest
    segment of the batch: one batch per class boundary, ragged lengths below it, with ties (values
    rounded to a coarse grid), without, and all-equal / reversed runs."""
    rng = np.random.default_rng(longest)
    units = []
    for i in range(40):
        nb = longest if i < 2 else int(rng.integers(max(1, longest // 2), longest + 1))
        nc = longest if i in (0, 3) else int(rng.integers(max(1, longest // 2), longest + 1))
        base = rng.normal(100, 10, nb)
        cur = rng.normal(100 + (i % 3), 10, nc)
        if i % 4 == 1:
            base, cur = np.round(base), np.round(cur)            # many ties, across and inside groups
        elif i % 4 == 2:
            base, cur = np.round(base / 8) * 8, np.round(cur / 8) * 8  # long runs
        elif i == 7:
            base[:] = 5.0
            cur[:] = 5.0
        elif i == 11:
            base, cur = np.sort(base)[::-1].copy(), np.sort(cur)
        units.append((rng.normal(100, 10, 40), base, cur))
    b = ctx.batch_from_units(units, service_id=[i // 10 for i in range(40)], metric_type=[i % 8 for i in range(40)])
    p = full_params(oracle, algs=abi.ALG_MA_ALL)
    p.min_mann_white = p.min_wilcoxon = p.min_kruskal = 1
    check_batch(ctx, oracle, b, p)
    b.free()


@pytest.mark.parametrize("pairwise", [abi.PW_NONE, abi.PW_ANY, abi.PW_MANN_WHITE, abi.PW_WILCOXON, abi.PW_KRUSKAL])
@pytest.mark.parametrize("algs", [abi.ALG_MA_ALL, abi.ALG_EWMA, abi.ALG_DES, abi.ALG_MA_ALL | abi.ALG_DES,
                                  abi.ALG_MA_ROLLING, abi.ALG_MA_ROLLING | abi.ALG_EWMA])
def test_parameter_combinations(ctx, oracle, pairwise, algs):
    b = ctx.batch_synth(1500, 200, 64, 64, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES, seed=7 + pairwise)
    p = full_params(oracle, algs, pairwise)
    check_batch(ctx, oracle, b, p)
    b.free()


@pytest.mark.parametrize("window", [1, 7, 31, 32, 33, 100, 5000])
def test_rolling_window_sizes(ctx, oracle, window):
    """Rolling band for windows shorter / longer than a staging chunk, the model window and the unit."""
    b = ctx.batch_synth(700, 150, 40, 70, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES, seed=window)
    p = full_params(oracle, abi.ALG_MA_ROLLING, abi.PW_NONE)
    p.ma_window = window
    check_batch(ctx, oracle, b, p)
    b.free()


@pytest.mark.parametrize("shape,period,n", [((400, 0, 90), 96, 600), ((10080, 0, 60), 1440, 150),
                                            ((2900, 30, 1500), 1440, 100), ((64, 0, 33), 32, 300),
                                            ((5000, 40, 40), 2048, 64), ((700, 0, 150), 100, 300),
                                            ((3000, 0, 700), 1000, 80)])
@pytest.mark.parametrize("seasonals", ["tensor_memory", "shared_memory"])
def test_holt_winters(ctx, oracle, shape, period, n, seasonals, monkeypatch):
    """BASELINE.json configs[3]: Holt-Winters seasonal (period 1440, 7-day history), continuous mode.
    The kernels scan each season in parallel, so floats are compared at 1e-5 relative.  Periods that
    are multiples of 32 keep the seasonals in tensor memory (k_hw_tm), the others (100, 1000 here) in
    shared memory (k_hw); FOREMAST_GPU_HW_SMEM forces the latter so both run on every shape."""
    if seasonals == "shared_memory":
        monkeypatch.setenv("FOREMAST_GPU_HW_SMEM", "1")
    b = ctx.batch_synth(n, *shape, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES, seed=period)
    p = full_params(oracle, abi.ALG_HW | abi.ALG_MA_ALL, abi.PW_ALL)
    p.period = period
    got, worst = check_batch(ctx, oracle, b, p)
    ran = (got["status"] >> (abi.S_ALG_SHIFT + 4)) & 1
    assert ran.mean() > 0.9
    print(f"HW shape={shape} period={period} worst rel err {worst:.3e}, anomalous {(got['n_anom'][:, 4] > 0).mean():.3f}")
    # bands of the HW model
    p.band_algorithm = abi.ALG_HW
    host = b.download()
    ref, up, lo = oracle.score(host, p, want_bands=True)
    res = ctx.score(b, p)
    for u in (0, n // 2, n - 1):
        gu, gl = res.bands(u)
        o = int(host.offsets[u]) + int(host.seg_lens[u][0]) + int(host.seg_lens[u][1])
        if ran[u]:
            np.testing.assert_allclose(gu, up[o:o + gu.size], rtol=1e-5)
            np.testing.assert_allclose(gl, lo[o:o + gl.size], rtol=1e-5, atol=1e-30)
    res.free()
    b.free()


@pytest.mark.parametrize("period,shape", [(256, (1041, 0, 133)), (384, (1553, 0, 200)), (512, (2065, 7, 600)),
                                          (1024, (4113, 0, 40)), (1440, (4320, 0, 1440)), (1440, (5000, 0, 3000)),
                                          (64, (300, 0, 1)), (2048, (6200, 0, 2100)),
                                          (168, (1100, 0, 200)), (100, (777, 0, 33)), (1000, (4100, 0, 1300)),
                                          (34, (180, 0, 40)), (1530, (5000, 0, 100)), (62, (400, 0, 77))])
def test_holt_winters_tile_classes(ctx, oracle, period, shape):
    """k_hw_tm takes different paths for a whole season of the model window, for the super-tile the model
    window ends in, for whole and ragged super-tiles of current samples and for the last super-tile of a
    unit, and its blocks of four seasonals end differently for every chunk length (period / 32 = 8, 12,
    16, 32, 45, 2, 64 here; then periods that are k x L with L < 32 lanes: 168 = 6 x 28, 100 = 4 x 25,
    1000 = 40 x 25, 34 = 2 x 17, 1530 = 51 x 30, 62 = 2 x 31).  Scores, bands and the anomaly list against the oracle on each."""
    n = 120
    b = ctx.batch_synth(n, *shape, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES, seed=period + shape[2])
    p = full_params(oracle, abi.ALG_HW, abi.PW_NONE)
    p.period = period
    got, worst = check_batch(ctx, oracle, b, p)
    ran = (got["status"] >> (abi.S_ALG_SHIFT + 4)) & 1
    assert ran.mean() > 0.5
    p.band_algorithm = abi.ALG_HW
    host = b.download()
    ref, up, lo = oracle.score(host, p, want_bands=True)
    res = ctx.score(b, p)
    for u in (0, n // 3, n - 1):
        gu, gl = res.bands(u)
        o = int(host.offsets[u]) + int(host.seg_lens[u][0]) + int(host.seg_lens[u][1])
        if ran[u]:
            np.testing.assert_allclose(gu, up[o:o + gu.size], rtol=1e-5)
            np.testing.assert_allclose(gl, lo[o:o + gl.size], rtol=1e-5, atol=1e-30)
    offsets, idx, val = res.anomalies(abi.ALG_HW)
    assert np.array_equal(np.diff(offsets.astype(np.int64)), ref["n_anom"][:, 4])
    for u in np.flatnonzero(ref["n_anom"][:, 4])[:40]:
        cur = host.segments(u)[2]
        sl = slice(int(offsets[u]), int(offsets[u + 1]))
        assert np.all(np.diff(idx[sl].astype(np.int64)) > 0) and np.array_equal(val[sl], cur[idx[sl]])
        assert int(idx[sl][0]) == int(ref["first_idx"][u, 4])
    res.free()
    b.free()


def test_ragged_large_batch_takes_the_ordered_band_path(ctx, oracle):
    """Batches of 131 072 units and more are scored by k_band in a per-tile order that puts units of similar
    length into the same warp (fm_launch_band_order).  Every unit here has its own lengths, so no tile gets the
    identity order; scores and verdicts against the oracle on all of them."""
    n = 140_000
    rng = np.random.default_rng(77)
    nh = rng.choice([0, 0, 0, 40, 100], n)
    nb, nc = rng.integers(20, 200, n), rng.integers(1, 160, n)
    tot = nh + nb + nc
    offs = np.zeros(n + 1, np.int64)
    np.cumsum(tot + (-tot) % 4, out=offs[1:])
    vals = (rng.normal(50.0, 5.0, int(offs[-1])) + (rng.random(int(offs[-1])) < 0.002) * 60.0).astype(np.float32)
    csr = abi.HostCsr(vals, offs, np.stack([nh, nb, nc], axis=1), np.arange(n) // 20, rng.integers(0, 10, n))
    b = ctx.batch_from_csr(csr)
    p = full_params(oracle, abi.ALG_MA_ALL | abi.ALG_MA_ROLLING | abi.ALG_EWMA | abi.ALG_DES, abi.PW_MANN_WHITE)
    p.ma_window = 10
    got, worst = check_batch(ctx, oracle, b, p)
    assert (got["n_anom"].sum(axis=0)[:4] > 0).all()
    b.free()


def test_mixed_sort_sizes_run_per_class(ctx, oracle):
    """A batch whose units need different sort sizes in the rank kernels (short canary windows with long ones and
    over-long ones among them) is grouped by size class and every class runs in its own network; before, one long unit
    put all of them into the largest.  Scores against the oracle for every pairwise mode."""
    rng = np.random.default_rng(21)
    n = 3000
    lens = rng.choice([1, 7, 30, 33, 64, 65, 100, 190, 257, 500, 760, 1000, 1500, 2048, 2100, 2600], (n, 2))
    units = []
    for nb, nc in lens:
        kind = rng.integers(0, 3)
        z = rng.normal(10, 2, nb + nc) if kind == 0 else np.round(rng.normal(10, 2, nb + nc)) if kind == 1 else np.full(nb + nc, 3.0)
        units.append(([], z[:nb].astype(np.float32), z[nb:].astype(np.float32)))
    sid = np.arange(n) // 10
    mts = rng.integers(0, 8, n)
    p = full_params(oracle, abi.ALG_MA_ALL, abi.PW_ALL)
    b = ctx.batch_from_units(units, service_id=sid, metric_type=mts)
    got, _ = check_batch(ctx, oracle, b, p)
    ran = got["status"] & (abi.S_MW_RUN | abi.S_WILCOXON_RUN | abi.S_KRUSKAL_RUN)
    assert (ran != 0).mean() > 0.5 and (ran == 0).any()   # the over-long ones are not ranked (D17)
    for pw in (abi.PW_WILCOXON, abi.PW_MANN_WHITE, abi.PW_KRUSKAL):
        p.pairwise = pw
        check_batch(ctx, oracle, b, p)
    b.free()


@pytest.mark.parametrize("period,hist", [(96, 500), (1440, 4400), (100, 777)])
def test_model_cache_continuous_mode(ctx, oracle, period, hist):
    """Continuous mode: fit Holt-Winters once, then score successive current windows from the cached
    state.  Must equal one oracle pass over history + all windows (sigma is the fit's)."""
    n, c1, c2 = 300, 37, 70
    full = oracle.synth(n, hist, 0, c1 + c2, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES, seed=period, units_per_service=6)
    units = [full.segments(u) for u in range(n)]
    p = full_params(oracle, abi.ALG_HW, abi.PW_NONE)
    p.period = period
    k1 = [min(c1, len(c)) for _, _, c in units]
    ref1 = oracle.score(abi.HostCsr.from_units([(h, [], c[:k]) for (h, _, c), k in zip(units, k1)], full.service_id, full.metric_type), p)[0]
    ref12 = oracle.score(full, p)[0]
    hb = ctx.batch_from_units([(h, [], []) for h, _, _ in units], full.service_id, full.metric_type)
    b1 = ctx.batch_from_units([([], [], c[:k]) for (_, _, c), k in zip(units, k1)], full.service_id, full.metric_type)
    b2 = ctx.batch_from_units([([], [], c[k:]) for (_, _, c), k in zip(units, k1)], full.service_id, full.metric_type)
    model = ctx.model_fit(hb, p)
    r1 = model.score(b1)
    g1 = r1.series()
    r2 = model.score(b2)
    g2 = r2.series()
    fitted = (ref12["status"] >> (abi.S_ALG_SHIFT + 4)) & 1
    assert fitted.mean() > 0.9
    ok1 = (fitted == 1) & (np.array(k1) > 0)
    assert np.array_equal(g1["n_anom"][ok1, 4], ref1["n_anom"][ok1, 4])
    assert np.array_equal(g1["first_idx"][ok1, 4], ref1["first_idx"][ok1, 4])
    np.testing.assert_allclose(g1["sigma"][ok1, 4], ref12["sigma"][ok1, 4], rtol=REL_TOL)
    n2 = np.array([len(c) - k for (_, _, c), k in zip(units, k1)])
    ok2 = (fitted == 1) & (n2 > 0)
    assert np.array_equal(g2["n_anom"][ok2, 4], (ref12["n_anom"][:, 4] - ref1["n_anom"][:, 4])[ok2])
    # units without two seasons of history (or without new samples) are unknown, the rest carry the HW bit
    assert np.all(g1["status"][fitted == 0] & abi.S_UNKNOWN) and np.all(g2["status"][n2 == 0] & abi.S_UNKNOWN)
    assert np.all((g2["status"][ok2] >> (abi.S_ALG_SHIFT + 4)) & 1)
    v = r2.verdicts()
    assert int(v["n_units"].sum()) == n and int(v["n_anom_units"].sum()) == int(((g2["status"] & abi.S_ANOMALY) != 0).sum())
    assert (g2["n_anom"][:, 4] > 0).any()
    from foremast_b200._lib import ForemastGpuError
    with pytest.raises(ForemastGpuError):
        r2.anomalies(abi.ALG_HW)
    for o in (r1, r2, model, hb, b1, b2):
        o.free()


def test_edge_cases(ctx, oracle):
    rng = np.random.default_rng(11)
    z = rng.normal(10, 1, 200).astype(np.float32)
    units = [
        ([], [], []),                       # empty unit
        (z[:50], [], []),                   # no current -> unknown
        ([], z[:4], z[4:40]),               # model window too short -> unknown, tests gated off
        (z[:50], z[50:55], z[60:90]),       # baseline below MIN_* -> no rank tests
        ([1.5] * 40, [1.5] * 30, [1.5] * 30),   # constant series, all ties
        ([0.0] * 40, [0.0] * 30, [0.0] * 29 + [5.0]),  # zeros then a spike
        (z[:10], z[10:30], z[30:50]),       # exactly at the minimum sizes
        (z[:1], z[:33], z[33:66]),          # hist too short, base is the model window
        ([], z[:100], z[100:101]),          # single current point
        ([-0.0, 0.0, 1.0] * 10, [-0.0] * 20, [0.0] * 20),  # signed zeros (canonicalised at ingest)
        ([float("nan"), 1.0, float("inf")] + list(z[:30]), z[30:60], z[60:90]),  # gaps dropped at ingest
    ]
    sid = [0, 0, 1, 1, 2, 2, 3, 3, 4, 5, 5]
    mt = [abi.MT_CPU, abi.MT_LATENCY, abi.MT_ERROR5XX, abi.MT_MEMORY, abi.MT_ERROR4XX, abi.MT_ERROR5XX,
          abi.MT_TPS, 15, 200, abi.MT_LATENCY, abi.MT_CPU]
    b = ctx.batch_from_units(units, service_id=sid, metric_type=mt)
    host = b.download()
    assert list(host.seg_lens[10]) == [31, 30, 30]
    p = full_params(oracle)
    got, _ = check_batch(ctx, oracle, b, p)
    assert got[0]["status"] & abi.S_UNKNOWN and got[1]["status"] & abi.S_UNKNOWN and got[2]["status"] & abi.S_UNKNOWN
    assert got[5]["status"] & abi.S_ANOMALY and not got[4]["status"] & abi.S_ANOMALY
    b.free()
    # empty batch
    e = ctx.batch_from_units([])
    r = ctx.score(e, p)
    assert r.series().size == 0 and r.verdicts().size == 0
    r.free()
    e.free()
    # ... also through the two bulk host paths, with no arrays at all (n_units = 0, NULL pointers)
    import ctypes as C
    L, h = ctx._L, C.c_void_p()
    empty = abi.Csr(0, None, None, None, None, None)
    assert L.fm_batch_from_csr(ctx._h, C.byref(empty), C.byref(h)) == 0
    L.fm_batch_free(h)
    assert L.fm_score_host(ctx._h, C.byref(empty), C.byref(p), 0, None) in (0, abi.E_INVALID)
    one = np.zeros(1, dtype=abi.VERDICT)
    assert L.fm_score_host(ctx._h, C.byref(empty), C.byref(p), 1, one.ctypes.data) == 0
    assert int(one[0]["n_units"]) == 0 and int(one[0]["verdict"]) == abi.V_UNKNOWN


def test_size_limits(ctx, oracle):
    """Longest inputs: a 50 000-sample history streams through the band kernel; rank segments above
    FM_RANK_MAX_SEGMENT (2048) samples make that ONE unit's rank tests "not run" (D17) while the rest of the
    unit and of the batch is scored as usual -- e.g. a 7-day baseline against a 30-minute current window."""
    rng = np.random.default_rng(2)
    z = rng.normal(50, 3, 60000).astype(np.float32)
    b = ctx.batch_from_units([(z[:50000], z[50000:52048], z[52048:54096]), (z[:37], z[100:140], z[200:260])],
                             service_id=[0, 0], metric_type=[abi.MT_CPU, abi.MT_LATENCY])
    p = full_params(oracle)
    check_batch(ctx, oracle, b, p)
    b.free()
    units = [(z[:100], z[100:2200], z[3000:3100]),          # baseline 2100 > 2048: MWU / KW off, Wilcoxon on 100 pairs
             (z[:0], z[5000:15080], z[20000:20030]),        # 7-day baseline vs 30 points, baseline is the model window
             (z[:64], z[1000:4000], z[6000:9000]),          # 3000 pairs: nothing ranked
             (z[:64], z[100:140], z[200:260]),              # an ordinary unit beside them
             (z[:64], z[300:2348], z[2400:4448])]           # exactly 2048 / 2048: everything runs
    big = ctx.batch_from_units(units, service_id=[0, 0, 1, 1, 2], metric_type=[abi.MT_CPU] * 5)
    for pairwise in (abi.PW_ALL, abi.PW_ANY, abi.PW_WILCOXON, abi.PW_MANN_WHITE, abi.PW_NONE):
        p.pairwise = pairwise
        got, _ = check_batch(ctx, oracle, big, p)
    p.pairwise = abi.PW_ALL
    got, _ = check_batch(ctx, oracle, big, p)
    run = lambda u, bit: bool(int(got["status"][u]) & bit)  # noqa: E731
    assert not run(0, abi.S_MW_RUN) and not run(0, abi.S_KRUSKAL_RUN) and run(0, abi.S_WILCOXON_RUN)
    assert not run(1, abi.S_MW_RUN) and run(1, abi.S_WILCOXON_RUN)
    assert not (run(2, abi.S_MW_RUN) or run(2, abi.S_KRUSKAL_RUN) or run(2, abi.S_WILCOXON_RUN))
    assert run(3, abi.S_MW_RUN) and run(3, abi.S_WILCOXON_RUN) and run(3, abi.S_KRUSKAL_RUN)
    assert run(4, abi.S_MW_RUN) and run(4, abi.S_WILCOXON_RUN) and run(4, abi.S_KRUSKAL_RUN)
    big.free()


def test_bulk_csr_ingest_and_realignment(ctx, oracle):
    h = oracle.synth(300, 50, 21, 22, flags=abi.SYNTH_ANOMALIES)
    p = full_params(oracle)
    b = ctx.batch_from_csr(h)
    check_batch(ctx, oracle, b, p)
    b.free()
    # unaligned CSR (units packed back to back) is re-packed by the library
    vals, offs, pos = [], [], 0
    for u in range(h.n_units):
        hs, bs, cs = h.segments(u)
        offs.append(pos)
        vals += [hs, bs, cs]
        pos += hs.size + bs.size + cs.size
    offs.append(pos)
    packed = abi.HostCsr(np.concatenate(vals), offs, h.seg_lens, h.service_id, h.metric_type)
    b2 = ctx.batch_from_csr(packed)
    ref, _, _ = oracle.score(h, p)
    r2 = ctx.score(b2, p)
    assert_scores_equal(r2.series(), ref)
    v = ctx.score_host(packed, p)
    assert v.tobytes() == oracle.verdicts(h, ref, h.n_services).tobytes()
    r2.free()
    b2.free()


def test_score_host_pipelined_chunks(ctx, oracle, monkeypatch):
    """fm_score_host on an aligned CSR: chunked upload overlapped with scoring must give the same
    verdicts as the oracle, for many small chunks and for one big one."""
    h = oracle.synth(3000, 64, 40, 40, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES, units_per_service=7)
    p = full_params(oracle)
    ref, _, _ = oracle.score(h, p)
    want = oracle.verdicts(h, ref, h.n_services).tobytes()
    for chunk in ("16384", "100000", None):
        if chunk:
            monkeypatch.setenv("FOREMAST_GPU_CHUNK_BYTES", chunk)
        else:
            monkeypatch.delenv("FOREMAST_GPU_CHUNK_BYTES", raising=False)
        assert ctx.score_host(h, p).tobytes() == want
        assert ctx.score_host(h, p, h.n_services + 5)[:h.n_services].tobytes() == want
    # units of several rank sort sizes: every upload chunk groups its own units by size class
    rng = np.random.default_rng(4)
    units = []
    for nb, nc in rng.choice([5, 30, 40, 70, 130, 300, 700], (1500, 2)):
        z = np.round(rng.normal(20, 3, nb + nc), int(rng.integers(0, 3))).astype(np.float32)
        units.append(([], z[:nb], z[nb:]))
    hm = abi.HostCsr.from_units(units, np.arange(1500) // 6, rng.integers(0, 8, 1500))
    ref, _, _ = oracle.score(hm, p)
    want = oracle.verdicts(hm, ref, hm.n_services).tobytes()
    for chunk in ("65536", None):
        if chunk:
            monkeypatch.setenv("FOREMAST_GPU_CHUNK_BYTES", chunk)
        else:
            monkeypatch.delenv("FOREMAST_GPU_CHUNK_BYTES", raising=False)
        assert ctx.score_host(hm, p).tobytes() == want


def test_bands_materialised(ctx, oracle):
    b = ctx.batch_synth(200, 300, 0, 45, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES)
    host = b.download()
    for alg in (abi.ALG_MA_ALL, abi.ALG_MA_ROLLING, abi.ALG_EWMA, abi.ALG_DES):
        p = full_params(oracle, abi.ALG_MA_ALL | abi.ALG_MA_ROLLING | abi.ALG_EWMA | abi.ALG_DES, abi.PW_NONE)
        p.band_algorithm = alg
        ref, up, lo = oracle.score(host, p, want_bands=True)
        res = ctx.score(b, p)
        assert_scores_equal(res.series(), ref)
        for u in (0, 7, 50, 199):
            gu, gl = res.bands(u)
            o = int(host.offsets[u]) + int(host.seg_lens[u][0]) + int(host.seg_lens[u][1])
            n = int(host.seg_lens[u][2])
            assert gu.size == n
            assert np.array_equal(gu, up[o:o + n]) and np.array_equal(gl, lo[o:o + n])
        res.free()
    b.free()


@pytest.mark.parametrize("alg", [abi.ALG_MA_ALL, abi.ALG_MA_ROLLING, abi.ALG_EWMA, abi.ALG_DES, abi.ALG_HW])
def test_anomaly_list(ctx, oracle, alg):
    """Compact (index, value) list of the flagged current points (the [t0, v0, t1, v1, ...] pairs of
    DeploymentController.go:443-452): recomputed on the host from the oracle's bands."""
    n = 5000
    b = ctx.batch_synth(n, 220, 0, 70, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES, seed=alg)
    p = full_params(oracle, abi.ALG_ALL, abi.PW_NONE)
    p.period = 48
    p.band_algorithm = alg
    host = b.download()
    ref, up, lo = oracle.score(host, p, want_bands=True)
    p.band_algorithm = 0
    res = ctx.score(b, p)
    a = {1: 0, 2: 1, 4: 2, 8: 3, 16: 4}[alg]
    offsets, idx, val = res.anomalies(alg)
    got = res.series()
    assert np.array_equal(np.diff(offsets.astype(np.int64)), got["n_anom"][:, a])
    assert int(offsets[-1]) == int(got["n_anom"][:, a].sum()) > 0
    checked = 0
    for u in np.flatnonzero(got["n_anom"][:, a])[:400]:
        rule = p.rules[int(host.metric_type[u])] if p.rules[int(host.metric_type[u])].bound else p.rules[0]
        cur = host.segments(u)[2]
        o = int(host.offsets[u]) + int(host.seg_lens[u][0]) + int(host.seg_lens[u][1])
        ub, lb = up[o:o + cur.size], lo[o:o + cur.size]
        flags = ((rule.bound & 1) != 0) & (cur > ub) | ((rule.bound & 2) != 0) & (cur < lb)
        # float32 bands can tie with the sample where the fp64 comparison does not; the counts are
        # the exact ones, so only check the list against them and against unambiguous points
        sl = slice(int(offsets[u]), int(offsets[u + 1]))
        assert np.all(np.diff(idx[sl].astype(np.int64)) > 0)
        assert np.array_equal(val[sl], cur[idx[sl]])
        if int(flags.sum()) == int(got["n_anom"][u, a]):
            assert np.array_equal(idx[sl], np.flatnonzero(flags))
            checked += 1
        i1, v1 = res.anomalies_unit(int(u))
        assert np.array_equal(i1, idx[sl]) and np.array_equal(v1, val[sl])
    assert checked >= min(10, np.count_nonzero(got["n_anom"][:, a])) * 0.8
    assert int(got["first_idx"][np.flatnonzero(got["n_anom"][:, a])[0], a]) == int(idx[int(offsets[np.flatnonzero(got["n_anom"][:, a])[0]])])
    res.free()
    b.free()


def test_device_csr_and_alignment_errors(ctx, oracle):
    import torch
    from foremast_b200._lib import ForemastGpuError
    h = oracle.synth(64, 40, 20, 20)
    dv = torch.from_numpy(h.values).cuda()
    do = torch.from_numpy(h.offsets).cuda()
    ds = torch.from_numpy(h.seg_lens.astype(np.int32)).cuda()
    di = torch.from_numpy(h.service_id.astype(np.int32)).cuda()
    dm = torch.from_numpy(h.metric_type).cuda()
    b = ctx.batch_from_device(64, dv.data_ptr(), do.data_ptr(), ds.data_ptr(), di.data_ptr(), dm.data_ptr(), keepalive=(dv, do, ds, di, dm))
    check_batch(ctx, oracle, b, full_params(oracle))
    b.free()
    bad = do.clone()
    bad[3] += 1
    with pytest.raises(ForemastGpuError) as ei:
        ctx.batch_from_device(64, dv.data_ptr(), bad.data_ptr(), ds.data_ptr(), di.data_ptr(), dm.data_ptr())
    assert ei.value.code == abi.E_ALIGN


def test_sharded_batches_merge_like_multi_gpu(ctx, oracle):
    """The N>1 path on one GPU: two shards with global unit ids (fm_batch_set_shard), their dense
    per-service counters merged exactly as foremast_b200.dist does across ranks (sum / max on
    tensors aliasing the library's buffers), verdicts finalised after the merge."""
    import torch
    from foremast_b200 import dist as fdist
    n, ups = 3000, 40
    p = full_params(oracle)
    whole = oracle.synth(n, 64, 40, 40, units_per_service=ups, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES)
    ref, _, _ = oracle.score(whole, p)
    want = oracle.verdicts(whole, ref, whole.n_services)
    results = []
    for rank in range(2):
        lo, hi = fdist.shard_range(n, rank, 2, ups)
        b = ctx.batch_synth(hi - lo, 64, 40, 40, units_per_service=ups, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES,
                            first_unit=lo)
        b.set_shard(lo, whole.n_services)
        r = ctx.score(b, p)
        r.sync()
        results.append((b, r))
    c0, w0 = fdist.alias_service_buffers(results[0][1])
    c1, w1 = fdist.alias_service_buffers(results[1][1])
    c0 += c1
    torch.maximum(w0, w1, out=w0)
    torch.cuda.synchronize()
    got = results[0][1].verdicts()
    assert got.tobytes() == want.tobytes()
    assert (got["verdict"] == abi.V_UNHEALTH).any() and int(got["worst_unit"][got["verdict"] == abi.V_UNHEALTH].max()) >= n // 2
    for b, r in results:
        r.free()
        b.free()


def test_full_size_properties(ctx, oracle):
    """BASELINE.json headline shape: 1M series-windows x 1440 points (720 baseline + 720 current),
    all scorers.  Size-independent properties on everything, and the bit-exact oracle check on ALL of it."""
    n = 1_000_000
    b = ctx.batch_synth(n, 0, 720, 720, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES)
    p = full_params(oracle)
    res = ctx.score(b, p)
    sc = res.series()
    st = sc["status"]
    ran = (st & abi.S_MW_RUN) != 0
    assert ran.mean() > 0.99
    assert np.all(sc["mw_2u1"][ran] >= 0) and np.all(sc["mw_2u1"][ran] <= 2 * 720 * 720)
    wn = sc["w_n"].astype(np.int64)
    assert np.array_equal(sc["w_2wplus"] + sc["w_2wminus"], wn * (wn + 1))
    assert np.all((sc["p_mw"] >= 0) & (sc["p_mw"] <= 1) & (sc["p_k"] >= 0) & (sc["p_k"] <= 1))
    v = res.verdicts()
    assert int(v["n_units"].sum()) == n
    assert int(v["n_anom_units"].sum()) == int(((st & abi.S_ANOMALY) != 0).sum())
    assert int(v["n_unknown"].sum()) == int(((st & abi.S_UNKNOWN) != 0).sum())
    assert np.array_equal(v["verdict"] == abi.V_UNHEALTH, v["n_anom_units"] > 0)
    # idempotence: scoring the same batch again gives the same bytes
    res2 = ctx.score(b, p)
    assert res2.series().tobytes() == sc.tobytes()
    res2.free()
    # bit-exact against the oracle on EVERY unit of the batch and on every verdict (bench.py does the same on
    # every rank of every run)
    host = b.download()
    ref, _, _ = oracle.score(host, p)
    assert_scores_equal(sc, ref)
    assert v.tobytes() == oracle.verdicts(host, ref, b.info()["n_services"]).tobytes()
    res.free()
    b.free()


def test_library_comm_single_rank_merge_is_identity(ctx, oracle):
    """fm_comm_init with a world of one (loads NCCL, builds a communicator) and fm_result_merge: the merged
    verdicts of a single shard are its own verdicts."""
    from foremast_b200.api import Context
    c2 = Context()
    try:
        c2.comm_init(Context.comm_unique_id(), 0, 1)
        assert c2.comm_info() == (0, 1)
        b = c2.batch_synth(1000, 64, 40, 40, units_per_service=7, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES)
        p = full_params(oracle)
        res = c2.score(b, p)
        res.merge()
        got = res.verdicts()
        host = b.download()
        ref, _, _ = oracle.score(host, p)
        assert got.tobytes() == oracle.verdicts(host, ref, b.info()["n_services"]).tobytes()
        ns = b.info()["n_services"]
        assert res.verdicts_range(3, ns - 5).tobytes() == got[3:ns - 2].tobytes()   # a slice = the same records, ids kept
        assert res.verdicts_range(ns, 0).size == 0
        with pytest.raises(Exception):
            res.verdicts_range(ns - 1, 2)
        res.free()
        b.free()
    finally:
        c2.close()


def test_library_comm_two_gpus_merge_matches_oracle_fold(oracle):
    """Two contexts on two devices of one process, one thread each: shard by service, fm_result_merge through
    the C ABI (NCCL on the library's streams), merged verdicts on BOTH ranks = the oracle fold of the whole."""
    import threading
    from foremast_b200 import dist as fdist
    from foremast_b200._lib import ForemastGpuError
    from foremast_b200.api import Context
    try:
        Context(1).close()
    except ForemastGpuError:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    n, ups, world = 4001, 7, 2
    whole = oracle.synth(n, 64, 40, 40, units_per_service=ups, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES)
    p = full_params(oracle)
    ref, _, _ = oracle.score(whole, p)
    want = oracle.verdicts(whole, ref, whole.n_services)
    uid = Context.comm_unique_id()
    out, errs = [None] * world, []

    def run(rank):
        try:
            c = Context(rank)
            c.comm_init(uid, rank, world)
            lo, hi = fdist.shard_range(n, rank, world, ups)
            b = c.batch_synth(hi - lo, 64, 40, 40, units_per_service=ups, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES,
                              first_unit=lo)
            b.set_shard(lo, whole.n_services)
            res = c.score(b, p)
            res.merge()
            out[rank] = res.verdicts().copy()
            s_lo, s_hi = -(-lo // ups), -(-hi // ups)
            assert res.verdicts_range(s_lo, s_hi - s_lo).tobytes() == out[rank][s_lo:s_hi].tobytes()
            res.free()
            b.free()
            c.close()
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    ts = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=120)
    assert not errs, errs
    for r in range(world):
        assert out[r] is not None and out[r].tobytes() == want.tobytes()


def test_mann_whitney_search_kernel_extremes(ctx, oracle):
    """k_mw sorts only the baseline and finds every current sample by a lower bound, with equal values among
    the current samples resolved by the duplicate rounds: shapes and value patterns at the edges of that scheme --
    very unequal lengths, a full tile, every current sample equal, every current sample equal to a long run of
    the baseline, current samples all below / all above / all inside one gap, negative values and zeros of both
    signs, and keys that collide in the id table by construction (many distinct values)."""
    rng = np.random.default_rng(99)
    z = rng.normal(0, 1, 5000).astype(np.float32)
    grid = (np.arange(2048) * 0.5 - 300).astype(np.float32)
    units = [
        (z[:20], z[:2048], z[2048:2049]),                                   # 2048 vs 1
        (z[:20], z[:1], z[1:2049]),                                         # 1 vs 2048
        (z[:20], grid, grid[::-1].copy()),                                  # full tiles, every value met exactly once
        (z[:20], grid, np.full(2048, grid[77], np.float32)),                # all current samples equal one baseline value
        (z[:20], np.full(1000, 3.25, np.float32), np.full(900, 3.25, np.float32)),   # one pooled tie group
        (z[:20], np.repeat(grid[:64], 16), np.repeat(grid[32:96], 9)),      # long runs on both sides, half of them shared
        (z[:20], z[:700], z[700:1400] - 100),                               # all current below the baseline
        (z[:20], z[:700], z[700:1400] + 100),                               # all above
        (z[:20], np.concatenate([z[:350] - 50, z[350:700] + 50]), rng.integers(0, 3, 700).astype(np.float32) * 0.25),  # one gap, heavy ties
        (z[:20], np.array([-0.0, 0.0, -1.0, 1.0] * 100, np.float32), np.array([0.0, -0.0, 0.0, 2.0, -2.0] * 90, np.float32)),
        (z[:20], z[:720], np.arange(720, dtype=np.float32) * 1e-3),         # 720 distinct keys: slot collisions, no ties
        (z[:20], np.arange(720, dtype=np.float32), np.arange(720, dtype=np.float32)[::-1].copy() + 0.5),  # interleaved, no ties
        (z[:20], z[:5], z[5:10]),                                           # min_kruskal only
    ]
    b = ctx.batch_from_units(units, service_id=[i // 4 for i in range(len(units))], metric_type=[abi.MT_LATENCY] * len(units))
    for pairwise in (abi.PW_ALL, abi.PW_MANN_WHITE, abi.PW_KRUSKAL, abi.PW_ANY):
        p = full_params(oracle, algs=abi.ALG_MA_ALL, pairwise=pairwise)
        got, _ = check_batch(ctx, oracle, b, p)
    p = full_params(oracle, algs=abi.ALG_MA_ALL, pairwise=abi.PW_ALL)
    got, _ = check_batch(ctx, oracle, b, p)
    n = np.array([[len(u[1]), len(u[2])] for u in units], dtype=np.int64)
    assert np.array_equal(got["mw_2u1"][:12] >= 0, np.ones(12, bool)) and np.all(got["mw_2u1"] <= 2 * n[:, 0] * n[:, 1])
    assert int(got["mw_ties"][4]) == 1900 ** 3 - 1900              # one tie group of 1900
    assert int(got["mw_2u1"][6]) == 2 * 700 * 700 and int(got["mw_2u1"][7]) == 0   # U of the baseline: all pairs / none
    b.free()


@pytest.mark.parametrize("alg", [abi.ALG_EWMA, abi.ALG_DES, abi.ALG_MA_ALL])
def test_forecast_band_filter_points_on_the_edge(ctx, oracle, alg):
    """In chunks that lie inside the current window k_band proves most points inside their EWMA / DES band from
    |x - forecast| alone and takes the oracle's arithmetic for the rest (fm_band.cu band_clear).  Here the 41st current
    point of every unit -- inside such a chunk -- is moved ONTO its band edge as the oracle computes it (the float
    next to it on either side, the upper or the lower edge), for noisy, integer-valued, offset and nearly constant
    series and tight to loose thresholds: counts and first indices must stay bit-identical."""
    rng = np.random.default_rng(300 + alg)
    n, n_model, n_cur, j = 600, 96, 96, 40
    kinds = [lambda m: rng.normal(100, 5, m), lambda m: np.round(rng.normal(50, 2, m)), lambda m: 1e6 + rng.normal(0, 3, m),
             lambda m: rng.normal(0, 1, m) * 1e-3, lambda m: 7.0 + (rng.random(m) < 0.05) * 0.5]
    base = [kinds[i % len(kinds)](n_model + n_cur).astype(np.float32) for i in range(n)]
    sid, mts = np.arange(n) // 6, rng.integers(0, 8, n)
    for k, bound in ((0.5, abi.BOUND_BOTH), (2.0, abi.BOUND_UPPER), (3.0, abi.BOUND_BOTH), (1.5, abi.BOUND_LOWER)):
        p = full_params(oracle, abi.ALG_MA_ALL | abi.ALG_EWMA | abi.ALG_DES, abi.PW_NONE)
        for mt in range(9):
            p.rules[mt].threshold, p.rules[mt].bound, p.rules[mt].min_lower_bound = k, bound, (0.0 if mt % 2 else -1e30)
        h = abi.HostCsr.from_units([([], z[:n_model], z[n_model:]) for z in base], sid, mts)
        p.band_algorithm = alg
        _, up, lo = oracle.score(h, p, want_bands=True)
        p.band_algorithm = 0
        moved = []
        for i, z in enumerate(base):
            o = int(h.offsets[i]) + n_model + j
            edge = np.float32(up[o] if i % 2 == 0 else lo[o])
            z = z.copy()
            z[n_model + j] = (edge, np.nextafter(edge, np.float32(np.inf)), np.nextafter(edge, np.float32(-np.inf)))[(i // 2) % 3]
            moved.append(z)
        b = ctx.batch_from_units([([], z[:n_model], z[n_model:]) for z in moved], service_id=sid, metric_type=mts)
        got, _ = check_batch(ctx, oracle, b, p)
        a = {abi.ALG_MA_ALL: 0, abi.ALG_EWMA: 2, abi.ALG_DES: 3}[alg]
        assert (got["n_anom"][:, a] > 0).any()
        b.free()


@pytest.mark.parametrize("window", [1, 2, 7, 30, 32, 33, 100])
def test_rolling_band_filter_adversarial(ctx, oracle, window):
    """The rolling band decides most points with a division-free filter and the rest with the oracle's exact
    arithmetic (fm_band.cu roll_clear / roll_exact).  Series built to sit where a sloppy filter would differ from
    the oracle: constant and two-valued windows (zero variance), huge offsets with tiny noise (cancellation in
    the variance), tiny and huge magnitudes, points exactly ON the band edge of a two-point window, ramps, and
    steps; every threshold from tight to loose, both bounds.  Counts and first indices must be bit-identical."""
    rng = np.random.default_rng(1000 + window)
    n_model, n_cur = 90, 150

    def unit(x):
        x = np.asarray(x, dtype=np.float32)
        return (x[:n_model], np.empty(0, np.float32), x[n_model:n_model + n_cur])
    t = np.arange(n_model + n_cur)
    series = [
        np.full(240, 5.0), np.zeros(240), np.full(240, -3.25),
        np.where(t % 2 == 0, 1.0, 3.0), np.where(t % 3 == 0, 10.0, 10.5),
        1e9 + rng.normal(0, 1, 240), 1e9 + rng.normal(0, 100, 240), -4e8 + rng.normal(0, 0.5, 240),
        rng.normal(0, 1, 240) * 1e-30, rng.normal(0, 1, 240) * 1e30, rng.normal(5, 1, 240) * 1e-38,
        t * 0.5, t * -2.0 + 7, np.where(t < 150, 1.0, 2.0), np.where(t < 130, 100.0, 100.0001),
        np.round(rng.normal(50, 2, 240)), np.round(rng.normal(50, 0.6, 240)),
        np.concatenate([np.full(120, 7.0), rng.normal(7, 1e-3, 120)]),
        np.concatenate([rng.normal(7, 1e-3, 120), np.full(120, 7.0)]),
        rng.normal(100, 5, 240), np.abs(rng.standard_cauchy(240)) * 3,
    ]
    # points exactly on the edge: after a window of two distinct values a and b the band edge with k = 1 and W = 2 is
    # mean + sd = max(a, b) exactly (in exact arithmetic); many such pairs in a row
    edge = np.tile(np.array([2.0, 4.0, 4.0, 2.0, 2.0, 4.0, 3.0, 5.0], np.float32), 30)
    series.append(edge)
    units = [unit(s) for s in series]
    b = ctx.batch_from_units(units, service_id=[i // 4 for i in range(len(units))], metric_type=[i % 9 for i in range(len(units))])
    for k, bound in ((0.0, abi.BOUND_BOTH), (0.5, abi.BOUND_BOTH), (1.0, abi.BOUND_UPPER), (1.0, abi.BOUND_BOTH), (2.0, abi.BOUND_LOWER),
                     (3.0, abi.BOUND_BOTH), (10.0, abi.BOUND_BOTH), (1e6, abi.BOUND_BOTH)):
        p = full_params(oracle, algs=abi.ALG_MA_ROLLING, pairwise=abi.PW_NONE)
        p.ma_window = window
        for mt in range(9):
            p.rules[mt].threshold, p.rules[mt].bound, p.rules[mt].min_lower_bound = k, bound, (0.0 if mt % 2 else -1e30)
        check_batch(ctx, oracle, b, p)
    b.free()

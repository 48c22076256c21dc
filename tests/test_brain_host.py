"""Host-side logic of the job-level worker (foremast_b200/brain.py): the config-string format of
foremast-service, the Prometheus matrix shape, the brain's environment block and the `reason` text
that foremast-trigger parses.  CPU only; the GPU end-to-end job is in test_gpu_brain.py."""
import html
import json
import re

import numpy as np

from foremast_b200 import _abi as abi
from foremast_b200 import brain
from oracle import oracle


def test_parse_config_roundtrip_of_service_format():
    # what convertMetricQuerys emits (foremast-service/cmd/manager/main.go:50-81)
    urls = {"error5xx": "http://prom:9090/api/v1/query_range?query=namespace_pod_http_errors_5xx%7Bpod%3D~%22a%7Cb%22%7D&start=START_TIME&end=END_TIME&step=60",
            "latency": "http://prom:9090/api/v1/query_range?query=lat&start=100&end=700&step=60"}
    cfg = brain.CONFIG_SEPARATOR.join(k + brain.KV_SEPARATOR + v for k, v in urls.items())
    assert " ||" in cfg and "== " in cfg
    got = brain.parse_config(cfg)
    assert got == urls
    u = brain.substitute_times(got["error5xx"], 1550000000, 1550001800.9)
    assert "start=1550000000&end=1550001800&step=60" in u
    assert brain.query_window(got["latency"]) == (100.0, 700.0, 60.0)
    assert brain.parse_config("") == {}


def test_parse_matrix_and_units():
    payload = {"status": "success", "data": {"resultType": "matrix", "result": [
        {"metric": {"pod": "a"}, "values": [[1000, "0.5"], [1060, "NaN"], [1120, "2"]]},
        {"metric": {"pod": "b"}, "values": [[1000, "1.5"]]}]}}
    m = brain.parse_matrix(json.dumps(payload))
    assert len(m) == 2 and m[0][1] == 1000.0 and m[0][2].dtype == np.float32
    assert np.isnan(m[0][2][1]) and m[0][2][2] == 2.0
    units, meta = brain.build_units({"cpu": [({}, 0.0, np.ones(5, np.float32))]}, {"cpu": m[:1]}, {"cpu": m})
    assert len(units) == 2 and meta[1]["labels"] == {"pod": "b"} and meta[0]["metric_type"] == abi.MT_CPU
    assert units[1][1] is units[0][1]  # fewer baseline series than current: the last one is reused
    assert [brain.metric_type_of(a) for a in ("error5xx", "http_4xx", "p99_latency", "cpu", "memory_used", "tps", "weird")] == \
        [abi.MT_ERROR5XX, abi.MT_ERROR4XX, abi.MT_LATENCY, abi.MT_CPU, abi.MT_MEMORY, abi.MT_TPS, abi.MT_OTHER]


def test_params_from_brain_yaml_env():
    """The literal env block of deploy/foremast/3_brain/foremast-brain.yaml:21-81 reproduces the defaults."""
    env = {"ML_ALGORITHM": "moving_average_all", "threshold": "2.0", "min_lower_bound": "0", "bound": "1",
           "metric_type_threshold_count": "5",
           "metric_type0": "error5xx", "threshold0": "2", "bound0": "1", "min_lower_bound0": "0",
           "metric_type1": "error4xx", "threshold1": "3", "bound1": "1", "min_lower_bound1": "0",
           "metric_type2": "latency", "threshold2": "10", "bound2": "3", "min_lower_bound2": "0",
           "metric_type3": "cpu", "threshold3": "5", "bound3": "1", "min_lower_bound3": "0",
           "metric_type4": "memory", "threshold4": "5", "bound4": "1", "min_lower_bound4": "0",
           "MIN_MANN_WHITE_DATA_POINTS": "20", "MIN_WILCOXON_DATA_POINTS": "20", "MIN_KRUSKAL_DATA_POINTS": "5"}
    blank = abi.Params()
    blank.rules[0].bound = 1
    p = brain.params_from_env(env, oracle.params_default())
    assert bytes(p) == bytes(oracle.params_default())
    q = brain.params_from_env({"ML_ALGORITHM": "holt_winters exponential_smoothing", "ML_PAIRWISE_ALGORITHM": "ANY",
                               "ML_PAIRWISE_THRESHOLD": "0.01"}, oracle.params_default())
    assert q.algorithms == abi.ALG_HW | abi.ALG_EWMA and q.pairwise == abi.PW_ANY and q.pairwise_threshold == 0.01


def test_reason_text_matches_trigger_regexes():
    pairs = brain.anomaly_pairs(1550000000.0, 60, [1, 21], [40.134, 40.466])
    assert pairs == [1550000060.0, 40.134, 1550001260.0, 40.466]
    r = brain.reason_text([{"alias": "error5xx", "tags": "{}", "pairs": pairs}])
    # foremast-trigger/pkg/foremasttrigger/trigger.go:295,309
    m = re.search(r"&quot;name&quot;\s*:\s*&quot;([\w\.]*)", r)
    assert m and m.group(1) == "error5xx"
    t = re.search(r"&quot;ts&quot;\s*:\s*\[(\d*).\d", r)
    assert t and t.group(1) == "1550000060"
    assert json.loads(html.unescape(r))["anomaly"][0]["values"] == [40.134, 40.466]  # trigger.go:363 un-escapes


def test_native_matrix_scanner_matches_json_module():
    """fm_parse_prometheus_matrix against the json module on awkward but legal responses."""
    rng = np.random.default_rng(3)
    result = []
    for i in range(7):
        n = int(rng.integers(0, 300))
        vals = rng.normal(1e3 ** rng.uniform(-2, 3), 5, n)
        pts = []
        for k, v in enumerate(vals):
            s = repr(float(v)) if k % 5 else f"{v:.6e}"
            if k % 7 == 3:   # the short decimals of counters and ratios (the scanner's exact fast path)
                s = ["1", "0", "0.5", "-3", "12.25", "153.7", "+7", "0.001", "000.125", "123456789012345", "-0.0", "5."][k % 12]
            if k % 41 == 40:
                s = ["NaN", "+Inf", "-Inf"][k % 3]
            pts.append([1550000000 + 60 * k + (0.123 if i == 3 else 0), s])
        result.append({"metric": {"__name__": "namespace_pod_x", "pod": f"p-{i}", "weird": "a\\\"b]}[{,"}, "values": pts})
    payload = {"status": "success", "warnings": ["x", {"y": [1, 2, {"z": "]"}]}],
               "data": {"resultType": "matrix", "extra": [[1, "2"]], "result": result}}
    for body in (json.dumps(payload), json.dumps(payload, indent=2), json.dumps(payload, separators=(",", ":"))):
        got = brain.parse_matrix(body.encode())
        want = brain.parse_matrix(json.loads(body))
        assert len(got) == len(want) == 7
        for (gl, gt, gv), (wl, wt, wv) in zip(got, want):
            assert gl == wl and gt == wt
            assert gv.tobytes() == wv.tobytes() or (np.array_equal(np.isnan(gv), np.isnan(wv)) and
                                                    np.array_equal(gv[~np.isnan(gv)], wv[~np.isnan(wv)]))
    assert brain.parse_matrix(b'{"status":"success","data":{"resultType":"matrix","result":[]}}') == []
    from foremast_b200._lib import ForemastGpuError
    import pytest
    for bad in (b"", b"[]", b'{"data":{"result":[{"values":[[1,2]]}]}}', b'{"status":"error"}'):
        with pytest.raises(ForemastGpuError):
            brain.parse_matrix(bad)


def test_missing_scrapes_are_seen_and_regridded():
    """query_range omits a missing scrape.  The scanner reports the holes (`reserved`), and with PROM_GRID the
    samples sit at round((ts - t0) / step): index i IS time t0 + i*step again ([10, 70, 190] used to
    parse as three uniform samples)."""
    def body(points):
        return json.dumps({"status": "success", "data": {"resultType": "matrix", "result": [
            {"metric": {"pod": "a"}, "values": [[t, str(v)] for t, v in points]}]}}).encode()
    pts = [(10, 1.0), (70, 2.0), (190, 4.0), (250, 5.0), (490, 9.0)]
    series, values = brain.parse_matrix_series(body(pts))
    assert int(series[0]["n_values"]) == 5 and float(series[0]["t0"]) == 10.0 and float(series[0]["step"]) == 60.0
    assert int(series[0]["reserved"]) == 4                      # slots 130, 310, 370, 430 have no sample
    assert list(values) == [1.0, 2.0, 4.0, 5.0, 9.0]            # plain mode keeps the samples as they came
    series, values = brain.parse_matrix_series(body(pts), abi.PROM_GRID)
    assert int(series[0]["n_values"]) == 9
    assert np.array_equal(np.isnan(values), [False, False, True, False, False, True, True, True, False])
    series, values = brain.parse_matrix_series(body(pts), abi.PROM_GRID | abi.PROM_FILL)
    assert list(values) == [1.0, 2.0, 2.0, 4.0, 5.0, 5.0, 5.0, 5.0, 9.0]
    # the first pair spans a gap: the smallest gap is the step, not the first one ...
    series, values = brain.parse_matrix_series(body([(0, 1.0), (120, 3.0), (180, 4.0), (240, 5.0)]), abi.PROM_GRID | abi.PROM_FILL)
    assert float(series[0]["step"]) == 60.0 and list(values) == [1.0, 1.0, 3.0, 4.0, 5.0]
    # ... and a step hint from the query URL settles it when every gap is a multiple
    series, values = brain.parse_matrix_series(body([(0, 1.0), (120, 3.0), (240, 5.0)]), abi.PROM_GRID, 60.0)
    assert int(series[0]["reserved"]) == 2 and int(series[0]["n_values"]) == 5
    # regular series: nothing changes, fractional timestamps included
    reg = [(1550000000.123 + 60 * i, i) for i in range(50)]
    s0, v0 = brain.parse_matrix_series(body(reg))
    s1, v1 = brain.parse_matrix_series(body(reg), abi.PROM_GRID | abi.PROM_FILL)
    assert int(s0[0]["reserved"]) == 0 and v0.tobytes() == v1.tobytes()
    # timestamps that go backwards are flagged, never re-gridded
    s2, v2 = brain.parse_matrix_series(body([(100, 1.0), (40, 2.0), (160, 3.0)]), abi.PROM_GRID)
    assert int(s2[0]["reserved"]) == 0xFFFFFFFF and list(v2) == [1.0, 2.0, 3.0]


def test_native_matrix_scanner_is_reentrant():
    """One response per thread (the worker's fan-out): same results as serially, and the calls really
    run side by side (ctypes drops the GIL)."""
    import time
    from concurrent.futures import ThreadPoolExecutor
    rng = np.random.default_rng(5)
    docs = []
    for d in range(16):
        result = [{"metric": {"pod": f"p{d}-{i}", "le": "0.5"},
                   "values": [[1550000000 + 60 * t, repr(float(v))] for t, v in enumerate(rng.normal(100, 7, 1440))]}
                  for i in range(8)]
        docs.append(json.dumps({"status": "success", "data": {"resultType": "matrix", "result": result}}).encode())
    t0 = time.perf_counter()
    serial = [brain.parse_matrix_native(d) for d in docs]
    t_serial = time.perf_counter() - t0
    with ThreadPoolExecutor(8) as ex:
        t0 = time.perf_counter()
        threaded = list(ex.map(brain.parse_matrix_native, docs))
        t_threads = time.perf_counter() - t0
    for a, b in zip(serial, threaded):
        assert len(a) == len(b) == 8
        for (la, ta, va), (lb, tb, vb) in zip(a, b):
            assert la == lb and ta == tb and va.tobytes() == vb.tobytes()
    mb = sum(len(d) for d in docs) / 1e6
    print(f"{mb:.1f} MB: serial {mb / t_serial:.0f} MB/s, 8 threads {mb / t_threads:.0f} MB/s")


def test_exporter_series_names_match_dashboard():
    """foremast-dashboard/src/config/metrics.js:20-31 names; _anomaly value = unix ts (metricReducer.js:79-80)."""
    from foremast_b200 import exporter
    txt = exporter.exposition("namespace_app_pod_http_server_requests_errors_5xx", "dev", "demo", upper=0.589, lower=0.0,
                              anomaly_ts=[1550000060, 1550001260])
    assert 'foremastbrain:namespace_app_pod_http_server_requests_errors_5xx_upper{exported_namespace="dev",app="demo"} 0.589' in txt
    assert 'foremastbrain:namespace_app_pod_http_server_requests_errors_5xx_lower{exported_namespace="dev",app="demo"} 0.0' in txt
    assert 'foremastbrain:namespace_app_pod_http_server_requests_errors_5xx_anomaly{exported_namespace="dev",app="demo"} 1550001260' in txt
    assert re.match(r"^foremastbrain.*", txt.splitlines()[1])   # custom-metrics-config-map.yaml:27
    assert exporter.exposition("m", "n", "a") == ""


def test_hpa_score_rule():
    """docs/dynamic_autoscaling.md:31-55: > 50 only for a rewarded scale-up, < 50 only for a rewarded scale-down."""
    d = lambda tps, lat: [{"metricType": "tps", "current": tps, "upper": 120.0, "lower": 80.0},  # noqa: E731
                          {"metricType": "latency", "current": lat, "upper": 300.0, "lower": 0.0}]
    assert brain.hpa_score(d(100, 100))[0] == 50            # load inside its band
    assert brain.hpa_score(d(140, 100))[0] == 50            # load up but the SLA holds: scaling up is not rewarded
    assert brain.hpa_score(d(140, 400))[0] == 75            # load up half a band width and latency out of its band
    assert brain.hpa_score(d(500, 400))[0] == 100
    assert brain.hpa_score(d(60, 100))[0] == 25             # load down, SLA holds: scale down
    assert brain.hpa_score(d(60, 400))[0] == 50             # load down but latency violated: no reward for scaling down
    assert brain.hpa_score([{"metricType": "cpu", "current": 1, "upper": 2, "lower": 0}])[0] == 50


def test_metrics_endpoint_serves_bands_and_hpa_score():
    import urllib.request
    from foremast_b200 import exporter
    srv = exporter.MetricsServer(port=0)
    try:
        srv.set_band("dev", "demo", "namespace_app_pod_http_server_requests_latency", upper=310.5, lower=0.0, anomaly_ts=[1550000060])
        srv.set_hpa_score("dev", "demo", 75)
        page = urllib.request.urlopen(f"http://127.0.0.1:{srv.port}/metrics", timeout=5).read().decode()
        assert 'foremastbrain:namespace_app_pod_http_server_requests_latency_upper{exported_namespace="dev",app="demo"} 310.5' in page
        assert 'foremastbrain:namespace_app_pod_http_server_requests_latency_anomaly{exported_namespace="dev",app="demo"} 1550000060' in page
        assert 'foremastbrain:namespace_app_per_pod:hpa_score{exported_namespace="dev",app="demo"} 75' in page
        import re
        assert all(re.match(r"^(#|foremastbrain[:\w]*\{)", ln) for ln in page.strip().split("\n"))  # adapter: ^foremastbrain.*
    finally:
        srv.close()


def test_regridding_fuzz_against_python():
    """Random scrape gaps, several series per response, integer and fractional timestamps: the scanner's grid
    placement (NaN holes, forward fill) against a straightforward Python re-grid of the same points."""
    rng = np.random.default_rng(17)
    for trial in range(40):
        step = float(rng.choice([15, 30, 60, 300]))
        t0 = 1550000000 + (0.25 if trial % 3 == 0 else 0.0)
        result, want_nan, want_fill, missing = [], [], [], []
        for s in range(int(rng.integers(1, 5))):
            n = int(rng.integers(2, 120))
            keep = rng.random(n) > rng.choice([0.0, 0.1, 0.4])
            keep[0] = keep[-1] = True
            if keep.sum() < 2:
                keep[1] = True
            # make sure two neighbours survive somewhere, so that the smallest gap IS the step (no hint given)
            j = int(rng.integers(0, n - 1))
            keep[j] = keep[j + 1] = True
            vals = np.round(rng.normal(10, 3, n), 3).astype(np.float32)
            result.append({"metric": {"pod": f"p{s}"}, "values": [[t0 + step * i, repr(float(vals[i]))] for i in range(n) if keep[i]]})
            g = np.where(keep, vals, np.nan).astype(np.float32)
            f = g.copy()
            for i in range(1, n):
                if np.isnan(f[i]):
                    f[i] = f[i - 1]
            want_nan.append(g)
            want_fill.append(f)
            missing.append(int((~keep).sum()))
        body = json.dumps({"status": "success", "data": {"resultType": "matrix", "result": result}}).encode()
        for flags, want in ((abi.PROM_GRID, want_nan), (abi.PROM_GRID | abi.PROM_FILL, want_fill)):
            for hint in (0.0, step):
                series, values = brain.parse_matrix_series(body, flags, hint)
                assert len(series) == len(want)
                for rec, w, miss in zip(series, want, missing):
                    got = values[int(rec["first_value"]):int(rec["first_value"]) + int(rec["n_values"])]
                    assert int(rec["reserved"]) == miss and float(rec["step"]) == step
                    assert got.shape == w.shape and np.array_equal(np.isnan(got), np.isnan(w))
                    assert np.array_equal(got[~np.isnan(got)], w[~np.isnan(w)])

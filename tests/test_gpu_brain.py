"""End-to-end canary job through the job-level worker on the GPU: the reference demo profile
(data1.txt healthy, data2.txt with two 40.x spikes; examples/demo/continuous/demo_v2.yaml:33)."""
import html
import json
import os
import re

import numpy as np
import pytest

from foremast_b200 import _abi as abi
from foremast_b200 import brain

pytestmark = pytest.mark.gpu


def matrix(series, t0=1550000000, step=60):
    return {"status": "success", "data": {"resultType": "matrix", "result": [
        {"metric": {"pod": f"pod-{i}"}, "values": [[t0 + k * step, repr(float(v))] for k, v in enumerate(s)]}
        for i, s in enumerate(series)]}}


def test_demo_job_goes_unhealthy_then_healthy(golden_dir):
    from foremast_b200.api import Context, default_params
    fx = json.load(open(os.path.join(golden_dir, "demo_fixture.json")))
    calm = [min(v, 0.5) for v in fx["data1"]]  # data1 itself peaks at 0.666, above its own 2-sigma band
    store = {"http://prom/base5xx": matrix([fx["data1"]]), "http://prom/cur5xx": matrix([fx["data2"], calm]),
             "http://prom/basecpu": matrix([np.full(40, 0.4)]), "http://prom/curcpu": matrix([np.full(40, 0.4)]),
             "http://prom/ok5xx": matrix([calm[::-1]])}
    doc = {"id": "job-1", "appName": "demo", "strategy": "canary", "status": "initial",
           "baselineConfig": "error5xx== http://prom/base5xx || cpu== http://prom/basecpu",
           "currentConfig": "error5xx== http://prom/cur5xx || cpu== http://prom/curcpu",
           "startTimeUnix": 1550000000, "endTimeUnix": 1550002520}
    with Context(0) as ctx:
        w = brain.BrainWorker(ctx, lambda url: store[url], default_params(), clock=lambda: 1550003000.0)
        out = w.process(doc)
        assert out["status"] == "completed_unhealth"   # converter.go:19 -> "anomaly" -> barrelman Unhealthy
        m = re.search(r"&quot;name&quot;\s*:\s*&quot;([\w\.]*)", out["reason"])
        t = re.search(r"&quot;ts&quot;\s*:\s*\[(\d*).\d", out["reason"])
        assert m.group(1) == "error5xx" and int(t.group(1)) == 1550000000 + 60 * 1
        info = json.loads(out["anomalyInfo"])["error5xx"]["values"]
        assert info[0::2] == [1550000060.0, 1550001260.0]            # spikes at idx 1 and 21
        assert np.allclose(info[1::2], [40.134, 40.466], rtol=1e-6)
        assert json.loads(html.unescape(out["reason"]))["anomaly"][0]["name"] == "error5xx"
        # healthy current window; before endTime the job stays in progress, after it completes
        doc["currentConfig"] = "error5xx== http://prom/ok5xx || cpu== http://prom/curcpu"
        assert w.process(doc)["status"] == "completed_health"
        w.clock = lambda: 1550001000.0
        assert w.process(doc)["status"] == "postprocess_inprogress"
        # no current metric -> unknown (docs/guides/design.md:25)
        doc["currentConfig"] = ""
        assert w.process(doc)["status"] == "completed_unknown"
        # broken store -> preprocess_failed
        doc["currentConfig"] = "cpu== http://prom/missing"
        assert w.process(doc)["status"] == "preprocess_failed"


def body(series, t0=1550000000, step=60, drop=()):
    """query_range response body (bytes); sample indices in `drop` are missing scrapes (simply absent)."""
    return json.dumps({"status": "success", "data": {"resultType": "matrix", "result": [
        {"metric": {"pod": f"pod-{i}"}, "values": [[t0 + k * step, repr(float(v))] for k, v in enumerate(s) if k not in drop]}
        for i, s in enumerate(series)]}}).encode()


def test_response_bodies_straight_into_a_batch():
    """fm_batch_add_prometheus_many: the three bodies of every job alias become units with the worker's pairing,
    parsed on a thread pool, identical to the per-series path (fm_batch_add_series) on the same data; missing
    scrapes are re-gridded (filled) so that indices stay times; a malformed body fails only its own job."""
    from foremast_b200.api import Context
    rng = np.random.default_rng(8)
    jobs, want_units, want_meta = [], [], []
    for j in range(23):
        nb, nc = int(rng.integers(1, 4)), int(rng.integers(1, 5))
        hist = rng.normal(50, 5, int(rng.integers(0, 400))).astype(np.float32)
        base = [rng.normal(50, 5, int(rng.integers(5, 80))).astype(np.float32) for _ in range(nb)]
        cur = [rng.normal(50, 5, int(rng.integers(5, 80))).astype(np.float32) for _ in range(nc)]
        jobs.append((j, 1 + j % 8, 60.0, body([hist]) if hist.size else None, body(base), body(cur)))
        for i, c in enumerate(cur):
            want_units.append((hist, base[min(i, nb - 1)], c))
            want_meta.append((j, 1 + j % 8))
    with Context(0) as ctx:
        b, per = ctx.batch_from_prometheus(jobs, threads=4)
        ref = ctx.batch_from_units(want_units, service_id=[m[0] for m in want_meta], metric_type=[m[1] for m in want_meta])
        hb, hr = b.download(), ref.download()
        assert int(per.sum()) == len(want_units)
        for f in ("values", "offsets", "seg_lens", "service_id", "metric_type"):
            assert np.array_equal(getattr(hb, f), getattr(hr, f)), f
        b.free()
        ref.free()
        # a missing scrape: sample 3 of the current window is absent -> with GRID|FILL its slot holds sample 2 again
        cur = np.arange(10, dtype=np.float32)
        b, per = ctx.batch_from_prometheus([(0, 1, 60.0, None, body([cur]), body([cur], drop={3}))])
        seg = b.download().segments(0)
        assert list(seg[2]) == [0, 1, 2, 2, 4, 5, 6, 7, 8, 9]
        b.free()
        b, per = ctx.batch_from_prometheus([(0, 1, 60.0, None, body([cur]), body([cur], drop={3}))], flags=0)
        assert list(b.download().segments(0)[2]) == [0, 1, 2, 4, 5, 6, 7, 8, 9]  # plain mode: samples as they came
        b.free()
        # one broken body among good ones
        b, per = ctx.batch_from_prometheus([(0, 1, 60.0, None, body([cur]), body([cur])),
                                            (1, 1, 60.0, None, b'{"data":{"result":[{"values":[[1,2]]}]}}', body([cur])),
                                            (2, 1, 60.0, None, body([cur]), body([cur, cur]))])
        assert list(per) == [1, -1, 2] and b.info()["n_units"] == 3
        b.free()


def test_worker_batches_jobs_and_owns_the_state_machine():
    """500 mixed job documents through ONE scoring pass give the same per-job status / reason / anomalyInfo as
    500 single calls; run_once drives the store through the brain's status transitions
    (converter.go:12-25), takes over stuck jobs after MAX_STUCK_IN_SECONDS and writes hpalogs for hpa jobs."""
    from foremast_b200.api import Context, default_params
    rng = np.random.default_rng(21)
    store_bodies, docs = {}, []
    t_end = 1550000000 + 60 * 40
    for j in range(500):
        kind = j % 5
        n_pods = 1 + j % 3
        base = [rng.normal(100, 4, 40) for _ in range(n_pods)]
        cur = [rng.normal(100, 4, 40) for _ in range(n_pods)]
        if kind == 1:
            cur[0][7] = 400.0                       # a spike -> unhealthy
        if kind == 2:
            cur = [c + 30 for c in cur]             # a level shift
        lat_b, lat_c = [rng.normal(200, 10, 40)], [rng.normal(200, 10, 40)]
        if kind == 4:
            lat_c[0][-1] = 900.0
        for name, series in (("b_tps", base), ("c_tps", cur), ("b_lat", lat_b), ("c_lat", lat_c)):
            store_bodies[f"http://prom/{j}/{name}"] = body(series)
        doc = {"id": f"job-{j}", "appName": f"app{j}", "strategy": ["canary", "rollover", "continuous", "canary", "hpa"][kind],
               "status": "initial", "startTimeUnix": 1550000000, "endTimeUnix": t_end if kind != 3 else t_end + 10_000,
               "baselineConfig": f"tps== http://prom/{j}/b_tps || latency== http://prom/{j}/b_lat",
               "currentConfig": f"tps== http://prom/{j}/c_tps || latency== http://prom/{j}/c_lat"}
        if j == 77:
            doc["currentConfig"] = ""               # no current metric -> unknown
        if j == 78:
            doc["currentConfig"] = "tps== http://prom/missing"   # fetch fails
        docs.append(doc)
    p = default_params()
    p.rules[abi.MT_TPS].threshold, p.rules[abi.MT_TPS].bound = 4.0, abi.BOUND_BOTH
    with Context(0) as ctx:
        w = brain.BrainWorker(ctx, lambda url: store_bodies[url], p, clock=lambda: float(t_end + 100))
        many = w.process_many(docs)
        singles = [w.process(d) for d in docs]
        assert len(many) == 500
        for a, b_ in zip(many, singles):
            for k in ("status", "reason", "anomalyInfo"):
                assert a.get(k) == b_.get(k), k
        st = [m["status"] for m in many]
        assert st[77] == "completed_unknown" and st[78] == "preprocess_failed"
        assert st.count("completed_unhealth") >= 150 and st.count("completed_health") >= 60
        assert st.count("postprocess_inprogress") > 0          # kind 3: endTime not reached yet
        assert all("hpalog" in m for j, m in enumerate(many) if j % 5 == 4 and j != 78)
        hl = many[4]["hpalog"]["hpalog"]
        assert 0 <= hl["hpascore"] <= 100 and {d["metricType"] for d in hl["details"]} == {"tps", "latency"}
        # ---- the store loop
        store = brain.DictStore(docs[:50])
        store.update("job-3", {"status": "preprocess_inprogress", "modified_at": float(t_end + 100 - 30)})    # someone else, recently
        store.update("job-8", {"status": "preprocess_inprogress", "modified_at": float(t_end + 100 - 500)})   # stuck
        assert w.run_once(store) == 49
        assert store.docs["job-3"]["status"] == "preprocess_inprogress"                  # left alone
        assert store.docs["job-8"]["status"] == many[8]["status"]                        # taken over and finished
        for j in range(50):
            if j != 3:
                assert store.docs[f"job-{j}"]["status"] == many[j]["status"], j
        assert len(store.hpalogs) == 10 and store.hpalogs[0]["job_id"] == "job-4"
        # postprocess_inprogress jobs are picked up again on the next cycle, finished ones are not
        again = w.run_once(store)
        assert again == sum(1 for j in range(50) if j != 3 and many[j]["status"] == "postprocess_inprogress")


def test_model_cache_is_keyed_and_capped():
    """MAX_CACHE_SIZE: models are found by key, a hit fits nothing, the least recently used one is evicted, and a
    cached model keeps scoring like a fresh pass over history + window (same flags as the oracle)."""
    from foremast_b200.api import Context
    from oracle import oracle
    p = oracle.params_default()
    p.algorithms, p.pairwise, p.period = abi.ALG_HW, abi.PW_NONE, 64
    with Context(0) as ctx:
        cache = brain.ModelCache(ctx, p, max_size=2)
        jobs = {}
        for j in range(3):
            full = oracle.synth(40, 64 * 5, 0, 50, flags=abi.SYNTH_ANOMALIES, seed=100 + j, units_per_service=8)
            jobs[f"job-{j}"] = (full, [full.segments(u) for u in range(40)])
        fits = []

        def history(key):
            def make():
                fits.append(key)
                full, units = jobs[key]
                return ctx.batch_from_units([(h, [], []) for h, _, _ in units], full.service_id, full.metric_type)
            return make

        def window(key, lo, hi):
            full, units = jobs[key]
            return ctx.batch_from_units([([], [], c[lo:hi]) for _, _, c in units], full.service_id, full.metric_type)

        got = {}
        for key, lo, hi in (("job-0", 0, 20), ("job-1", 0, 20), ("job-0", 20, 50), ("job-2", 0, 50), ("job-1", 20, 50)):
            b = window(key, lo, hi)
            r = cache.score(key, history(key), b)
            got.setdefault(key, []).append(r.series()["n_anom"][:, 4].copy())
            r.free()
            b.free()
        # job-0 hit its model on the second cycle; job-2 evicted job-1 (least recently used), so job-1 was fitted twice
        assert fits == ["job-0", "job-1", "job-2", "job-1"]
        assert (cache.hits, cache.misses, cache.evictions, len(cache)) == (1, 4, 2, 2)
        assert "job-1" in cache and "job-2" in cache and "job-0" not in cache
        # two cached cycles of job-0 = one oracle pass over history + both windows
        full, _ = jobs["job-0"]
        ref = oracle.score(full, p)[0]
        fitted = ((ref["status"] >> (abi.S_ALG_SHIFT + 4)) & 1) == 1
        assert fitted.all() and np.array_equal((got["job-0"][0] + got["job-0"][1])[fitted], ref["n_anom"][fitted, 4])
        cache.clear()
        assert len(cache) == 0

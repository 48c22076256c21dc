"""Job-level mirror of the foremast-brain pass, on top of libforemast_gpu.

This is the Python twin of the worker INTEGRATION.md describes for foremast-service: it takes one
job document in the shape foremast-service stores in Elasticsearch
(foremast-service/pkg/models/models.go:102-124), fetches the three range queries, scores them on
the GPU and returns the fields the brain writes back: `status`, `statusCode`, `reason`,
`anomalyInfo` (status vocabulary: foremast-service/pkg/converter/converter.go:10-28).

Nothing here scores on the CPU: series go into a Context (C ABI) and verdicts come back.
"""
import html
import json
import time
from urllib.parse import parse_qs, urlparse

import numpy as np

from . import _abi as abi

# foremast-service/cmd/manager/main.go:29,32
CONFIG_SEPARATOR = " ||"
KV_SEPARATOR = "== "
STEP_SECONDS = 60  # foremast-barrelman/pkg/client/metrics/metricsquery.go:63

# ML_ALGORITHM values (AI_MODEL names used by the brain; docs/guides/design.md:68-71)
ALGORITHMS = {
    "moving_average_all": abi.ALG_MA_ALL, "moving_average": abi.ALG_MA_ROLLING,
    "exponential_smoothing": abi.ALG_EWMA, "double_exponential_smoothing": abi.ALG_DES,
    "holt_winters": abi.ALG_HW, "holtwinters": abi.ALG_HW,
}


def parse_config(config):
    """'alias== url || alias== url' -> {alias: url}  (convertMetricQuerys, main.go:50-81)."""
    out = {}
    if not config:
        return out
    for part in config.split(CONFIG_SEPARATOR):
        part = part.strip()
        if not part:
            continue
        alias, sep, url = part.partition(KV_SEPARATOR)
        if not sep:
            raise ValueError(f"malformed metric config entry: {part!r}")
        out[alias.strip()] = url.strip()
    return out


def substitute_times(url, start, end):
    """continuous / hpa jobs carry START_TIME / END_TIME placeholders (main.go:59-63)."""
    return url.replace("START_TIME", str(int(start))).replace("END_TIME", str(int(end)))


def query_window(url):
    """(start, end, step) of a query_range URL built by prometheus.BuildURL (prometheushelper.go:13-43)."""
    q = parse_qs(urlparse(url).query)

    def num(k, default):
        try:
            return float(q[k][0])
        except (KeyError, ValueError, IndexError):
            return default
    return num("start", 0.0), num("end", 0.0), num("step", float(STEP_SECONDS))


def parse_matrix_series(body, flags=0, step_hint=0.0):
    """Response body (bytes) -> (fm_prom_series records, float32 values) through the library's scanner.
    flags = abi.PROM_GRID [| abi.PROM_FILL] places the samples on the query grid (missing scrapes become
    NaN, or the previous sample), so that sample index and time stay aligned; `reserved` of a record tells
    how many grid slots the series lacks."""
    import ctypes as C
    from ._lib import ForemastGpuError, load
    L = load()
    if isinstance(body, str):
        body = body.encode()
    ns, nv = C.c_uint32(), C.c_uint64()
    rc = L.fm_parse_prometheus_matrix_grid(body, len(body), flags, step_hint, None, 0, C.byref(ns), None, 0, C.byref(nv))
    if rc not in (abi.OK, abi.E_RANGE):
        raise ForemastGpuError(rc, "malformed Prometheus response")
    series = np.zeros(ns.value, dtype=abi.PROM_SERIES)
    values = np.empty(nv.value, np.float32)
    rc = L.fm_parse_prometheus_matrix_grid(body, len(body), flags, step_hint, series.ctypes.data, ns.value, C.byref(ns),
                                           values.ctypes.data, nv.value, C.byref(nv))
    if rc:
        raise ForemastGpuError(rc, "malformed Prometheus response")
    return series, values


def parse_matrix_native(body, flags=0, step_hint=0.0):
    """Response body (bytes) -> [(labels, t0, float32 values)] through the library's one-pass
    scanner (fm_parse_prometheus_matrix): no DOM, samples parsed straight into one float array."""
    if isinstance(body, str):
        body = body.encode()
    series, values = parse_matrix_series(body, flags, step_hint)
    out = []
    for s in series:
        labels = json.loads(body[int(s["metric_off"]):int(s["metric_off"] + s["metric_len"])]) if s["metric_len"] else {}
        v0 = int(s["first_value"])
        out.append((labels, float(s["t0"]), values[v0:v0 + int(s["n_values"])]))
    return out


def parse_matrix(payload):
    """Prometheus query_range matrix -> [(labels, t0, float32 values)], one per result (per pod).
    Shape per foremast-dashboard/src/services/ApiService.js:36-39 / metricReducer.js:67-69:
    data.result[].values = [[ts, "v"], ...].  Non-numeric samples become NaN (dropped at ingest).
    Raw response bodies take the native scanner; already-decoded dicts are walked in Python."""
    if isinstance(payload, (str, bytes)):
        return parse_matrix_native(payload)
    res = (payload or {}).get("data", {}).get("result", []) or []
    out = []
    for r in res:
        vals = r.get("values") or []
        t0 = float(vals[0][0]) if vals else 0.0
        v = np.empty(len(vals), np.float32)
        for i, pt in enumerate(vals):
            try:
                v[i] = float(pt[1])
            except (TypeError, ValueError):
                v[i] = np.nan
        out.append((r.get("metric", {}), t0, v))
    return out


def metric_type_of(alias):
    """metric_type rows of deploy/foremast/3_brain/foremast-brain.yaml:34-73, matched on the alias."""
    a = alias.lower()
    for name, code in abi.MT_NAMES.items():
        if name in a:
            return code
    if "5xx" in a:
        return abi.MT_ERROR5XX
    if "4xx" in a:
        return abi.MT_ERROR4XX
    if "mem" in a:
        return abi.MT_MEMORY
    return abi.MT_OTHER


def params_from_env(env, defaults):
    """Overrides `defaults` (fm_params) from the brain's environment block (foremast-brain.yaml:21-81,
    foremast-brain/README.md:20-38)."""
    p = defaults
    alg = env.get("ML_ALGORITHM")
    if alg:
        mask = 0
        for name in alg.replace(",", " ").split():
            if name.lower() not in ALGORITHMS:
                raise ValueError(f"unsupported ML_ALGORITHM {name!r}")
            mask |= ALGORITHMS[name.lower()]
        p.algorithms = mask
    for key, field, cast in (("threshold", "threshold", float), ("ML_THRESHOLD", "threshold", float),
                             ("min_lower_bound", "min_lower_bound", float),
                             ("bound", "bound", int), ("ML_BOUND", "bound", int)):
        if key in env:
            setattr(p.rules[abi.MT_OTHER], field, cast(env[key]))
    for i in range(int(env.get("metric_type_threshold_count", 0))):
        name = env.get(f"metric_type{i}")
        if name is None:
            continue
        mt = abi.MT_NAMES.get(name.lower())
        if mt is None:
            continue
        row = p.rules[mt]
        row.threshold = float(env.get(f"threshold{i}", row.threshold or p.rules[0].threshold))
        row.bound = int(env.get(f"bound{i}", row.bound or p.rules[0].bound))
        row.min_lower_bound = float(env.get(f"min_lower_bound{i}", row.min_lower_bound))
    pw = env.get("ML_PAIRWISE_ALGORITHM")
    if pw:
        p.pairwise = abi.PW_NAMES[pw.upper()]
    if "ML_PAIRWISE_THRESHOLD" in env:
        p.pairwise_threshold = float(env["ML_PAIRWISE_THRESHOLD"])
    for key, field in (("MIN_MANN_WHITE_DATA_POINTS", "min_mann_white"), ("MIN_WILCOXON_DATA_POINTS", "min_wilcoxon"),
                       ("MIN_KRUSKAL_DATA_POINTS", "min_kruskal"),
                       ("MIN_HISTORICAL_DATA_POINT_TO_MEASURE", "min_historical")):
        if key in env:
            setattr(p, field, int(env[key]))
    return p


def build_units(hist, base, cur):
    """Pair the three query results of every alias into scoring units.
    hist/base/cur: {alias: [(labels, t0, values), ...]}.  One unit per current series (pod); its
    baseline is the baseline series of the same rank (old pod i vs new pod i), its history the
    app-level historical series (metricsquery.go:79-99).  Returns (units, meta)."""
    units, meta = [], []
    for alias in sorted(cur):
        h = hist.get(alias) or []
        b = base.get(alias) or []
        hv = h[0][2] if h else np.empty(0, np.float32)
        for i, (labels, t0, cv) in enumerate(cur[alias]):
            bv = b[min(i, len(b) - 1)][2] if b else np.empty(0, np.float32)
            units.append((hv, bv, cv))
            meta.append({"alias": alias, "labels": labels, "t0": t0, "metric_type": metric_type_of(alias)})
    return units, meta


def anomaly_pairs(t0, step, idx, values):
    """[t0, v0, t1, v1, ...] exactly as barrelman's convertToAnomaly decodes it
    (foremast-barrelman/pkg/controller/DeploymentController.go:443-452)."""
    out = []
    for i, v in zip(idx, values):
        out += [float(t0 + int(i) * step), float(v)]
    return out


def reason_text(anomalies):
    """HTML-escaped JSON naming the anomalous metrics and timestamps.  foremast-trigger digs the
    metric name and the first timestamp out of it with
    `&quot;name&quot;\\s*:\\s*&quot;([\\w\\.]*)` and `&quot;ts&quot;\\s*:\\s*\\[(\\d*).\\d`
    (foremast-trigger/pkg/foremasttrigger/trigger.go:295,309) and un-escapes it for its TSV (:363)."""
    body = [{"name": a["alias"], "tags": a["tags"], "ts": [float(t) for t in a["pairs"][0::2]],
             "values": a["pairs"][1::2]} for a in anomalies]
    return html.escape(json.dumps({"anomaly": body}), quote=True)


# status vocabulary of the job document, foremast-service/pkg/converter/converter.go:12-25
ST_INITIAL, ST_PRE_INPROGRESS, ST_PRE_COMPLETED, ST_PRE_FAILED = "initial", "preprocess_inprogress", "preprocess_completed", "preprocess_failed"
ST_POST_INPROGRESS = "postprocess_inprogress"
MAX_STUCK_IN_SECONDS = 90  # deploy/foremast/3_brain/foremast-brain.yaml:80-81
HPA_METRIC = "namespace_app_per_pod:hpa_score"  # examples/hpa/README.MD:59 (foremastbrain:namespace_app_per_pod:hpa_score)


class DictStore:
    """Dict-backed stand-in of the job store the brain polls and writes back to: the `documents` index
    (ByStatus, "used by backend python model", foremast-service/pkg/search/elasticsearchstore.go:154-180)
    and the `hpalogs` index (GetLogs, :203-228).  The worker only needs these three calls; an
    Elasticsearch-backed store has the same shape."""

    def __init__(self, docs=()):
        self.docs = {d["id"]: dict(d) for d in docs}
        self.hpalogs = []

    def by_status(self, statuses, limit=None):
        out = [dict(d) for d in self.docs.values() if d.get("status", ST_INITIAL) in statuses]
        return out[:limit] if limit else out

    def update(self, doc_id, fields):
        self.docs[doc_id].update(fields)

    def add_hpalog(self, log):
        self.hpalogs.append(log)


def hpa_score(details, sla_metric="latency", load_metric="tps"):
    """The 0-100 score the HPA object targets at 50 (examples/hpa/README.MD:25-75), from the rule of
    docs/dynamic_autoscaling.md:31-55.  details = [{"metricType", "current", "upper", "lower"}, ...]
    (models.go:198-205).  Trend of the load metric: above its band = UP, below = DOWN, inside = none.
    Reward R(action): while the SLA metric stays inside its band, scaling DOWN is rewarded and UP is not;
    once it leaves the band, the other way round.  Score > 50 only for a rewarded UP, < 50 only for a
    rewarded DOWN, by how far the load left its band (in band widths, capped at one) [decision D18]."""
    by = {}
    for d in details:
        by.setdefault(d["metricType"], d)
    load, sla = by.get(load_metric), by.get(sla_metric)
    if load is None:
        return 50, "no load metric"
    meets_sla = sla is None or sla["current"] <= sla["upper"]
    width = max(load["upper"] - load["lower"], 1e-12)
    if load["current"] > load["upper"] and not meets_sla:
        s = min(1.0, (load["current"] - load["upper"]) / width)
        return int(round(50 + 50 * s)), f"{load_metric} above its band and {sla_metric} above its band: scale up"
    if load["current"] < load["lower"] and meets_sla:
        s = min(1.0, (load["lower"] - load["current"]) / width)
        return int(round(50 - 50 * s)), f"{load_metric} below its band and {sla_metric} within its band: scale down"
    return 50, "no rewarded scaling action"


class BrainWorker:
    """The brain's loop at the Elasticsearch seam, batched: many job documents -> ONE batch -> one scoring
    pass (service_id = job index), then per job the fields the brain writes back.

    process_many(docs) is the batch form; process(doc) the single-job form (same code, one job);
    run_once(store) claims open jobs from a store (status transitions initial -> preprocess_inprogress ->
    preprocess_completed -> postprocess_inprogress | completed_*, converter.go:12-25, stuck jobs taken over
    after MAX_STUCK_IN_SECONDS) and writes the results back.  `fetch(url)` returns a response body (bytes /
    str) or an already decoded dict.  Response bodies go into the batch through the library's scanner
    (fm_batch_add_prometheus_many): no Python per sample."""

    def __init__(self, ctx, fetch, params, step=STEP_SECONDS, clock=time.time, max_stuck_seconds=MAX_STUCK_IN_SECONDS,
                 gap_flags=abi.PROM_GRID | abi.PROM_FILL, parse_threads=0):
        self.ctx, self.fetch, self.params, self.step, self.clock = ctx, fetch, params, step, clock
        self.max_stuck_seconds, self.gap_flags, self.parse_threads = max_stuck_seconds, gap_flags, parse_threads

    # ---- fetch
    def _bodies(self, config, start, end):
        """{alias: (body bytes, step seconds)} of one config string."""
        out = {}
        for alias, url in parse_config(config).items():
            u = substitute_times(url, start, end)
            body = self.fetch(u)
            if isinstance(body, dict):
                body = json.dumps(body)
            if isinstance(body, str):
                body = body.encode()
            out[alias] = (body, query_window(u)[2])
        return out

    # ---- the batched pass
    def process_many(self, docs, store=None):
        now = self.clock()
        results = [None] * len(docs)
        jobs, owners = [], []  # one fm_prom_job per (document, alias); owners[k] = (doc index, alias, current body)
        ends = []
        for di, doc in enumerate(docs):
            start = doc.get("startTimeUnix", now - 30 * 60)
            end = doc.get("endTimeUnix", now)
            ends.append(end)
            try:
                cur = self._bodies(doc.get("currentConfig", ""), start, min(end, now))
                base = self._bodies(doc.get("baselineConfig", ""), start, end)
                hist = self._bodies(doc.get("historicalConfig", ""), now - 7 * 86400, now)
            except Exception as e:  # bad config / store error: converter.go:22-23
                results[di] = {"status": ST_PRE_FAILED, "statusCode": "500", "reason": f"metric fetch failed: {e}"}
                continue
            for alias in sorted(cur):
                cb, step = cur[alias]
                jobs.append((di, metric_type_of(alias), step or self.step, hist.get(alias, (None, 0))[0],
                             base.get(alias, (None, 0))[0], cb))
                owners.append((di, alias, cb, step or self.step))
            if store is not None:
                store.update(doc["id"], {"status": ST_PRE_COMPLETED, "modified_at": now})
        per_job = np.zeros(0, np.int64)
        batch = None
        if jobs:
            batch, per_job = self.ctx.batch_from_prometheus(jobs, self.gap_flags, self.parse_threads)
        n_units = int(np.clip(per_job, 0, None).sum())
        first = np.concatenate([[0], np.cumsum(np.clip(per_job, 0, None))]).astype(np.int64)
        for k, n in enumerate(per_job):  # a response that does not parse fails its document, not the batch
            if n < 0 and results[owners[k][0]] is None:
                results[owners[k][0]] = {"status": ST_PRE_FAILED, "statusCode": "500",
                                         "reason": f"malformed Prometheus response for {owners[k][1]}"}
        verdicts = scores = None
        anomalies = {}
        if n_units:
            batch.set_shard(0, len(docs))
            res = self.ctx.score(batch, self.params)
            verdicts, scores = res.verdicts(), res.series()
            if (verdicts["verdict"] == abi.V_UNHEALTH).any():
                for a in range(abi.N_ALG):
                    if self.params.algorithms & (1 << a):
                        anomalies[1 << a] = res.anomalies(1 << a)  # (offsets, idx, val) of the whole batch
            res.free()
        if batch is not None:
            batch.free()
        unit_owner = np.repeat(np.arange(len(owners)), np.clip(per_job, 0, None)) if len(owners) else np.zeros(0, np.int64)
        for di, doc in enumerate(docs):
            if results[di] is not None:
                continue
            mine = [k for k, o in enumerate(owners) if o[0] == di and per_job[k] > 0]
            if not mine:
                results[di] = {"status": "completed_unknown", "statusCode": "200", "reason": "no current metric"}
                continue
            v = verdicts[di]
            units = np.flatnonzero(np.isin(unit_owner, mine))
            found = []
            if int(v["verdict"]) == abi.V_UNHEALTH:
                for alg, (offsets, idx, val) in anomalies.items():
                    cnt = np.diff(offsets.astype(np.int64))
                    for u in units[cnt[units] > 0]:
                        k = int(unit_owner[u])
                        _, alias, cb, step = owners[k]
                        series, _ = parse_matrix_series(cb, self.gap_flags, step)  # labels and t0 of the flagged series only
                        srec = series[int(u - first[k])]
                        labels = json.loads(cb[int(srec["metric_off"]):int(srec["metric_off"] + srec["metric_len"])]) if srec["metric_len"] else {}
                        sl = slice(int(offsets[u]), int(offsets[u + 1]))
                        found.append({"alias": alias, "tags": json.dumps(labels, sort_keys=True), "algorithm": abi.ALG_NAMES[alg],
                                      "pairs": anomaly_pairs(float(srec["t0"]), step, idx[sl], val[sl])})
                    if found:
                        break  # the first enabled algorithm that flagged something names the reason
            status = abi.V_STATUS[int(v["verdict"])]
            if status == "completed_health" and now < ends[di]:
                status = ST_POST_INPROGRESS  # keep watching until endTime (docs/guides/design.md:43)
            out = {"status": status, "statusCode": "200", "modified_at": now,
                   "processingContent": json.dumps({"units": int(units.size),
                                                    "differs": int(((scores["status"][units] & abi.S_DIFFERS) != 0).sum())})}
            if found:
                out["reason"] = reason_text(found)
                info = {}
                for a in found:  # one entry per metric (models.go:78-81): the first anomalous series of the alias
                    info.setdefault(a["alias"], {"tags": a["tags"], "values": a["pairs"]})
                out["anomalyInfo"] = json.dumps(info)
            elif status == "completed_unknown":
                out["reason"] = "current metric is not available"
            if doc.get("strategy") == "hpa":
                out["hpalog"] = self._hpalog(doc, mine, owners, first, scores, now)
            results[di] = out
        if store is not None:
            for doc, out in zip(docs, results):
                fields = {k: v for k, v in out.items() if k != "hpalog"}
                store.update(doc["id"], fields)
                if "hpalog" in out:
                    store.add_hpalog(out["hpalog"])
        return results

    def _hpalog(self, doc, mine, owners, first, scores, now):
        """hpalogs document of an `hpa` job (models.go:193-208): per metric the latest sample of the first
        current series against its moving-average band (ML_ALGORITHM=moving_average_all, the documented
        default), and the score they give."""
        details = []
        for k in mine:
            _, alias, cb, step = owners[k]
            sc = scores[int(first[k])]
            series, values = parse_matrix_series(cb, self.gap_flags, step)
            n0 = int(series[0]["n_values"])
            vals = values[int(series[0]["first_value"]):int(series[0]["first_value"]) + n0]
            vals = vals[np.isfinite(vals)]
            if not vals.size or not int(sc["status"]) & (1 << abi.S_ALG_SHIFT):
                continue
            rule = self.params.rules[metric_type_of(alias)] if self.params.rules[metric_type_of(alias)].bound else self.params.rules[0]
            up = float(sc["ma_mean"]) + float(sc["k_eff"]) * float(sc["ma_std"])
            lo = max(float(rule.min_lower_bound), float(sc["ma_mean"]) - float(sc["k_eff"]) * float(sc["ma_std"]))
            details.append({"metricType": alias, "current": float(vals[-1]), "upper": up, "lower": lo})
        load = next((d["metricType"] for d in details if any(t in d["metricType"].lower() for t in ("tps", "count", "traffic"))), "tps")
        sla = next((d["metricType"] for d in details if "latency" in d["metricType"].lower()), "latency")
        score, why = hpa_score(details, sla_metric=sla, load_metric=load)
        return {"job_id": doc.get("id", ""), "timestamp": float(now),
                "hpalog": {"hpascore": score, "reason": why, "details": details}}

    def process(self, doc):
        return self.process_many([doc])[0]

    # ---- the poll loop at the store
    def claim(self, store, limit=None):
        """Jobs to work on now: new ones, ones still being watched (postprocess_inprogress), and ones another
        worker reserved but left untouched for max_stuck_seconds (docs/guides/design.md:39); each is
        reserved by writing preprocess_inprogress + modified_at."""
        now = self.clock()
        docs = []
        for d in store.by_status({ST_INITIAL, ST_POST_INPROGRESS, ST_PRE_INPROGRESS, ST_PRE_COMPLETED}):
            stuck = d.get("status") in (ST_PRE_INPROGRESS, ST_PRE_COMPLETED)
            if stuck and now - float(d.get("modified_at", 0)) <= self.max_stuck_seconds:
                continue  # somebody else is on it
            docs.append(d)
            if limit and len(docs) >= limit:
                break
        for d in docs:
            store.update(d["id"], {"status": ST_PRE_INPROGRESS, "modified_at": now})
        return docs

    def run_once(self, store, limit=None):
        docs = self.claim(store, limit)
        if docs:
            self.process_many(docs, store)
        return len(docs)


class ModelCache:
    """MAX_CACHE_SIZE (foremast-brain/README.md:30, "Max cached model size"): the fitted models the brain keeps
    between the cycles of continuous / hpa jobs, here Holt-Winters states in HBM (fm_model_fit), looked up by
    a key (the job id), at most `max_size` of them, the least recently used one freed first.

        cache = ModelCache(ctx, params, max_size=int(env.get("MAX_CACHE_SIZE", 64)))
        result = cache.score(job_id, make_history_batch, current_batch)

    make_history_batch() is only called on a miss (it fetches the 7-day history); the current batch must
    hold the same units in the same order as that history (fm_model_score)."""

    def __init__(self, ctx, params, max_size=64):
        import collections
        assert max_size >= 1
        self.ctx, self.params, self.max_size = ctx, params, max_size
        self._models = collections.OrderedDict()
        self.hits = self.misses = self.evictions = 0

    def __len__(self):
        return len(self._models)

    def __contains__(self, key):
        return key in self._models

    def score(self, key, make_history_batch, current_batch):
        m = self._models.get(key)
        if m is None:
            self.misses += 1
            hb = make_history_batch()
            try:
                m = self.ctx.model_fit(hb, self.params)
            finally:
                hb.free()
            self._models[key] = m
            while len(self._models) > self.max_size:
                _, old = self._models.popitem(last=False)
                old.free()
                self.evictions += 1
        else:
            self.hits += 1
            self._models.move_to_end(key)
        return m.score(current_batch)

    def drop(self, key):
        m = self._models.pop(key, None)
        if m is not None:
            m.free()

    def clear(self):
        for m in self._models.values():
            m.free()
        self._models.clear()

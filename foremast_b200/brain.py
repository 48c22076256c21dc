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


def parse_matrix_native(body):
    """Response body (bytes) -> [(labels, t0, float32 values)] through the library's one-pass
    scanner (fm_parse_prometheus_matrix): no DOM, samples parsed straight into one float array."""
    import ctypes as C
    from ._lib import ForemastGpuError, load
    L = load()
    if isinstance(body, str):
        body = body.encode()
    ns, nv = C.c_uint32(), C.c_uint64()
    rc = L.fm_parse_prometheus_matrix(body, len(body), None, 0, C.byref(ns), None, 0, C.byref(nv))
    if rc not in (abi.OK, abi.E_RANGE):
        raise ForemastGpuError(rc, "malformed Prometheus response")
    series = np.zeros(ns.value, dtype=abi.PROM_SERIES)
    values = np.empty(nv.value, np.float32)
    rc = L.fm_parse_prometheus_matrix(body, len(body), series.ctypes.data, ns.value, C.byref(ns),
                                      values.ctypes.data, nv.value, C.byref(nv))
    if rc:
        raise ForemastGpuError(rc, "malformed Prometheus response")
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


class BrainWorker:
    """process(document) -> the fields the brain writes back into the job document."""

    def __init__(self, ctx, fetch, params, step=STEP_SECONDS, clock=time.time):
        self.ctx, self.fetch, self.params, self.step, self.clock = ctx, fetch, params, step, clock

    def _fetch_all(self, config, start, end):
        out = {}
        for alias, url in parse_config(config).items():
            out[alias] = parse_matrix(self.fetch(substitute_times(url, start, end)))
        return out

    def process(self, doc):
        now = self.clock()
        start = doc.get("startTimeUnix", now - 30 * 60)
        end = doc.get("endTimeUnix", now)
        try:
            cur = self._fetch_all(doc.get("currentConfig", ""), start, min(end, now))
            base = self._fetch_all(doc.get("baselineConfig", ""), start, end)
            hist = self._fetch_all(doc.get("historicalConfig", ""), now - 7 * 86400, now)
        except Exception as e:  # bad config / store error: converter.go:22-23
            return {"status": "preprocess_failed", "statusCode": "500", "reason": f"metric fetch failed: {e}"}
        units, meta = build_units(hist, base, cur)
        if not units:
            return {"status": "completed_unknown", "statusCode": "200", "reason": "no current metric"}
        batch = self.ctx.batch_from_units(units, service_id=[0] * len(units),
                                          metric_type=[m["metric_type"] for m in meta])
        res = self.ctx.score(batch, self.params)
        verdict = res.verdicts()[0]
        scores = res.series()
        anomalies = []
        if int(verdict["verdict"]) == abi.V_UNHEALTH:
            for a in range(abi.N_ALG):
                if not self.params.algorithms & (1 << a):
                    continue
                offsets, idx, val = res.anomalies(1 << a)
                for u in np.flatnonzero(np.diff(offsets.astype(np.int64))):
                    sl = slice(int(offsets[u]), int(offsets[u + 1]))
                    m = meta[u]
                    anomalies.append({"alias": m["alias"], "tags": json.dumps(m["labels"], sort_keys=True),
                                      "algorithm": abi.ALG_NAMES[1 << a],
                                      "pairs": anomaly_pairs(m["t0"], self.step, idx[sl], val[sl])})
                if anomalies:
                    break  # the first enabled algorithm that flagged something names the reason
        res.free()
        batch.free()
        status = abi.V_STATUS[int(verdict["verdict"])]
        if status == "completed_health" and now < end:
            status = "postprocess_inprogress"  # keep watching until endTime (docs/guides/design.md:43)
        out = {"status": status, "statusCode": "200", "modified_at": now,
               "processingContent": json.dumps({"units": len(units), "differs": int(((scores["status"] & abi.S_DIFFERS) != 0).sum())})}
        if anomalies:
            out["reason"] = reason_text(anomalies)
            info = {}
            for a in anomalies:  # one entry per metric (models.go:78-81): the first anomalous series of the alias
                info.setdefault(a["alias"], {"tags": a["tags"], "values": a["pairs"]})
            out["anomalyInfo"] = json.dumps(info)
        elif status == "completed_unknown":
            out["reason"] = "current metric is not available"
        return out

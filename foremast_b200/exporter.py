"""Prometheus exposition of the brain's output series.

The brain serves `foremastbrain:<metric>_upper`, `_lower` and `_anomaly` gauges on :8000/metrics
(deploy/foremast/3_brain/foremast-brain.yaml:87-89,119-122); the dashboard reads them by exactly
these names with the label `exported_namespace` (foremast-dashboard/src/config/metrics.js:20-31)
and treats the VALUE of an `_anomaly` sample as the unix timestamp of the anomalous point
(src/reducers/metricReducer.js:79-80); the custom-metrics adapter matches `^foremastbrain.*` with
`exported_namespace` and `app` (deploy/custom-metrics/custom-metrics-config-map.yaml:27-35).
"""


def _labels(namespace, app, extra=None):
    items = {"exported_namespace": namespace, "app": app}
    items.update(extra or {})
    return ",".join('%s="%s"' % (k, str(v).replace("\\", "\\\\").replace('"', '\\"')) for k, v in items.items())


def exposition(metric, namespace, app, upper=None, lower=None, anomaly_ts=(), extra_labels=None):
    """Text-format lines for one (app, metric): latest upper / lower band and anomaly timestamps."""
    lab = _labels(namespace, app, extra_labels)
    base = "foremastbrain:" + metric
    lines = []
    if upper is not None:
        lines += [f"# TYPE {base}_upper gauge", f"{base}_upper{{{lab}}} {float(upper)!r}"]
    if lower is not None:
        lines += [f"# TYPE {base}_lower gauge", f"{base}_lower{{{lab}}} {float(lower)!r}"]
    if len(anomaly_ts):
        lines.append(f"# TYPE {base}_anomaly gauge")
        # one sample per scrape: the most recent anomalous point (its value IS the timestamp)
        lines.append(f"{base}_anomaly{{{lab}}} {int(max(anomaly_ts))}")
    return "\n".join(lines) + ("\n" if lines else "")


def from_result(result, unit, metric, namespace, app, t0, step=60, algorithm=None):
    """Exposition for one scored unit: last band values of `params.band_algorithm` and the
    timestamps of its anomalous points (`algorithm` defaults to the band algorithm)."""
    up, lo = result.bands(unit)
    idx = ()
    if algorithm is not None:
        idx, _ = result.anomalies_unit(unit)
    ts = [t0 + int(i) * step for i in idx]
    return exposition(metric, namespace, app, up[-1] if up.size else None, lo[-1] if lo.size else None, ts)


def hpa_exposition(namespace, app, score):
    """The gauge the HPA object reads through the custom-metrics adapter:
    `foremastbrain:namespace_app_per_pod:hpa_score` (examples/hpa/README.MD:59), labels as above."""
    name = "foremastbrain:namespace_app_per_pod:hpa_score"
    return f"# TYPE {name} gauge\n{name}{{{_labels(namespace, app)}}} {int(score)}\n"


class MetricsServer:
    """The brain's `:8000/metrics` endpoint (deploy/foremast/3_brain/foremast-brain.yaml:87-89, scraped by the
    ServiceMonitor at :119-122): a registry of exposition blocks, one per (namespace, app, metric), that the
    worker overwrites after every scoring cycle, served as one text page.  Standard library only."""

    def __init__(self, host="127.0.0.1", port=8000):
        import threading
        from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
        self._blocks, self._lock = {}, threading.Lock()
        outer = self

        class Handler(BaseHTTPRequestHandler):
            def do_GET(self):  # noqa: N802
                if self.path.split("?")[0] != "/metrics":
                    self.send_error(404)
                    return
                body = outer.render().encode()
                self.send_response(200)
                self.send_header("Content-Type", "text/plain; version=0.0.4")
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

            def log_message(self, *a):
                pass

        self._httpd = ThreadingHTTPServer((host, port), Handler)
        self.port = self._httpd.server_address[1]
        self._thread = threading.Thread(target=self._httpd.serve_forever, daemon=True)
        self._thread.start()

    def set(self, namespace, app, metric, text):
        with self._lock:
            self._blocks[(namespace, app, metric)] = text

    def set_band(self, namespace, app, metric, upper=None, lower=None, anomaly_ts=()):
        self.set(namespace, app, metric, exposition(metric, namespace, app, upper, lower, anomaly_ts))

    def set_hpa_score(self, namespace, app, score):
        self.set(namespace, app, "hpa_score", hpa_exposition(namespace, app, score))

    def render(self):
        with self._lock:
            return "".join(self._blocks[k] for k in sorted(self._blocks))

    def close(self):
        self._httpd.shutdown()
        self._httpd.server_close()

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

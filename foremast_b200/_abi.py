"""ctypes / numpy mirrors of the POD types in include/foremast_gpu.h.

Shared by the product binding (foremast_b200._lib) and the test-only oracle binding
(oracle/oracle.py) so that both sides are fed byte-identical parameter blocks.
"""
import ctypes as C

import numpy as np

ABI_VERSION = 1

# error codes
OK, E_INVALID, E_NOMEM, E_CUDA, E_NO_DEVICE, E_ALIGN, E_STATE, E_RANGE, E_UNSUPPORTED, E_COMM, E_TOO_LARGE = (
    0, -1, -2, -3, -4, -5, -6, -7, -8, -9, -10)
SCAN_STATIC_MAX_COUNTS = 16777216

# ML_ALGORITHM bits (deploy/foremast/3_brain/foremast-brain.yaml:24-25, docs/guides/design.md:68-71)
ALG_MA_ALL, ALG_MA_ROLLING, ALG_EWMA, ALG_DES, ALG_HW = 1, 2, 4, 8, 16
N_ALG = 5
ALG_ALL = 0x1F
ALG_NAMES = {ALG_MA_ALL: "moving_average_all", ALG_MA_ROLLING: "moving_average",
             ALG_EWMA: "exponential_smoothing", ALG_DES: "double_exponential_smoothing",
             ALG_HW: "holt_winters"}

# ML_PAIRWISE_ALGORITHM (foremast-brain/README.md:34)
PW_NONE, PW_ALL, PW_ANY, PW_MANN_WHITE, PW_WILCOXON, PW_KRUSKAL = 0, 1, 2, 3, 4, 5
PW_NAMES = {"NONE": PW_NONE, "ALL": PW_ALL, "ANY": PW_ANY, "MANN_WHITE": PW_MANN_WHITE,
            "WILCOXON": PW_WILCOXON, "KRUSKAL": PW_KRUSKAL}

BOUND_UPPER, BOUND_LOWER, BOUND_BOTH = 1, 2, 3

MT_OTHER, MT_ERROR5XX, MT_ERROR4XX, MT_LATENCY, MT_CPU, MT_MEMORY, MT_TPS, MT_GC, MT_THREADS = range(9)
MAX_METRIC_TYPES = 16
# metric_type names as they appear in foremast-brain.yaml:34-73
MT_NAMES = {"error5xx": MT_ERROR5XX, "error4xx": MT_ERROR4XX, "latency": MT_LATENCY, "cpu": MT_CPU,
            "memory": MT_MEMORY, "tps": MT_TPS, "gc": MT_GC, "threads": MT_THREADS}

COMM_ID_BYTES = 128
V_HEALTH, V_UNKNOWN, V_UNHEALTH = 0, 1, 2
# the statuses the brain writes into the ES document (foremast-service/pkg/converter/converter.go:16-21)
V_STATUS = {V_HEALTH: "completed_health", V_UNKNOWN: "completed_unknown", V_UNHEALTH: "completed_unhealth"}

S_UNKNOWN, S_ANOMALY, S_DIFFERS, S_MW_RUN, S_WILCOXON_RUN, S_KRUSKAL_RUN, S_MODEL_BASE = (
    1, 2, 4, 8, 16, 32, 64)
S_ALG_SHIFT = 8

SYNTH_RAGGED, SYNTH_ANOMALIES = 1, 2

NONE_IDX = 0xFFFFFFFF


class MetricRule(C.Structure):
    _fields_ = [("threshold", C.c_double), ("min_lower_bound", C.c_double),
                ("bound", C.c_uint32), ("reserved", C.c_uint32)]


class Params(C.Structure):
    _fields_ = [("algorithms", C.c_uint32), ("pairwise", C.c_uint32),
                ("pairwise_threshold", C.c_double), ("differ_scale", C.c_double),
                ("min_mann_white", C.c_uint32), ("min_wilcoxon", C.c_uint32),
                ("min_kruskal", C.c_uint32), ("min_historical", C.c_uint32),
                ("min_anomaly_points", C.c_uint32), ("ma_window", C.c_uint32),
                ("period", C.c_uint32), ("band_algorithm", C.c_uint32),
                ("alpha", C.c_double), ("beta", C.c_double), ("gamma", C.c_double),
                ("rules", MetricRule * MAX_METRIC_TYPES)]

    def copy(self):
        q = Params()
        C.memmove(C.byref(q), C.byref(self), C.sizeof(Params))
        return q


class Csr(C.Structure):
    _fields_ = [("n_units", C.c_uint64), ("values", C.c_void_p), ("offsets", C.c_void_p),
                ("seg_lens", C.c_void_p), ("service_id", C.c_void_p), ("metric_type", C.c_void_p)]


class SeriesMeta(C.Structure):
    _fields_ = [("service_id", C.c_uint32), ("metric_type", C.c_uint8), ("reserved", C.c_uint8 * 3)]


class SynthConfig(C.Structure):
    _fields_ = [("n_units", C.c_uint64), ("n_hist", C.c_uint32), ("n_base", C.c_uint32),
                ("n_cur", C.c_uint32), ("units_per_service", C.c_uint32),
                ("n_metric_types", C.c_uint32), ("seed", C.c_uint32), ("flags", C.c_uint32),
                ("_pad", C.c_uint32), ("first_unit", C.c_uint64)]


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("flags", C.c_uint32)]


SERIES_SCORE = np.dtype([
    ("status", "<u4"), ("n_anom", "<u4", (N_ALG,)), ("first_idx", "<u4", (N_ALG,)), ("w_n", "<u4"),
    ("mw_2u1", "<i8"), ("mw_2r1", "<i8"), ("mw_ties", "<i8"),
    ("w_2wplus", "<i8"), ("w_2wminus", "<i8"), ("w_ties", "<i8"),
    ("p_mw", "<f8"), ("p_w", "<f8"), ("p_k", "<f8"), ("kw_h", "<f8"), ("k_eff", "<f8"),
    ("ma_mean", "<f8"), ("ma_std", "<f8"), ("sigma", "<f8", (N_ALG,)),
], align=True)
assert SERIES_SCORE.itemsize == 192

VERDICT = np.dtype([
    ("service_id", "<u4"), ("verdict", "<u4"), ("n_units", "<u4"), ("n_unknown", "<u4"),
    ("n_anom_units", "<u4"), ("n_anom_points", "<u4"), ("worst_unit", "<u4"),
    ("worst_unit_points", "<u4"),
])
assert VERDICT.itemsize == 32
assert C.sizeof(Params) == 464
assert C.sizeof(SynthConfig) == 48


PROM_SERIES = np.dtype([("metric_off", "<u8"), ("metric_len", "<u8"), ("t0", "<f8"), ("step", "<f8"),
                        ("first_value", "<u8"), ("n_values", "<u4"), ("reserved", "<u4")])
assert PROM_SERIES.itemsize == 48


PROM_GRID, PROM_FILL = 1, 2


class PromJob(C.Structure):
    """fm_prom_job: the three query_range response bodies of one job alias."""
    _fields_ = [("meta", SeriesMeta), ("reserved", C.c_uint32), ("step", C.c_double),
                ("hist_json", C.c_char_p), ("base_json", C.c_char_p), ("cur_json", C.c_char_p),
                ("hist_len", C.c_uint64), ("base_len", C.c_uint64), ("cur_len", C.c_uint64)]


class HostCsr:
    """A host-side ragged CSR batch (numpy arrays) in the layout of `fm_csr`."""

    def __init__(self, values, offsets, seg_lens, service_id=None, metric_type=None):
        self.values = np.ascontiguousarray(values, dtype=np.float32)
        self.offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        self.seg_lens = np.ascontiguousarray(seg_lens, dtype=np.uint32).reshape(-1, 3)
        n = self.seg_lens.shape[0]
        assert self.offsets.shape[0] == n + 1
        self.service_id = np.ascontiguousarray(
            service_id if service_id is not None else np.zeros(n), dtype=np.uint32)
        self.metric_type = np.ascontiguousarray(
            metric_type if metric_type is not None else np.zeros(n), dtype=np.uint8)
        assert self.service_id.shape[0] == n and self.metric_type.shape[0] == n

    @property
    def n_units(self):
        return self.seg_lens.shape[0]

    @property
    def n_services(self):
        return int(self.service_id.max()) + 1 if self.n_units else 0

    def as_struct(self):
        return Csr(self.n_units, self.values.ctypes.data, self.offsets.ctypes.data,
                   self.seg_lens.ctypes.data, self.service_id.ctypes.data,
                   self.metric_type.ctypes.data)

    def segments(self, u):
        o = int(self.offsets[u])
        nh, nb, nc = (int(x) for x in self.seg_lens[u])
        v = self.values
        return v[o:o + nh], v[o + nh:o + nh + nb], v[o + nh + nb:o + nh + nb + nc]

    @classmethod
    def from_units(cls, units, service_id=None, metric_type=None, align=4):
        """units: list of (hist, base, cur) float sequences."""
        offs, lens, chunks, pos = [], [], [], 0
        for h, b, c in units:
            # same canonicalisation as the library's ingest: -0.0 -> +0.0 (DECISIONS D11)
            h, b, c = (np.asarray(x, dtype=np.float32).ravel() + np.float32(0) for x in (h, b, c))
            offs.append(pos)
            lens.append((h.size, b.size, c.size))
            n = h.size + b.size + c.size
            pad = (-n) % align
            chunks += [h, b, c, np.zeros(pad, np.float32)]
            pos += n + pad
        offs.append(pos)
        chunks.append(np.zeros(32, np.float32))
        values = np.concatenate(chunks) if chunks else np.zeros(32, np.float32)
        return cls(values, offs, np.array(lens, dtype=np.uint32).reshape(-1, 3), service_id, metric_type)

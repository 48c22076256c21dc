"""ctypes binding of the CPU oracle (oracle/libfm_oracle.so).  TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED (see oracle/DECISIONS.md): the reference's scorer source is not in the mount and
none of its tests pins a score.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from foremast_b200 import _abi as abi

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libfm_oracle.so")
_lib = None


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("fm_oracle.c", "fm_synth.c", "fm_oracle.h", "Makefile")]
    srcs.append(os.path.join(_HERE, "..", "include", "foremast_gpu.h"))
    stale = (not os.path.exists(_SO)) or any(
        os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs)
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "-B", "libfm_oracle.so"], check=True,
                       stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.fmo_params_default.argtypes = [C.POINTER(abi.Params)]
        L.fmo_params_default.restype = None
        L.fmo_score.argtypes = [C.POINTER(abi.Csr), C.POINTER(abi.Params), C.c_void_p, C.c_void_p,
                                C.c_void_p, C.c_int]
        L.fmo_verdicts.argtypes = [C.POINTER(abi.Csr), C.c_void_p, C.c_uint32, C.c_uint64, C.c_void_p]
        L.fmo_synth_stride.argtypes = [C.POINTER(abi.SynthConfig)]
        L.fmo_synth_stride.restype = C.c_uint64
        L.fmo_synth.argtypes = [C.POINTER(abi.SynthConfig)] + [C.c_void_p] * 5 + [C.c_int]
        L.fmo_max_threads.restype = C.c_int
        _lib = L
    return _lib


def params_default():
    p = abi.Params()
    lib().fmo_params_default(C.byref(p))
    return p


def max_threads():
    return int(lib().fmo_max_threads())


def score(csr, params, want_bands=False, threads=0):
    """Returns (scores[SERIES_SCORE], upper, lower); bands are indexed like csr.values."""
    scores = np.zeros(csr.n_units, dtype=abi.SERIES_SCORE)
    up = lo = None
    if want_bands:
        up = np.full(csr.values.shape, np.nan, np.float32)
        lo = np.full(csr.values.shape, np.nan, np.float32)
    st = csr.as_struct()
    rc = lib().fmo_score(C.byref(st), C.byref(params), scores.ctypes.data,
                         up.ctypes.data if want_bands else None,
                         lo.ctypes.data if want_bands else None, threads)
    if rc:
        raise RuntimeError(f"fmo_score failed: {rc}")
    return scores, up, lo


def verdicts(csr, scores, n_services, first_unit=0):
    out = np.zeros(n_services, dtype=abi.VERDICT)
    st = csr.as_struct()
    rc = lib().fmo_verdicts(C.byref(st), scores.ctypes.data, n_services, first_unit, out.ctypes.data)
    if rc:
        raise RuntimeError(f"fmo_verdicts failed: {rc}")
    return out


def synth(n_units, n_hist, n_base, n_cur, units_per_service=80, n_metric_types=8,
          seed=0xF03E5A57, flags=0, first_unit=0, threads=0):
    cfg = abi.SynthConfig(n_units, n_hist, n_base, n_cur, units_per_service, n_metric_types,
                          seed & 0xFFFFFFFF, flags, 0, first_unit)
    stride = int(lib().fmo_synth_stride(C.byref(cfg)))
    values = np.empty(n_units * stride + 32, np.float32)
    offsets = np.empty(n_units + 1, np.int64)
    seg = np.empty((n_units, 3), np.uint32)
    sid = np.empty(n_units, np.uint32)
    mt = np.empty(n_units, np.uint8)
    rc = lib().fmo_synth(C.byref(cfg), values.ctypes.data, offsets.ctypes.data, seg.ctypes.data,
                         sid.ctypes.data, mt.ctypes.data, threads)
    if rc:
        raise RuntimeError(f"fmo_synth failed: {rc}")
    return abi.HostCsr(values, offsets, seg, sid, mt)

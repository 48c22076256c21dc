"""ctypes binding of libforemast_gpu.so (include/foremast_gpu.h).

There is no CPU path: if the library is missing or no sm_100 GPU is visible the binding raises.
"""
import ctypes as C
import os

from . import _abi as abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FOREMAST_GPU_LIB", os.path.join(_HERE, "libforemast_gpu.so"))

# every symbol include/foremast_gpu.h declares: (name, restype, argtypes)
_P = C.c_void_p
_PP = C.POINTER(C.c_void_p)
SYMBOLS = {
    "fm_init": (C.c_int, [C.POINTER(abi.Config), _PP]),
    "fm_shutdown": (None, [_P]),
    "fm_version": (C.c_int, []),
    "fm_strerror": (C.c_char_p, [C.c_int]),
    "fm_last_error": (C.c_char_p, [_P]),
    "fm_last_error_copy": (C.c_int, [_P, C.c_char_p, C.c_uint32]),
    "fm_params_default": (None, [C.POINTER(abi.Params)]),
    "fm_device_info": (C.c_int, [_P, C.POINTER(C.c_int32), C.POINTER(C.c_uint64), C.c_char_p, C.c_uint32]),
    "fm_timer_start": (C.c_int, [_P]),
    "fm_timer_stop": (C.c_int, [_P, C.POINTER(C.c_float)]),
    "fm_host_alloc": (C.c_int, [_P, C.c_uint64, _PP]),
    "fm_host_free": (None, [_P, _P]),
    "fm_parse_prometheus_matrix": (C.c_int, [C.c_char_p, C.c_uint64, _P, C.c_uint32, C.POINTER(C.c_uint32), _P,
                                              C.c_uint64, C.POINTER(C.c_uint64)]),
    "fm_parse_prometheus_matrix_grid": (C.c_int, [C.c_char_p, C.c_uint64, C.c_uint32, C.c_double, _P, C.c_uint32,
                                                   C.POINTER(C.c_uint32), _P, C.c_uint64, C.POINTER(C.c_uint64)]),
    "fm_batch_add_prometheus": (C.c_int, [_P, C.POINTER(abi.PromJob), C.c_uint32, C.POINTER(C.c_uint32)]),
    "fm_batch_add_prometheus_many": (C.c_int, [_P, _P, C.c_uint32, C.c_uint32, C.c_uint32, _P]),
    "fm_batch_begin": (C.c_int, [_P, C.c_uint64, C.c_uint64, _PP]),
    "fm_batch_add_series": (C.c_int, [_P, C.POINTER(abi.SeriesMeta), _P, C.c_uint32, _P, C.c_uint32, _P, C.c_uint32]),
    "fm_batch_seal": (C.c_int, [_P]),
    "fm_batch_from_csr": (C.c_int, [_P, C.POINTER(abi.Csr), _PP]),
    "fm_batch_from_device": (C.c_int, [_P, C.POINTER(abi.Csr), _PP]),
    "fm_batch_synth": (C.c_int, [_P, C.POINTER(abi.SynthConfig), _PP]),
    "fm_batch_set_shard": (C.c_int, [_P, C.c_uint64, C.c_uint32]),
    "fm_batch_info": (C.c_int, [_P, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)]),
    "fm_batch_download": (C.c_int, [_P, _P, _P, _P, _P, _P]),
    "fm_batch_free": (None, [_P]),
    "fm_score": (C.c_int, [_P, _P, C.POINTER(abi.Params), _PP]),
    "fm_model_fit": (C.c_int, [_P, _P, C.POINTER(abi.Params), _PP]),
    "fm_model_score": (C.c_int, [_P, _P, _P, _PP]),
    "fm_model_free": (None, [_P]),
    "fm_score_host": (C.c_int, [_P, C.POINTER(abi.Csr), C.POINTER(abi.Params), C.c_uint32, _P]),
    "fm_result_sync": (C.c_int, [_P]),
    "fm_result_verdicts": (C.c_int, [_P, _P, C.c_uint64, C.POINTER(C.c_uint64)]),
    "fm_result_verdicts_range": (C.c_int, [_P, C.c_uint64, C.c_uint64, _P]),
    "fm_result_series": (C.c_int, [_P, C.c_uint64, C.c_uint64, _P]),
    "fm_result_bands": (C.c_int, [_P, C.c_uint64, _P, _P, C.c_uint32, C.POINTER(C.c_uint32)]),
    "fm_result_anomalies_build": (C.c_int, [_P, C.c_uint32, C.POINTER(C.c_uint64)]),
    "fm_result_anomalies_unit": (C.c_int, [_P, C.c_uint64, _P, _P, C.c_uint32, C.POINTER(C.c_uint32)]),
    "fm_result_anomalies_all": (C.c_int, [_P, _P, _P, _P, C.c_uint64]),
    "fm_scan_counts_static": (C.c_int, [_P, C.c_uint64, _P, _P]),
    "fm_comm_unique_id": (C.c_int, [_P, C.c_uint32]),
    "fm_comm_init": (C.c_int, [_P, _P, C.c_int32, C.c_int32]),
    "fm_comm_info": (C.c_int, [_P, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "fm_comm_destroy": (None, [_P]),
    "fm_result_merge": (C.c_int, [_P]),
    "fm_result_service_buffers": (C.c_int, [_P, _PP, _PP, C.POINTER(C.c_uint32)]),
    "fm_result_timings": (C.c_int, [_P, C.POINTER(C.c_float), C.c_uint32]),
    "fm_result_free": (None, [_P]),
}

_lib = None


class ForemastGpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libforemast_gpu error {code}: {msg}")
        self.code = code


def load():
    """Loads the shared library and binds every declared symbol; raises if anything is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} not found: build it with `python -m foremast_b200.build` "
                "(libforemast_gpu is the only scoring path; there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)  # AttributeError if the export is missing
            fn.restype = res
            fn.argtypes = args
        if L.fm_version() != abi.ABI_VERSION:
            raise ImportError("libforemast_gpu ABI version mismatch")
        _lib = L
    return _lib

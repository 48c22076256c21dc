"""CPU-only checks of the drop-in boundary: the C-ABI library loads without a GPU and exports
every symbol include/foremast_gpu.h declares; no compute calls here."""
import ctypes as C
import os
import re

import pytest

from foremast_b200 import _abi as abi
from foremast_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build.build_library()
    return _lib.load()


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "foremast_gpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fm_[a-z_0-9]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    names = declared_symbols()
    assert len(names) >= 25
    raw = C.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), f"{n} declared in foremast_gpu.h but not exported"
        assert n in _lib.SYMBOLS, f"{n} not bound in foremast_b200/_lib.py"
    assert set(_lib.SYMBOLS) == set(names)


def test_version_errors_and_defaults(lib):
    assert lib.fm_version() == abi.ABI_VERSION
    assert b"no sm_100" in lib.fm_strerror(abi.E_NO_DEVICE)
    p = abi.Params()
    lib.fm_params_default(C.byref(p))
    # deploy/foremast/3_brain/foremast-brain.yaml:24-79
    assert p.algorithms == abi.ALG_MA_ALL
    assert (p.min_mann_white, p.min_wilcoxon, p.min_kruskal) == (20, 20, 5)
    rows = {abi.MT_ERROR5XX: (2.0, 1), abi.MT_ERROR4XX: (3.0, 1), abi.MT_LATENCY: (10.0, 3),
            abi.MT_CPU: (5.0, 1), abi.MT_MEMORY: (5.0, 1), abi.MT_OTHER: (2.0, 1)}
    for mt, (thr, bound) in rows.items():
        assert p.rules[mt].threshold == thr and p.rules[mt].bound == bound and p.rules[mt].min_lower_bound == 0.0
    # the product defaults and the oracle defaults are written independently; they must agree
    from oracle import oracle
    q = oracle.params_default()
    assert bytes(p) == bytes(q)


def test_no_cpu_fallback_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    h = C.c_void_p()
    assert lib.fm_init(C.byref(abi.Config(-1, 0)), C.byref(h)) == abi.E_NO_DEVICE
    from foremast_b200.api import Context
    from foremast_b200._lib import ForemastGpuError
    with pytest.raises(ForemastGpuError):
        Context()


def test_product_does_not_reference_oracle():
    """The shipped path must not import, link or call anything under oracle/."""
    pkg = os.path.join(ROOT, "foremast_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "fmo_" not in txt and "import oracle" not in txt and "from oracle" not in txt, f

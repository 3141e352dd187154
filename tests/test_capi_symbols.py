"""The C-ABI shared libraries load here (no GPU) and export every symbol include/*.h declares;
compute entry points fail loudly without a device (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header, prefix):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(%s\w+)\s*\(" % prefix, src)))


def test_b200spmv_exports_all_declared_symbols(native):
    from spmv_research_b200 import capi
    L = ctypes.CDLL(capi.SO_PATH)
    names = _declared("b200spmv.h", "b200spmv_")
    assert len(names) >= 25
    for n in names:
        assert hasattr(L, n), "libb200spmv.so does not export %s" % n
    assert b"sm_100a" in capi.lib().b200spmv_version()


def test_b200amg_exports_all_declared_symbols(native):
    from spmv_research_b200 import amg
    L = ctypes.CDLL(amg._SO)
    for n in _declared("b200amg.h", "b200amg_"):
        assert hasattr(L, n), n


def test_sm100a_only_and_blackwell_loads(native):
    """The cubin inside the library is sm_100a and carries the cache-hinted streaming / gather loads."""
    import shutil
    import subprocess
    from spmv_research_b200 import capi
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    lst = subprocess.run([cuobjdump, "-lelf", capi.SO_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", lst))
    assert archs == {"sm_100a"}, archs
    sass = subprocess.run([cuobjdump, "-sass", capi.SO_PATH], capture_output=True, text=True).stdout
    assert "LDG.E.NA.64.CONSTANT" in sass         # L1 no-allocate streaming loads of the matrix arrays
    assert "LDG.E.EL.64.CONSTANT" in sass         # x gathers with the evict-last hint


def test_no_cpu_fallback(native):
    from spmv_research_b200 import capi
    import spmv_research_b200 as S
    if capi.device_count() > 0:
        pytest.skip("a GPU is present")
    rp = np.array([0, 1], np.int32)
    with pytest.raises(S.B200Error):
        S.csr_to_format(rp, np.zeros(1, np.int32), np.ones(1), 1, 1, 1)
    with pytest.raises(S.B200Error):
        capi.partition(rp, 1, 1, 2)


def test_symmetric_unexpanded_rejected(native):
    """Same behaviour as every general format of the reference (K/csr.cpp:224-225)."""
    import spmv_research_b200 as S
    with pytest.raises(S.B200Error, match="symmetric matrices have to be expanded"):
        S.csr_to_format(np.array([0, 1], np.int32), np.zeros(1, np.int32), np.ones(1), 1, 1, 1, symmetric=1, symmetry_expanded=0)


def test_b200mtx_exports_all_declared_symbols(native):
    from spmv_research_b200 import mtx
    L = ctypes.CDLL(mtx._SO)
    names = _declared("b200mtx.h", "b200mtx_")
    assert set(names) >= {"b200mtx_read", "b200mtx_free", "b200mtx_last_error"}
    for n in names:
        assert hasattr(L, n), n

"""Matrix Market ingest: the host reader (libb200mtx.so, include/b200mtx.h) mirrors the reference's
mtx_read + mtx_values_convert_to_real (bench.cpp:128-206); COO -> CSR then runs on the GPU (capi.coo_to_csr /
b200spmv_create_coo), mirroring coo_to_csr (bench.cpp:208-229)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "csrc", "mtx.c")
_SO = os.path.join(_HERE, "lib", "libb200mtx.so")
FIELDS = ("real", "integer", "pattern", "complex")


class MtxError(RuntimeError):
    pass


class _Info(C.Structure):
    _fields_ = [("m", C.c_int64), ("n", C.c_int64), ("nnz", C.c_int64), ("nnz_file", C.c_int64), ("nnz_diag", C.c_int64),
                ("nnz_non_diag", C.c_int64), ("symmetric", C.c_int32), ("skew_symmetry", C.c_int32),
                ("hermitian_symmetry", C.c_int32), ("field", C.c_int32)]


def build(force=False):
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(_SRC):
        os.makedirs(os.path.dirname(_SO), exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-fopenmp", "-fPIC", "-shared", "-Wall", "-Wextra", "-o", _SO, _SRC, "-lm"])
    return _SO


_lib = None


def _L():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise RuntimeError("libb200mtx.so is not built: run `python -c 'import __graft_entry__ as g; g.build()'`")
        L = C.CDLL(_SO)
        L.b200mtx_read.restype = C.c_int
        L.b200mtx_read.argtypes = [C.c_char_p, C.c_int32, C.POINTER(_Info), C.POINTER(C.POINTER(C.c_int32)),
                                   C.POINTER(C.POINTER(C.c_int32)), C.POINTER(C.POINTER(C.c_double))]
        L.b200mtx_free.argtypes = [C.c_void_p]
        L.b200mtx_last_error.restype = C.c_char_p
        _lib = L
    return _lib


def read_mtx(path, expand_symmetry=True):
    """(row, col, val, info): base-0 COO in the reference's order (file entries, then the mirrors), values as doubles
    (pattern -> 1.0, complex -> |z|, integer widened), info = dict of the header facts."""
    L = _L()
    info = _Info()
    r, c, v = C.POINTER(C.c_int32)(), C.POINTER(C.c_int32)(), C.POINTER(C.c_double)()
    if L.b200mtx_read(os.fsencode(path), 1 if expand_symmetry else 0, C.byref(info), C.byref(r), C.byref(c), C.byref(v)):
        raise MtxError(L.b200mtx_last_error().decode())
    try:
        k = int(info.nnz)
        row = np.ctypeslib.as_array(r, shape=(max(k, 1),))[:k].copy()
        col = np.ctypeslib.as_array(c, shape=(max(k, 1),))[:k].copy()
        val = np.ctypeslib.as_array(v, shape=(max(k, 1),))[:k].copy()
    finally:
        for p in (r, c, v):
            L.b200mtx_free(p)
    d = {k_: int(getattr(info, k_)) for k_, _ in _Info._fields_}
    d["field"] = FIELDS[d["field"]]
    return row, col, val, d


def mtx_to_format(path, **opts):
    """The reference's import_file() + csr_to_format() for a .mtx file: host parse, COO -> CSR on the GPU, format handle.
    Returns (MatrixFormat, info)."""
    from . import formats
    row, col, val, info = read_mtx(path, True)
    return formats.coo_to_format(row, col, val, info["m"], info["n"], **opts), info

"""Synthetic matrix source: ctypes binding of libb200amg.so (csrc/amg.c).

Host-side mirror of the reference harness's synthetic import
(benchmark_code/BENCH/src/bench.cpp:242-283 -> artificial_matrix_generation(), 11 parameters
`rows cols avg std distribution placement bw skew neigh crs seed`), bit-identical output.
"""
import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "lib", "libb200amg.so")
_SRC = os.path.join(_HERE, "csrc", "amg.c")


class _Params(C.Structure):
    _fields_ = [("nr_rows", C.c_int64), ("nr_cols", C.c_int64), ("avg", C.c_double), ("std", C.c_double),
                ("distribution", C.c_char), ("placement", C.c_char), ("pad_", C.c_char * 6),
                ("bw", C.c_double), ("skew", C.c_double), ("neigh", C.c_double), ("crs", C.c_double),
                ("seed", C.c_uint32), ("pad2_", C.c_uint32)]


def build(force=False):
    """gcc build of the generator; -ffp-contract=off keeps the double rounding sequence of the reference."""
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(_SRC):
        os.makedirs(os.path.dirname(_SO), exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-fopenmp", "-ffp-contract=off", "-fPIC", "-shared", "-Wall", "-Wextra",
                               "-o", _SO, _SRC, "-lm"])
    return _SO


_lib = None


def _L():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise RuntimeError("libb200amg.so is not built: run `python -c 'import __graft_entry__ as g; g.build()'`")
        L = C.CDLL(_SO)
        i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
        L.b200amg_row_ptr.restype = C.c_int64
        L.b200amg_row_ptr.argtypes = [C.POINTER(_Params), i32p, C.POINTER(C.c_double)]
        L.b200amg_bw_corrected.restype = C.c_double
        L.b200amg_bw_corrected.argtypes = [C.POINTER(_Params), C.c_int64]
        L.b200amg_fill.restype = C.c_int
        L.b200amg_fill.argtypes = [C.POINTER(_Params), i32p, C.c_double, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
        L.b200amg_fnv1a64.restype = C.c_uint64
        L.b200amg_fnv1a64.argtypes = [C.c_void_p, C.c_int64]
        L.b200amg_last_error.restype = C.c_char_p
        _lib = L
    return _lib


@dataclass(frozen=True)
class MatrixParams:
    """One line of a reference parameter file (matrix_generation_parameters/*.txt)."""
    rows: int
    cols: int
    avg: float
    std: float
    distribution: str = "normal"
    placement: str = "random"
    bw: float = 0.3
    skew: float = 0.0
    neigh: float = 0.5
    crs: float = 0.5
    seed: int = 14
    name: str = ""

    @staticmethod
    def parse(line):
        t = line.split()
        if len(t) < 11:
            raise ValueError("parameter line needs 11 fields: rows cols avg std dist placement bw skew neigh crs seed [name]")
        return MatrixParams(int(t[0]), int(t[1]), float(t[2]), float(t[3]), t[4], t[5], float(t[6]), float(t[7]),
                            float(t[8]), float(t[9]), int(t[10]), t[11] if len(t) > 11 else "")

    def line(self):
        return "%d %d %g %g %s %s %g %g %g %g %d" % (self.rows, self.cols, self.avg, self.std, self.distribution,
                                                      self.placement, self.bw, self.skew, self.neigh, self.crs, self.seed)

    def _c(self):
        return _Params(self.rows, self.cols, self.avg, self.std, self.distribution[:1].encode(), self.placement[:1].encode(),
                       b"", self.bw, self.skew, self.neigh, self.crs, self.seed & 0xFFFFFFFF, 0)


def row_ptr(p):
    """Pass 1: (row_ptr int32[m+1], bw_corrected)."""
    rp = np.zeros(p.rows + 1, np.int32)
    bw = C.c_double()
    cp = p._c()
    nnz = _L().b200amg_row_ptr(C.byref(cp), rp, C.byref(bw))
    if nnz < 0:
        raise RuntimeError(_L().b200amg_last_error().decode())
    return rp, bw.value


def bw_corrected(p, nnz):
    """The corrected band width of a line whose pass 1 (shared by every line with the same rows / avg / std / distribution /
    skew / seed) gave nnz non-zeros."""
    cp = p._c()
    return float(_L().b200amg_bw_corrected(C.byref(cp), int(nnz)))


def fill(p, rp, bw_corrected, i_s, i_e, col_out=None, val_out=None):
    """Pass 2 for rows [i_s, i_e): (col_ind int32, values f64), optionally into caller (e.g. pinned) buffers."""
    cnt = int(rp[i_e]) - int(rp[i_s])
    col = np.empty(cnt, np.int32) if col_out is None else col_out
    val = np.empty(cnt, np.float64) if val_out is None else val_out
    assert col.size >= cnt and val.size >= cnt
    cp = p._c()
    rc = _L().b200amg_fill(C.byref(cp), rp, bw_corrected, i_s, i_e, col.ctypes.data, val.ctypes.data)
    if rc:
        raise RuntimeError(_L().b200amg_last_error().decode())
    return col[:cnt], val[:cnt]


def fill_device(p, d_row_ptr, bw_corrected, i_s, i_e, d_col, d_val, stream=None):
    """Pass 2 on the GPU (libb200spmv.so: b200spmv_amg_fill_device, csrc/amg_device.cu): the non-zeros of rows [i_s, i_e)
    written straight into HBM, bit-identical to fill().  d_row_ptr: the WHOLE row_ptr in device memory; d_col / d_val:
    device buffers of row_ptr[i_e] - row_ptr[i_s] entries (torch tensors or raw addresses).  No CPU fallback."""
    from . import capi
    cp = p._c()
    capi.check(capi.lib().b200spmv_amg_fill_device(C.cast(C.pointer(cp), C.c_void_p), capi._ptr(d_row_ptr), bw_corrected,
                                                   int(i_s), int(i_e), capi._ptr(d_col), capi._ptr(d_val), stream))


def generate_device(p, i_s=0, i_e=None, device="cuda:0"):
    """Rows [i_s, i_e) generated into HBM: pass 1 on the host (the libm-dependent row lengths), pass 2 on the GPU.
    Returns (row_ptr numpy int32[m+1] of the whole matrix, d_row_ptr, d_col, d_val) -- the last three torch tensors on
    `device`, d_row_ptr whole, ready for capi.Handle(d_row_ptr[i_s:i_e+1], d_col, d_val, ..., device_arrays=True)."""
    import torch
    i_e = p.rows if i_e is None else i_e
    rp, bw = row_ptr(p)
    cnt = int(rp[i_e]) - int(rp[i_s])
    with torch.cuda.device(device):
        d_rp = torch.from_numpy(rp).to(device)
        d_col = torch.empty(cnt, dtype=torch.int32, device=device)
        d_val = torch.empty(cnt, dtype=torch.float64, device=device)
        fill_device(p, d_rp, bw, i_s, i_e, d_col, d_val, torch.cuda.current_stream().cuda_stream)
    return rp, d_rp, d_col, d_val


def generate(p):
    """Whole matrix: (row_ptr, col_ind, values) exactly as the reference generator returns them."""
    rp, bw = row_ptr(p)
    col, val = fill(p, rp, bw, 0, p.rows)
    return rp, col, val


def fnv1a64(a):
    a = np.ascontiguousarray(a)
    return int(_L().b200amg_fnv1a64(a.ctypes.data, a.nbytes))

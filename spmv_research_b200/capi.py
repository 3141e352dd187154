"""ctypes binding of libb200spmv.so -- the C-ABI declared in include/b200spmv.h.

This module is plumbing only: it never computes anything itself and has no fallback.  If the
CUDA library is missing or no sm_100 device is present, the calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200SPMV_LIB: another build of the same library (tools/sweep_exp.py times experimental kernel builds, lib/exp/*.so);
# unset = the shipped library.  Still a CUDA library either way: no CPU path exists.
SO_PATH = os.environ.get("B200SPMV_LIB") or os.path.join(_HERE, "lib", "libb200spmv.so")

CSR_ADAPTIVE = 0
SELL_C_SIGMA = 1
AUTO = 2
SKIP_X_UPLOAD = 1
SKIP_Y_DOWNLOAD = 2
X_ROW_ORDER, X_SLICE_WIDTH, X_SLICE_PTR, X_SELL_COL, X_SELL_VAL, X_TILE_ROW, X_SPLIT_ROW, X_SPLIT_TILE = range(1, 9)
_EXPORT_DTYPE = {X_ROW_ORDER: np.int32, X_SLICE_WIDTH: np.int32, X_SLICE_PTR: np.int64, X_SELL_COL: np.int32,
                 X_SELL_VAL: np.float64, X_TILE_ROW: np.int32, X_SPLIT_ROW: np.int32, X_SPLIT_TILE: np.int32}


class B200Error(RuntimeError):
    pass


class Opts(C.Structure):
    _fields_ = [("format", C.c_int32), ("device", C.c_int32), ("sell_C", C.c_int32), ("sell_sigma", C.c_int32),
                ("csr_nnz_per_thread", C.c_int32), ("csr_ext", C.c_int32), ("x_hint", C.c_int32), ("sell_sigma_alt", C.c_int32), ("sell_variant", C.c_int32), ("csr_compress_rows", C.c_int32), ("reserved0", C.c_int32), ("keep_diag", C.c_int32), ("pin_host_vectors", C.c_int32), ("col_panels", C.c_int32), ("host_pipeline", C.c_int32), ("reserved", C.c_int32 * 1)]


class Launch(C.Structure):
    _fields_ = [("scale", C.c_void_p), ("norm2", C.c_void_p), ("y_peer", C.c_void_p * 8), ("n_peer", C.c_int32), ("pad_", C.c_int32),
                ("peer_row_lo", C.c_int64), ("peer_row_hi", C.c_int64)]


class PowerOpts(C.Structure):
    _fields_ = [("exchange", C.c_int32), ("chunks", C.c_int32), ("full_exchange", C.c_int32), ("use_graph", C.c_int32),
                ("scalar_sync", C.c_int32), ("reserved", C.c_int32 * 3)]


X_NCCL, X_FUSED = 0, 1


class PcgResult(C.Structure):
    _fields_ = [("iterations", C.c_int64), ("products", C.c_int64), ("restarts", C.c_int64), ("err", C.c_double),
                ("err_explicit", C.c_double), ("err_best", C.c_double), ("eps", C.c_double), ("converged", C.c_int32), ("pad_", C.c_int32)]


_lib = None


def lib():
    """Load libb200spmv.so (built in-tree by spmv_research_b200.build); fail loudly if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise B200Error("%s is missing: the CUDA extension was not built (python -m spmv_research_b200.build). "
                        "There is no CPU fallback." % SO_PATH)
    L = C.CDLL(SO_PATH)
    vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
    sig = {
        "b200spmv_default_opts": (None, [C.POINTER(Opts)]),
        "b200spmv_create": (C.c_int, [C.POINTER(vp), vp, vp, vp, i64, i64, i64, C.POINTER(Opts)]),
        "b200spmv_create_device": (C.c_int, [C.POINTER(vp), vp, vp, vp, i64, i64, i64, C.POINTER(Opts)]),
        "b200spmv_destroy": (None, [vp]),
        "b200spmv_spmv_host": (C.c_int, [vp, vp, vp]),
        "b200spmv_spmv_host_ex": (C.c_int, [vp, vp, vp, i32]),
        "b200spmv_spmv_device": (C.c_int, [vp, vp, vp, vp]),
        "b200spmv_spmv_device_ex": (C.c_int, [vp, vp, vp, vp, C.POINTER(Launch)]),
        "b200spmv_launches_per_spmv": (i32, [vp]),
        "b200spmv_format": (i32, [vp]),
        "b200spmv_col_range": (C.c_int, [vp, C.POINTER(i64), C.POINTER(i64)]),
        "b200spmv_stats": (C.c_int, [vp, C.POINTER(C.c_double), i32]),
        "b200spmv_export": (C.c_int, [vp, i32, vp, i64, C.POINTER(i64)]),
        "b200spmv_partition_device": (C.c_int, [vp, i64, i64, i32, vp]),
        "b200spmv_partition": (C.c_int, [vp, i64, i64, i32, vp]),
        "b200spmv_partition_device_noalloc": (C.c_int, [vp, i64, i64, i32, vp, vp]),
        "b200spmv_merge_coords": (C.c_int, [vp, i32, vp]),
        "b200spmv_device_count": (C.c_int, [C.POINTER(i32)]),
        "b200spmv_set_device": (C.c_int, [i32]),
        "b200spmv_malloc": (C.c_int, [C.POINTER(vp), i64]),
        "b200spmv_free": (C.c_int, [vp]),
        "b200spmv_memcpy_h2d": (C.c_int, [vp, vp, i64]),
        "b200spmv_memcpy_d2h": (C.c_int, [vp, vp, i64]),
        "b200spmv_malloc_host": (C.c_int, [C.POINTER(vp), i64]),
        "b200spmv_free_host": (C.c_int, [vp]),
        "b200spmv_synchronize": (C.c_int, []),
        "b200spmv_ipc_export": (C.c_int, [vp, vp]),
        "b200spmv_ipc_import": (C.c_int, [vp, C.POINTER(vp)]),
        "b200spmv_ipc_close": (C.c_int, [vp]),
        "b200spmv_time_device": (C.c_int, [vp, vp, vp, i32, i64, vp]),
        "b200spmv_pcg": (C.c_int, [vp, vp, vp, i64, C.POINTER(PcgResult)]),
        "b200spmv_bicgstab": (C.c_int, [vp, vp, vp, i64, vp, C.POINTER(PcgResult)]),
        "b200spmv_coo_to_csr": (C.c_int, [vp, vp, vp, i64, i64, i64, vp, vp, vp, C.POINTER(i64), vp]),
        "b200spmv_create_coo": (C.c_int, [C.POINTER(vp), vp, vp, vp, i64, i64, i64, C.POINTER(Opts)]),
        "b200spmv_f32_create": (C.c_int, [C.POINTER(vp), vp, vp, vp, i64, i64, i64, i32]),
        "b200spmv_f32_spmv_host": (C.c_int, [vp, vp, vp]),
        "b200spmv_f32_spmv_device": (C.c_int, [vp, vp, vp, vp]),
        "b200spmv_f32_time_device": (C.c_int, [vp, vp, vp, i32, i64, vp]),
        "b200spmv_f32_destroy": (None, [vp]),
        "b200spmv_amg_fill_device": (C.c_int, [vp, vp, C.c_double, i64, i64, vp, vp, vp]),
        "b200spmv_comm_unique_id": (C.c_int, [vp]),
        "b200spmv_comm_create": (C.c_int, [C.POINTER(vp), vp, i32, i32, i32]),
        "b200spmv_comm_destroy": (None, [vp]),
        "b200spmv_power_create": (C.c_int, [C.POINTER(vp), vp, vp, vp, i64, C.POINTER(PowerOpts)]),
        "b200spmv_power_set_start": (C.c_int, [vp, vp]),
        "b200spmv_power_iterate": (C.c_int, [vp, i32, C.POINTER(C.c_double), C.POINTER(C.c_float)]),
        "b200spmv_power_get_x": (C.c_int, [vp, vp, C.POINTER(C.c_double)]),
        "b200spmv_power_info": (C.c_int, [vp, vp, i32]),
        "b200spmv_power_destroy": (None, [vp]),
        "b200spmv_chunk_schedule": (C.c_int, [i32, i32, i64, vp, vp, vp, vp]),
        "b200spmv_last_error": (C.c_char_p, []),
        "b200spmv_version": (C.c_char_p, []),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)
        f.restype = res
        f.argtypes = args
    _lib = L
    return L


def check(rc):
    if rc:
        raise B200Error(lib().b200spmv_last_error().decode(errors="replace"))


def _ptr(a):
    """Address of a numpy array / torch tensor / raw int."""
    if a is None:
        return None
    if isinstance(a, int):
        return a
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    if hasattr(a, "data_ptr"):
        return a.data_ptr()
    raise TypeError("cannot take the address of %r" % type(a))


def default_opts(**kw):
    o = Opts()
    lib().b200spmv_default_opts(C.byref(o))
    for k, v in kw.items():
        if not hasattr(o, k):
            raise TypeError("unknown option %s" % k)
        setattr(o, k, v)
    return o


class Handle:
    """Owns one b200spmv_t*."""

    def __init__(self, row_ptr, col_ind, values, m, n, nnz, opts=None, device_arrays=False, coo=False):
        """CSR arrays on the host (default) or in HBM (device_arrays); coo=True: row_ptr is the COO row index array (host)."""
        self._h = C.c_void_p()
        self.m, self.n, self.nnz = int(m), int(n), int(nnz)
        self._opts = opts if opts is not None else default_opts()
        fn = lib().b200spmv_create_coo if coo else (lib().b200spmv_create_device if device_arrays else lib().b200spmv_create)
        check(fn(C.byref(self._h), _ptr(row_ptr), _ptr(col_ind), _ptr(values), self.m, self.n, self.nnz, C.byref(self._opts)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value and _lib is not None:
            _lib.b200spmv_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def spmv_host(self, x, y, flags=0):
        assert x.dtype == np.float64 and y.dtype == np.float64 and x.size >= self.n and y.size >= self.m
        check(lib().b200spmv_spmv_host_ex(self._h, _ptr(x), _ptr(y), flags))

    def spmv_device(self, dx, dy, stream=None, scale=None, norm2=None, peers=(), peer_rows=None):
        if scale is None and norm2 is None and not peers:
            check(lib().b200spmv_spmv_device(self._h, _ptr(dx), _ptr(dy), stream))
            return
        l = Launch()
        l.scale = _ptr(scale)
        l.norm2 = _ptr(norm2)
        l.n_peer = len(peers)
        if peer_rows is not None:
            l.peer_row_lo, l.peer_row_hi = int(peer_rows[0]), int(peer_rows[1])
            if l.peer_row_hi <= l.peer_row_lo:
                l.n_peer = 0                      # nothing of this shard is referenced by anybody
        for i, p in enumerate(peers):
            l.y_peer[i] = _ptr(p)
        check(lib().b200spmv_spmv_device_ex(self._h, _ptr(dx), _ptr(dy), stream, C.byref(l)))

    @property
    def launches_per_spmv(self):
        return int(lib().b200spmv_launches_per_spmv(self._h))

    @property
    def format(self):
        """The format that runs (AUTO handles report what they resolved to)."""
        return int(lib().b200spmv_format(self._h))

    def col_range(self):
        """[lo, hi) of the column indices the matrix references."""
        lo, hi = C.c_int64(), C.c_int64()
        check(lib().b200spmv_col_range(self._h, C.byref(lo), C.byref(hi)))
        return lo.value, hi.value

    def stats(self):
        out = (C.c_double * 13)()
        check(lib().b200spmv_stats(self._h, out, 13))
        keys = ("mem_footprint", "units", "nsplit", "entries", "unit_size", "ext_or_sigma", "convert_ms",
                "trial_ms_csr", "trial_ms_sell", "trial_ms_sell_alt", "compressed_rows", "n_empty", "col_panels")
        return dict(zip(keys, list(out)))

    def export(self, what):
        cnt = C.c_int64()
        check(lib().b200spmv_export(self._h, what, None, 0, C.byref(cnt)))
        a = np.zeros(max(cnt.value, 1), _EXPORT_DTYPE[what])
        check(lib().b200spmv_export(self._h, what, a.ctypes.data, a.nbytes, C.byref(cnt)))
        return a[:cnt.value]

    def merge_coords(self, W):
        out = np.zeros((W + 1, 2), np.int64)
        check(lib().b200spmv_merge_coords(self._h, W, out.ctypes.data))
        return out

    def pcg(self, b, max_iterations):
        """Jacobi-preconditioned CG of the reference harness (bench_cg.cpp:93-320) on the device; returns (x, result dict)."""
        b = np.ascontiguousarray(b, np.float64)
        assert b.size >= self.m
        x = np.zeros(self.m)
        r = PcgResult()
        check(lib().b200spmv_pcg(self._h, _ptr(b), _ptr(x), int(max_iterations), C.byref(r)))
        return x, {k: getattr(r, k) for k, _ in PcgResult._fields_ if k != "pad_"}

    def bicgstab(self, b, max_iterations, trace=False):
        """Jacobi-preconditioned BiCGSTAB of the reference harness (bench_bicg.cpp:150-456) on the device: exactly
        max_iterations iterations (the reference has no stop test); returns (x_best, result dict [+ "trace": ||r_k||])."""
        b = np.ascontiguousarray(b, np.float64)
        assert b.size >= self.m
        x = np.zeros(self.m)
        r = PcgResult()
        tr = np.zeros(int(max_iterations) + 1) if trace else None
        check(lib().b200spmv_bicgstab(self._h, _ptr(b), _ptr(x), int(max_iterations), _ptr(tr) if trace else None, C.byref(r)))
        out = {k: getattr(r, k) for k, _ in PcgResult._fields_ if k != "pad_"}
        if trace:
            out["trace"] = tr
        return x, out

    def time_device(self, dx, dy, iters, flush_bytes=0):
        ms = np.zeros(iters, np.float32)
        check(lib().b200spmv_time_device(self._h, _ptr(dx), _ptr(dy), iters, flush_bytes, ms.ctypes.data))
        return ms


def partition(sums, N, total, W, device=False):
    """Reference nnz-balanced partitioner evaluated on the GPU: int64[W, 2] of [i_s, i_e)."""
    out = np.zeros((W, 2), np.int64)
    fn = lib().b200spmv_partition_device if device else lib().b200spmv_partition
    check(fn(_ptr(sums), int(N), int(total), int(W), out.ctypes.data))
    return out


NOALLOC_MAX_WORKERS = 4096


def partition_noalloc(d_sums, N, total, W, stream=None):
    """b200spmv_partition_device_noalloc: the reference partitioner on DEVICE prefix sums with its scratch in a static device
    array (no cudaMalloc inside the call; W <= NOALLOC_MAX_WORKERS): int64[W, 2] of [i_s, i_e)."""
    out = np.zeros((max(int(W), 1), 2), np.int64)
    check(lib().b200spmv_partition_device_noalloc(_ptr(d_sums), int(N), int(total), int(W), out.ctypes.data, stream))
    return out


class HandleF32:
    """Owns one b200spmv_f32_t* (the `_f` build of the CSR format: double matrix in, float vectors)."""

    def __init__(self, row_ptr, col_ind, values, m, n, nnz, device=-1):
        self._h = C.c_void_p()
        self.m, self.n, self.nnz = int(m), int(n), int(nnz)
        check(lib().b200spmv_f32_create(C.byref(self._h), _ptr(row_ptr), _ptr(col_ind), _ptr(values), self.m, self.n, self.nnz, device))

    def close(self):
        if self._h:
            lib().b200spmv_f32_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def spmv_host(self, x, y):
        check(lib().b200spmv_f32_spmv_host(self._h, _ptr(x), _ptr(y)))

    def spmv_device(self, dx, dy, stream=None):
        check(lib().b200spmv_f32_spmv_device(self._h, _ptr(dx), _ptr(dy), stream))

    def time_device(self, dx, dy, iters, flush_bytes=0):
        ms = (C.c_float * iters)()
        check(lib().b200spmv_f32_time_device(self._h, _ptr(dx), _ptr(dy), iters, int(flush_bytes), ms))
        return [float(v) for v in ms]


def coo_to_csr_device(d_row, d_col, d_val, m, n, nnz, d_row_ptr, d_col_out, d_val_out, stream=None):
    """b200spmv_coo_to_csr on device buffers (torch tensors or raw addresses); returns the number of duplicate pairs."""
    d = C.c_int64()
    check(lib().b200spmv_coo_to_csr(_ptr(d_row), _ptr(d_col), _ptr(d_val), int(m), int(n), int(nnz), _ptr(d_row_ptr), _ptr(d_col_out),
                                    _ptr(d_val_out), C.byref(d), stream))
    return d.value


def device_count():
    n = C.c_int32()
    try:
        check(lib().b200spmv_device_count(C.byref(n)))
    except B200Error:
        return 0
    return n.value


def malloc(nbytes):
    """cudaMalloc through the library (a base allocation: its IPC handle maps exactly this address)."""
    p = C.c_void_p()
    check(lib().b200spmv_malloc(C.byref(p), int(nbytes)))
    return p.value


def free(ptr):
    check(lib().b200spmv_free(C.c_void_p(ptr)))


def ipc_export(ptr):
    buf = (C.c_ubyte * 64)()
    check(lib().b200spmv_ipc_export(C.c_void_p(ptr), buf))
    return bytes(buf)


def ipc_import(handle_bytes):
    buf = (C.c_ubyte * 64).from_buffer_copy(handle_bytes)
    p = C.c_void_p()
    check(lib().b200spmv_ipc_import(buf, C.byref(p)))
    return p.value


# ---- multi-GPU power iteration (csrc/power.cu): thin callers, no arithmetic here ------------------------------------------
def comm_unique_id():
    buf = (C.c_ubyte * 128)()
    check(lib().b200spmv_comm_unique_id(buf))
    return bytes(buf)


class Comm:
    """Owns one b200spmv_comm_t* (an NCCL communicator created inside the library)."""

    def __init__(self, uid, rank, world, device=-1):
        self._c = C.c_void_p()
        self.rank, self.world = int(rank), int(world)
        buf = (C.c_ubyte * 128).from_buffer_copy(uid)
        check(lib().b200spmv_comm_create(C.byref(self._c), buf, self.rank, self.world, int(device)))

    def close(self):
        if getattr(self, "_c", None) is not None and self._c.value and _lib is not None:
            _lib.b200spmv_comm_destroy(self._c)
            self._c = C.c_void_p()

    __del__ = close


class Power:
    """Owns one b200spmv_power_t*: the row-sharded power iteration of include/b200spmv.h."""

    def __init__(self, handle, comm, bounds, n, exchange=X_NCCL, chunks=0, full_exchange=False, use_graph=False, scalar_sync=0):
        self._p = C.c_void_p()
        self._keep = (handle, comm)                       # the shard and the communicator must outlive the object
        self.n = int(n)
        b = np.ascontiguousarray(np.asarray(bounds, np.int64).reshape(-1))
        o = PowerOpts(exchange=int(exchange), chunks=int(chunks), full_exchange=int(bool(full_exchange)), use_graph=int(bool(use_graph)),
                      scalar_sync=int(scalar_sync))
        check(lib().b200spmv_power_create(C.byref(self._p), handle._h, comm._c if comm is not None else None, b.ctypes.data, self.n, C.byref(o)))

    def set_start(self, x0):
        x0 = np.ascontiguousarray(x0, np.float64)
        assert x0.size == self.n
        check(lib().b200spmv_power_set_start(self._p, x0.ctypes.data))

    def iterate(self, iters):
        """Returns (eigenvalue estimate, device ms of the iterations)."""
        eig, ms = C.c_double(), C.c_float()
        check(lib().b200spmv_power_iterate(self._p, int(iters), C.byref(eig), C.byref(ms)))
        return eig.value, ms.value

    def get_x(self):
        """(current un-normalised iterate as a host array, its scale)."""
        x = np.empty(self.n)
        s = C.c_double()
        check(lib().b200spmv_power_get_x(self._p, x.ctypes.data, C.byref(s)))
        return x, s.value

    def info(self):
        v = (C.c_int64 * 7)()
        check(lib().b200spmv_power_info(self._p, v, 7))
        return dict(zip(("chunks", "panels", "band_lo", "band_hi", "graph", "launches_per_iteration", "nccl_calls_per_iteration"), [int(t) for t in v]))

    def close(self):
        if getattr(self, "_p", None) is not None and self._p.value and _lib is not None:
            _lib.b200spmv_power_destroy(self._p)
            self._p = C.c_void_p()

    __del__ = close


def chunk_schedule(tile_rows, K, m):
    """b200spmv_chunk_schedule: tile_rows = one int32 array per column panel; returns (unit[P, K + 1], final_row[K + 1])."""
    P = len(tile_rows)
    arrs = [np.ascontiguousarray(a, np.int32) for a in tile_rows]
    ptrs = (C.c_void_p * P)(*[a.ctypes.data for a in arrs])
    nt = np.array([a.size - 1 for a in arrs], np.int64)
    unit = np.zeros((P, K + 1), np.int64)
    fin = np.zeros(K + 1, np.int64)
    check(lib().b200spmv_chunk_schedule(P, int(K), int(m), ptrs, nt.ctypes.data, unit.ctypes.data, fin.ctypes.data))
    return unit, fin

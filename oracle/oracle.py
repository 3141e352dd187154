"""ctypes bindings for oracle/liboracle.so (CPU restatement) and oracle/_ref/*.so (the
reference's own code compiled from /root/reference by oracle/Makefile).

TEST INFRASTRUCTURE ONLY: import from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Never from the product package.
"""
import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")


def build(force=False):
    """Compile liboracle.so (and oracle/_ref when the reference tree is present)."""
    so = os.path.join(HERE, "liboracle.so")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(os.path.join(HERE, "spmv_oracle.c")):
        subprocess.check_call(["make", "-s", "-C", HERE, "liboracle.so"])
    if os.path.isdir("/root/reference/benchmark_code"):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])
    return so


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        L.oracle_csr_spmv_scalar.argtypes = [_i32p, _i32p, _f64p, _f64p, _f64p, C.c_int64, C.c_int64]
        L.oracle_row_abs_bound.argtypes = [_i32p, _i32p, _f64p, _f64p, _f64p, C.c_int64]
        L.oracle_csr_spmv_scalar_f32.argtypes = [_i32p, _i32p, _f32p, _f32p, _f32p, C.c_int64, C.c_int64]
        L.oracle_row_abs_bound_f32.argtypes = [_i32p, _i32p, _f32p, _f32p, _f64p, C.c_int64]
        L.oracle_partition_prefix_sums.argtypes = [C.c_int64, C.c_int64, _i32p, C.c_int64, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.oracle_csr_spmv_omp.argtypes = [_i32p, _i32p, _f64p, _f64p, _f64p, C.c_int64, C.c_int64, _i64p, C.c_int]
        L.oracle_max_threads.restype = C.c_int
        L.oracle_merge_path_search.argtypes = [C.c_int64, _i32p, C.c_int64, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.oracle_merge_csrmv.argtypes = [C.c_int64, _i32p, _i32p, _f64p, _f64p, _f64p, C.c_int64, C.c_int64, _i64p]
        L.oracle_sellcs_convert.argtypes = [_i32p, _i32p, _f64p, C.c_int64, C.c_int64, C.c_int64, _i32p, _i32p, _i64p, C.c_void_p, C.c_void_p]
        L.oracle_sellcs_convert.restype = C.c_int64
        L.oracle_sellcs_spmv.argtypes = [C.c_int64, C.c_int64, _i32p, _i32p, _i64p, _i32p, _f64p, _f64p, _f64p]
        L.oracle_check_accuracy.argtypes = [_i32p, _i32p, _f64p, _f64p, _f64p, C.c_int64, C.c_double, C.c_void_p, _f64p]
        L.oracle_adaptive_rowblocks.argtypes = [_i32p, C.c_int64, C.c_int64, C.c_int64, _i32p]
        L.oracle_adaptive_rowblocks.restype = C.c_int64
        L.oracle_pcg.argtypes = [_i32p, _i32p, _f64p, C.c_int64, _f64p, _f64p, C.c_int64, _f64p]
        L.oracle_pcg.restype = C.c_int
        L.oracle_bicgstab.argtypes = [_i32p, _i32p, _f64p, C.c_int64, _f64p, _f64p, C.c_int64, _f64p, _f64p]
        L.oracle_bicgstab.restype = C.c_int
        L.oracle_sell_plain_convert.argtypes = [_i32p, _i32p, _f64p, C.c_int64, C.c_int64, C.c_int64, _i32p, C.c_void_p, C.c_void_p]
        L.oracle_sell_plain_convert.restype = C.c_int64
        L.oracle_sell_sorted_convert.argtypes = [_i32p, _i32p, _f64p, C.c_int64, C.c_int64, C.c_int64, _i32p, _i32p, _i32p, C.c_void_p, C.c_void_p]
        L.oracle_sell_sorted_convert.restype = C.c_int64
        L.oracle_coo_to_csr.argtypes = [_i32p, _i32p, _f64p, C.c_int64, C.c_int64, C.c_int64, _i32p, _i32p, _f64p]
        L.oracle_coo_to_csr.restype = C.c_int64
        L.oracle_sell_plain_spmv.argtypes = [C.c_int64, C.c_int64, _i32p, _i32p, _f64p, _f64p, _f64p]
        L.oracle_fnv1a64.argtypes = [C.c_void_p, C.c_int64]
        L.oracle_fnv1a64.restype = C.c_uint64
        _lib = L
    return _lib


# ---------------------------------------------------------------- restatement wrappers
def _pad1(a, dtype):
    """ctypes ndpointer rejects zero-size views of None; keep at least one element."""
    a = np.ascontiguousarray(a, dtype=dtype)
    return a if a.size else np.zeros(1, dtype)


def csr_spmv_f32(row_ptr, col_ind, val, x, y_prefill=1.0):
    """The reference's `_f` build of the serial CSR kernel: val (double, narrowed here with the C cast the plugin's constructor
    applies) and x as float32, float accumulation in ascending j.  Returns float32 y."""
    m = len(row_ptr) - 1
    y = np.full(m, y_prefill, np.float32)
    if m:
        lib().oracle_csr_spmv_scalar_f32(row_ptr, _pad1(col_ind, np.int32), _pad1(np.asarray(val).astype(np.float32), np.float32),
                                         np.ascontiguousarray(x, np.float32), y, 0, m)
    return y


def row_abs_bound_f32(row_ptr, col_ind, val, x):
    """sum_j |a_ij x_j| per row, in double, from the float32-narrowed operands."""
    m = len(row_ptr) - 1
    b = np.zeros(m, np.float64)
    if m:
        lib().oracle_row_abs_bound_f32(row_ptr, _pad1(col_ind, np.int32), _pad1(np.asarray(val).astype(np.float32), np.float32),
                                       np.ascontiguousarray(x, np.float32), b, m)
    return b


def csr_spmv(row_ptr, col_ind, val, x, y_prefill=1.0):
    m = len(row_ptr) - 1
    y = np.full(m, y_prefill, np.float64)
    if m:
        lib().oracle_csr_spmv_scalar(row_ptr, _pad1(col_ind, np.int32), _pad1(val, np.float64), x, y, 0, m)
    return y


def row_abs_bound(row_ptr, col_ind, val, x):
    m = len(row_ptr) - 1
    b = np.zeros(max(m, 1), np.float64)
    lib().oracle_row_abs_bound(row_ptr, _pad1(col_ind, np.int32), _pad1(val, np.float64), x, b, m)
    return b[:m]


def csr_spmv_omp(row_ptr, col_ind, val, x, y, bounds, threads):
    m = len(row_ptr) - 1
    lib().oracle_csr_spmv_omp(row_ptr, col_ind, val, x, y, m, int(row_ptr[m]) - int(row_ptr[0]), bounds, threads)


def max_threads():
    return lib().oracle_max_threads()


def partition(num_workers, sums, N, total):
    """All W ranges of loop_partitioner_balance_prefix_sums; returns int64 array [W, 2]."""
    out = np.zeros((num_workers, 2), np.int64)
    s, e = C.c_int64(), C.c_int64()
    for p in range(num_workers):
        lib().oracle_partition_prefix_sums(num_workers, p, sums, N, total, C.byref(s), C.byref(e))
        out[p] = (s.value, e.value)
    return out


def merge_path_search(diagonal, row_ptr, m, nnz):
    r, z = C.c_int64(), C.c_int64()
    row_end = np.ascontiguousarray(row_ptr[1:]) if m else np.zeros(1, np.int32)
    lib().oracle_merge_path_search(diagonal, row_end, m, nnz, C.byref(r), C.byref(z))
    return r.value, z.value


def merge_csrmv(T, row_ptr, col_ind, val, x, y_prefill=1.0):
    m = len(row_ptr) - 1
    nnz = int(row_ptr[m])
    y = np.full(m, y_prefill, np.float64)
    coords = np.zeros(2 * (T + 1), np.int64)
    lib().oracle_merge_csrmv(T, row_ptr, _pad1(col_ind, np.int32), _pad1(val, np.float64), x, y, m, nnz, coords)
    return y, coords.reshape(T + 1, 2)


def sellcs_convert(row_ptr, col_ind, val, C_, sigma):
    m = len(row_ptr) - 1
    ns = (m + C_ - 1) // C_
    row_order = np.zeros(max(m, 1), np.int32)
    width = np.zeros(max(ns, 1), np.int32)
    sptr = np.zeros(ns + 1, np.int64)
    ci, va = _pad1(col_ind, np.int32), _pad1(val, np.float64)
    total = lib().oracle_sellcs_convert(row_ptr, ci, va, m, C_, sigma, row_order, width, sptr, None, None)
    col = np.zeros(max(total, 1), np.int32)
    v = np.zeros(max(total, 1), np.float64)
    lib().oracle_sellcs_convert(row_ptr, ci, va, m, C_, sigma, row_order, width, sptr, col.ctypes.data, v.ctypes.data)
    return dict(row_order=row_order[:m], slice_width=width[:ns], slice_ptr=sptr, col=col[:total], val=v[:total], C=C_, sigma=sigma, m=m)


def sellcs_spmv(S, x, y_prefill=1.0):
    y = np.full(S["m"], y_prefill, np.float64)
    if S["m"]:
        lib().oracle_sellcs_spmv(S["m"], S["C"], _pad1(S["row_order"], np.int32), _pad1(S["slice_width"], np.int32), S["slice_ptr"],
                                 _pad1(S["col"], np.int32), _pad1(S["val"], np.float64), x, y)
    return y


def check_accuracy(row_ptr, col_ind, val, x, y, eps=1e-10):
    m = len(row_ptr) - 1
    out = np.zeros(4, np.float64)
    gold = np.zeros(max(m, 1), np.float64)
    lib().oracle_check_accuracy(row_ptr, _pad1(col_ind, np.int32), _pad1(val, np.float64), x, _pad1(y, np.float64), m, eps, gold.ctypes.data, out)
    return dict(max_rel=out[0], mae=out[1], max_ae=out[2], mse=out[3], gold=gold[:m])


def adaptive_rowblocks(row_ptr, block_nnz, multiblock):
    m = len(row_ptr) - 1
    rb = np.zeros(m + multiblock + 2, np.int32)
    cnt = lib().oracle_adaptive_rowblocks(row_ptr, m, block_nnz, multiblock, rb)
    return rb[:cnt]


def pcg(row_ptr, col_ind, val, b, max_iterations):
    """Serial restatement of the reference's preconditioned_cg (bench_cg.cpp:93-320): (x_best, info dict)."""
    m = len(row_ptr) - 1
    x = np.zeros(max(m, 1))
    out = np.zeros(5)
    rc = lib().oracle_pcg(row_ptr, _pad1(col_ind, np.int32), _pad1(val, np.float64), m, np.ascontiguousarray(b, np.float64), x,
                          int(max_iterations), out)
    if rc == 1:
        raise ValueError("bad K, zero in diagonal")
    if rc:
        raise MemoryError("oracle_pcg")
    return x[:m], dict(iterations=int(out[0]), err=out[1], err_explicit=out[2], err_best=out[3], eps=out[4])


def bicgstab(row_ptr, col_ind, val, b, max_iterations):
    """Serial restatement of the reference's preconditioned_bicgstab (bench_bicg.cpp:150-456): (x_best, info dict).
    No stop test, like the reference: exactly max_iterations iterations; info["trace"][k] = ||r_k|| at the top of iteration k."""
    m = len(row_ptr) - 1
    x = np.zeros(max(m, 1))
    out = np.zeros(5)
    trace = np.zeros(max(int(max_iterations), 1))
    rc = lib().oracle_bicgstab(row_ptr, _pad1(col_ind, np.int32), _pad1(val, np.float64), m, np.ascontiguousarray(b, np.float64), x,
                               int(max_iterations), trace, out)
    if rc == 1:
        raise ValueError("bad K, zero in diagonal")
    if rc:
        raise MemoryError("oracle_bicgstab")
    return x[:m], dict(iterations=int(out[0]), err=out[1], err_explicit=out[2], err_best=out[3], eps=out[4],
                       trace=trace[:int(max_iterations)])


def sell_plain_convert(row_ptr, col_ind, val, VL, nthreads):
    """The reference's plain SELL layout (K/sell.cpp:98-192): dict(row_cluster_ptr, ja, a, VL, m).
    The layout depends on nthreads (padding entries carry the last real column the converting thread saw)."""
    m = len(row_ptr) - 1
    S = (m + VL - 1) // VL
    rcp = np.zeros(S + 1, np.int32)
    ci, v = _pad1(col_ind, np.int32), _pad1(val, np.float64)
    total = lib().oracle_sell_plain_convert(row_ptr, ci, v, m, VL, nthreads, rcp, None, None)
    ja = np.zeros(max(total, 1), np.int32)
    a = np.zeros(max(total, 1), np.float64)
    lib().oracle_sell_plain_convert(row_ptr, ci, v, m, VL, nthreads, rcp, ja.ctypes.data, a.ctypes.data)
    return dict(row_cluster_ptr=rcp, ja=ja[:total], a=a[:total], VL=VL, m=m)


def sell_sorted_convert(row_ptr, col_ind, val, VL, nthreads):
    """The reference's per-thread-sorted SELL layout (K/sell_sorted.cpp:125-298): the plain SELL dict plus permutation /
    rev_permutation (rows sorted by ascending length inside each converting thread's range)."""
    m = len(row_ptr) - 1
    S = (m + VL - 1) // VL
    rcp = np.zeros(S + 1, np.int32)
    perm, rev = np.zeros(max(m, 1), np.int32), np.zeros(max(m, 1), np.int32)
    ci, v = _pad1(col_ind, np.int32), _pad1(val, np.float64)
    total = lib().oracle_sell_sorted_convert(row_ptr, ci, v, m, VL, nthreads, perm, rev, rcp, None, None)
    ja = np.zeros(max(total, 1), np.int32)
    a = np.zeros(max(total, 1), np.float64)
    lib().oracle_sell_sorted_convert(row_ptr, ci, v, m, VL, nthreads, perm, rev, rcp, ja.ctypes.data, a.ctypes.data)
    return dict(row_cluster_ptr=rcp, ja=ja[:total], a=a[:total], VL=VL, m=m, permutation=perm[:m], rev_permutation=rev[:m])


def sell_sorted_spmv(layout, x):
    """y[rev_permutation[i]] = lane i of the plain product over the reordered layout (K/sell_sorted.cpp:338-419)."""
    m = layout["m"]
    ys = sell_plain_spmv(layout, x)
    y = np.zeros(m)
    y[layout["rev_permutation"]] = ys[:m]
    return y


def sell_plain_spmv(layout, x):
    """y of the plain SELL product (fused multiply-adds in ascending k, K/sell.cpp:232-273); S * VL entries, the first m are y."""
    m, VL = layout["m"], layout["VL"]
    y = np.zeros(((m + VL - 1) // VL) * VL + 1)
    lib().oracle_sell_plain_spmv(m, VL, layout["row_cluster_ptr"], _pad1(layout["ja"], np.int32), _pad1(layout["a"], np.float64),
                                 np.ascontiguousarray(x, np.float64), y)
    return y[:-1]


def coo_to_csr(row, col, val, m, n):
    """COO -> CSR as the reference's import_file() (bench.cpp:208-229, csr_gen.c:178-213): (row_ptr, col, val, n_duplicates)."""
    nnz = len(row)
    rp = np.zeros(m + 1, np.int32)
    co, vo = np.zeros(max(nnz, 1), np.int32), np.zeros(max(nnz, 1), np.float64)
    d = lib().oracle_coo_to_csr(_pad1(np.ascontiguousarray(row, np.int32), np.int32), _pad1(np.ascontiguousarray(col, np.int32), np.int32),
                                _pad1(np.ascontiguousarray(val, np.float64), np.float64), m, n, nnz, rp, co, vo)
    if d < 0:
        raise ValueError("COO index out of range")
    return rp, co[:nnz], vo[:nnz], int(d)


def ref_mtx_to_csr(path, threads=2):
    """The reference's own ingest (mtx_read + mtx_values_convert_to_real + coo_to_csr) through oracle/_ref/ref_mtx_dump.exe:
    dict(m, n, nnz, symmetric, nnz_diag, nnz_non_diag, row_ptr, col, val), or None when the tool is not built."""
    import subprocess
    import tempfile
    exe = os.path.join(HERE, "_ref", "ref_mtx_dump.exe")
    if not os.path.exists(exe):
        return None
    with tempfile.TemporaryDirectory() as td:
        out = os.path.join(td, "csr.bin")
        env = dict(os.environ, OMP_NUM_THREADS=str(threads))
        p = subprocess.run([exe, path, out], capture_output=True, text=True, env=env)
        if p.returncode != 0:
            raise RuntimeError("ref_mtx_dump failed: %s %s" % (p.stdout[-500:], p.stderr[-500:]))
        b = open(out, "rb").read()
    h = np.frombuffer(b[:48], np.int64)
    m, n, nnz = int(h[0]), int(h[1]), int(h[2])
    o = 48
    rp = np.frombuffer(b[o:o + 4 * (m + 1)], np.int32).copy()
    o += 4 * (m + 1)
    ci = np.frombuffer(b[o:o + 4 * nnz], np.int32).copy()
    o += 4 * nnz
    va = np.frombuffer(b[o:o + 8 * nnz], np.float64).copy()
    return dict(m=m, n=n, nnz=nnz, symmetric=int(h[3]), nnz_diag=int(h[4]), nnz_non_diag=int(h[5]), row_ptr=rp, col=ci, val=va)


def set_threads(k):
    """omp_set_num_threads for every OpenMP library of this process (the oracle and oracle/_ref share libgomp)."""
    C.CDLL("libgomp.so.1").omp_set_num_threads(int(k))


def fnv1a64(a):
    a = np.ascontiguousarray(a)
    return int(lib().oracle_fnv1a64(a.ctypes.data, a.nbytes))


def within_bound(y, y_ref, bound, rel=1e-12):
    """North-star parity criterion: |y - y_ref| <= rel * sum_j |a_ij x_j| per row."""
    return bool(np.all(np.abs(y - y_ref) <= rel * bound))


# ---------------------------------------------------------------- reference (oracle/_ref)
def ref_available():
    return all(os.path.exists(os.path.join(HERE, "_ref", n)) for n in
               ("libref_csr.so", "libref_merge.so", "libref_sellcs32.so", "libref_sellcs64.so", "libref_amg.so"))


_ref = {}


def ref(name):
    if name not in _ref:
        _ref[name] = C.CDLL(os.path.join(HERE, "_ref", "libref_%s.so" % name))
    return _ref[name]


class RefCSR:
    """The reference's CPU CSR plugin (K/csr.cpp) behind its own csr_to_format()/spmv().
    CAUTION: the plugin keeps its per-thread row ranges in file-scope arrays (thread_i_s / thread_i_e, K/csr.cpp:36-44), so
    only the MOST RECENTLY created object computes a complete product: create, use, then create the next."""

    def __init__(self, row_ptr, col_ind, val, n):
        L = ref("csr")
        L.ref_csr_create.restype = C.c_void_p
        L.ref_csr_create.argtypes = [_i32p, _i32p, _f64p, C.c_long, C.c_long, C.c_long]
        L.ref_csr_spmv.argtypes = [C.c_void_p, _f64p, _f64p]
        L.ref_csr_spmv_serial.argtypes = [C.c_void_p, _f64p, _f64p]
        L.ref_csr_format_name.restype = C.c_char_p
        L.ref_csr_format_name.argtypes = [C.c_void_p]
        self.L = L
        self.m = len(row_ptr) - 1
        self.h = L.ref_csr_create(row_ptr, _pad1(col_ind, np.int32), _pad1(val, np.float64), self.m, n, int(row_ptr[-1]))

    def spmv(self, x, y):
        self.L.ref_csr_spmv(self.h, x, y)

    def spmv_serial(self, x, y):
        self.L.ref_csr_spmv_serial(self.h, x, y)

    @property
    def format_name(self):
        return self.L.ref_csr_format_name(self.h).decode()


class RefCSRf32(RefCSR):
    """The `_f` build of the same plugin (oracle/_ref/libref_csr_f.so): double matrix in, float x / y."""

    def __init__(self, row_ptr, col_ind, val, n):
        L = ref("csr_f")
        L.ref_csr_create.restype = C.c_void_p
        L.ref_csr_create.argtypes = [_i32p, _i32p, _f64p, C.c_long, C.c_long, C.c_long]
        L.ref_csr_spmv.argtypes = [C.c_void_p, _f32p, _f32p]
        L.ref_csr_spmv_serial.argtypes = [C.c_void_p, _f32p, _f32p]
        L.ref_csr_format_name.restype = C.c_char_p
        L.ref_csr_format_name.argtypes = [C.c_void_p]
        self.L = L
        self.m = len(row_ptr) - 1
        self.h = L.ref_csr_create(row_ptr, _pad1(col_ind, np.int32), _pad1(val, np.float64), self.m, n, int(row_ptr[-1]))


class RefSELL:
    """The reference's plain SELL plugin (K/sell.cpp) behind its own csr_to_format()/spmv(), plus its layout arrays.
    Like RefCSR it keeps per-thread state in a file-scope array (tds): one live object at a time."""

    LIB, PFX = "sell", "ref_sell_"

    def __init__(self, row_ptr, col_ind, val, n):
        L = ref(self.LIB)
        if self.PFX != "ref_sell_":
            for f in ("create", "spmv", "vec_len", "num_row_clusters", "nnz_ext", "row_cluster_ptr", "ja", "a"):
                setattr(L, "ref_sell_" + f, getattr(L, self.PFX + f))
        L.ref_sell_create.restype = C.c_void_p
        L.ref_sell_create.argtypes = [_i32p, _i32p, _f64p, C.c_long, C.c_long, C.c_long]
        L.ref_sell_spmv.argtypes = [C.c_void_p, _f64p, _f64p]
        L.ref_sell_vec_len.restype = C.c_long
        for f in ("ref_sell_num_row_clusters", "ref_sell_nnz_ext"):
            getattr(L, f).restype = C.c_long
            getattr(L, f).argtypes = [C.c_void_p]
        self.L = L
        self.m = len(row_ptr) - 1
        self.VL = int(L.ref_sell_vec_len())
        self.h = L.ref_sell_create(row_ptr, _pad1(col_ind, np.int32), _pad1(val, np.float64), self.m, n, int(row_ptr[-1]))
        S, total = L.ref_sell_num_row_clusters(self.h), L.ref_sell_nnz_ext(self.h)

        def grab(fn, count, ctype):
            f = getattr(L, fn)
            f.restype = C.POINTER(ctype)
            f.argtypes = [C.c_void_p]
            return np.ctypeslib.as_array(f(self.h), shape=(max(count, 1),))[:count].copy()

        self.row_cluster_ptr = grab("ref_sell_row_cluster_ptr", S + 1, C.c_int32)
        self.ja = grab("ref_sell_ja", total, C.c_int32)
        self.a = grab("ref_sell_a", total, C.c_double)

    def spmv(self, x, y):
        """y needs ceil(m / VL) * VL entries (whole vectors are stored)."""
        self.L.ref_sell_spmv(self.h, x, y)


class RefSELLSorted(RefSELL):
    """The reference's per-thread-sorted SELL plugin (K/sell_sorted.cpp); spmv needs a y of m entries."""
    LIB, PFX = "sell_sorted", "ref_sells_"

    def __init__(self, row_ptr, col_ind, val, n):
        super().__init__(row_ptr, col_ind, val, n)
        for fn in ("permutation", "rev_permutation"):
            f = getattr(self.L, "ref_sells_" + fn)
            f.restype = C.POINTER(C.c_int)
            f.argtypes = [C.c_void_p]
            setattr(self, fn, np.ctypeslib.as_array(f(self.h), shape=(max(self.m, 1),))[:self.m].copy())


def ref_partition(num_workers, sums, N, total):
    L = ref("csr")
    L.ref_partition_prefix_sums.argtypes = [C.c_long, C.c_long, _i32p, C.c_long, C.c_long, C.POINTER(C.c_long), C.POINTER(C.c_long)]
    out = np.zeros((num_workers, 2), np.int64)
    s, e = C.c_long(), C.c_long()
    for p in range(num_workers):
        L.ref_partition_prefix_sums(num_workers, p, sums, N, total, C.byref(s), C.byref(e))
        out[p] = (s.value, e.value)
    return out


def ref_merge_path_search(diagonal, row_ptr, m, nnz):
    L = ref("merge")
    L.ref_merge_path_search.argtypes = [C.c_int32, _i32p, C.c_int32, C.c_int32, _i32p]
    out = np.zeros(2, np.int32)
    L.ref_merge_path_search(diagonal, np.ascontiguousarray(row_ptr[1:]), m, nnz, out)
    return int(out[0]), int(out[1])


def ref_merge_csrmv(T, row_ptr, col_ind, val, x, y_prefill=1.0):
    L = ref("merge")
    L.ref_merge_csrmv.argtypes = [C.c_int, _i32p, _i32p, _f64p, _f64p, _f64p, C.c_int32, C.c_int32]
    m = len(row_ptr) - 1
    y = np.full(m, y_prefill, np.float64)
    L.ref_merge_csrmv(T, row_ptr, col_ind, val, x, y, m, int(row_ptr[m]))
    return y


def ref_sellcs_convert(row_ptr, col_ind, val, n, C_, sigma, index64=True):
    L = ref("sellcs64" if index64 else "sellcs32")
    it = np.int64 if index64 else np.int32
    ip = np.ctypeslib.ndpointer(it, flags="C_CONTIGUOUS")
    L.ref_sellcs_create.restype = C.c_void_p
    L.ref_sellcs_create.argtypes = [C.c_int64, C.c_int64, ip, ip, _f64p, C.c_uint64, C.c_uint64]
    m = len(row_ptr) - 1
    h = L.ref_sellcs_create(m, n, np.ascontiguousarray(row_ptr, it), np.ascontiguousarray(col_ind, it), val, C_, sigma)
    for f in ("ref_sellcs_nslices", "ref_sellcs_padded"):
        getattr(L, f).restype = C.c_int64
        getattr(L, f).argtypes = [C.c_void_p]
    ns, total = L.ref_sellcs_nslices(h), L.ref_sellcs_padded(h)

    def grab(fn, count, ctype, dtype):
        f = getattr(L, fn)
        f.restype = C.POINTER(ctype)
        f.argtypes = [C.c_void_p]
        return np.ctypeslib.as_array(f(h), shape=(max(count, 1),))[:count].astype(dtype, copy=True)

    ict = C.c_int64 if index64 else C.c_int32
    out = dict(row_order=grab("ref_sellcs_row_order", m, ict, np.int32),
               slice_width=grab("ref_sellcs_slice_widths", ns, ict, np.int32),
               slice_ptr=grab("ref_sellcs_slice_pointers", ns + 1, ict, np.int64),
               col=grab("ref_sellcs_column_indices", total, ict, np.int32),
               val=grab("ref_sellcs_values", total, C.c_double, np.float64), C=C_, sigma=sigma, m=m)
    L.ref_sellcs_free.argtypes = [C.c_void_p]
    L.ref_sellcs_free(h)
    return out


def ref_amg_generate(rows, cols, avg, std, dist, placement, bw, skew, neigh, crs, seed):
    L = ref("amg")
    L.ref_amg_generate.restype = C.c_void_p
    L.ref_amg_generate.argtypes = [C.c_long, C.c_long, C.c_double, C.c_double, C.c_char_p, C.c_uint, C.c_char_p,
                                   C.c_double, C.c_double, C.c_double, C.c_double]
    d, p = C.create_string_buffer(dist.encode()), C.create_string_buffer(placement.encode())
    h = L.ref_amg_generate(rows, cols, avg, std, d, seed, p, bw, skew, neigh, crs)
    L.ref_amg_nnz.restype = C.c_int64
    L.ref_amg_nnz.argtypes = [C.c_void_p]
    nnz = L.ref_amg_nnz(h)
    outs = []
    for fn, ct, cnt in (("ref_amg_row_ptr", C.c_int, rows + 1), ("ref_amg_col_ind", C.c_int, nnz), ("ref_amg_values", C.c_double, nnz)):
        f = getattr(L, fn)
        f.restype = C.POINTER(ct)
        f.argtypes = [C.c_void_p]
        outs.append(np.ctypeslib.as_array(f(h), shape=(max(cnt, 1),))[:cnt].copy())
    L.ref_amg_free.argtypes = [C.c_void_p]
    L.ref_amg_free(h)
    return outs[0].astype(np.int32), outs[1].astype(np.int32), outs[2]

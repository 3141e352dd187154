"""Host-side mirror of the reference's per-format plugin interface, over the C-ABI.

Reference (relative to /root/reference/benchmark_code/BENCH/src/spmv_kernels/):
  struct Matrix_Format { format_name, m, n, nnz, mem_footprint, csr_mem_footprint,
                         spmv(x, y), statistics_start(), statistics_print_data(buf, n) }   spmv_kernel.h:9-26
  csr_to_format(row_ptr, col_ind, values, m, n, nnz, symmetric, symmetry_expanded)          spmv_kernel.h:28
Same names, argument meaning and error behaviour; the compiled equivalent that links into the
reference harness is host/b200_plugin.cpp.  All arithmetic happens in libb200spmv.so.
"""
import numpy as np

from . import capi

FORMAT_NAMES = {capi.CSR_ADAPTIVE: "B200_CSR_ADAPTIVE", capi.SELL_C_SIGMA: "B200_SELL_C_SIGMA"}


class MatrixFormat:
    """Mirror of `struct Matrix_Format`; `spmv(x, y)` follows the reference GPU-plugin
    convention when `cache_x` is set (x uploaded on the first call only, GPU_clean/cuda_csr.cu:353-382)
    and otherwise moves x and y across PCIe on every call."""

    def __init__(self, handle, fmt, cache_x=False):
        self._h = handle
        self.format = fmt
        self.m, self.n, self.nnz = handle.m, handle.n, handle.nnz
        st = handle.stats()
        self.mem_footprint = st["mem_footprint"]
        self.csr_mem_footprint = self.nnz * (8 + 4) + (self.m + 1) * 4      # spmv_kernel.h:24
        self.requested_format = fmt
        self.format = handle.format                     # AUTO resolves to CSR or SELL at create time
        self.format_name = ("B200_AUTO:" if fmt == capi.AUTO else "") + FORMAT_NAMES[self.format]
        self.col_panels = int(st["col_panels"])
        self.sigma = int(st["ext_or_sigma"]) if self.format == capi.SELL_C_SIGMA else None
        self.trial_ms = {k[9:]: round(st[k], 5) for k in ("trial_ms_csr", "trial_ms_sell", "trial_ms_sell_alt") if st[k] > 0}
        self.convert_ms = st["convert_ms"]
        self.cache_x = cache_x
        self._x_seen = None
        self.num_loops = 0

    # -- reference interface ------------------------------------------------------------------
    def spmv(self, x, y):
        flags = 0
        if self.cache_x and self._x_seen == x.ctypes.data:
            flags |= capi.SKIP_X_UPLOAD
        self._h.spmv_host(x, y, flags)
        self._x_seen = x.ctypes.data
        self.num_loops += 1

    def statistics_start(self):
        self.num_loops = 0

    def statistics_print_data(self):
        return ""

    # -- device-resident path (iterated SpMV) -------------------------------------------------
    def spmv_device(self, dx, dy, stream=None, **kw):
        self._h.spmv_device(dx, dy, stream, **kw)

    @property
    def handle(self):
        return self._h

    @property
    def gpu_launches_per_spmv(self):
        return self._h.launches_per_spmv

    def compulsory_bytes(self):
        """B_c = 8 nnz + 4 nnz + 4 (m+1) + 8 n + 8 m   (identical for CSR and SELL)."""
        return 12 * self.nnz + 4 * (self.m + 1) + 8 * self.n + 8 * self.m

    def close(self):
        self._h.close()


def statistics_print_labels():
    return ""


def csr_to_format(row_ptr, col_ind, values, m, n, nnz, symmetric=0, symmetry_expanded=0, *, format=capi.CSR_ADAPTIVE,
                  cache_x=False, **opts):
    """Mirror of the reference factory.  Inputs are borrowed for the call only (the harness frees
    them right after, bench.cpp:605-629); the library uploads its own copy."""
    if symmetric and not symmetry_expanded:
        # same message as K/csr.cpp:224-225
        raise capi.B200Error("symmetric matrices have to be expanded to be supported by this format")
    row_ptr = np.ascontiguousarray(row_ptr, np.int32)
    col_ind = np.ascontiguousarray(col_ind, np.int32)
    values = np.ascontiguousarray(values, np.float64)
    if row_ptr.size != m + 1 or col_ind.size < nnz or values.size < nnz:
        raise capi.B200Error("CSR array sizes do not match m / nnz")
    o = capi.default_opts(format=format, **opts)
    h = capi.Handle(row_ptr, col_ind if nnz else None, values if nnz else None, m, n, nnz, o)
    return MatrixFormat(h, format, cache_x=cache_x)


def csr_to_format_device(d_row_ptr, d_col_ind, d_values, m, n, nnz, *, format=capi.CSR_ADAPTIVE, **opts):
    """Same, from arrays already resident in HBM (torch tensors or raw device addresses)."""
    o = capi.default_opts(format=format, **opts)
    h = capi.Handle(d_row_ptr, d_col_ind, d_values, m, n, nnz, o, device_arrays=True)
    return MatrixFormat(h, format)


def coo_to_format(row, col, val, m, n, *, format=capi.CSR_ADAPTIVE, cache_x=False, **opts):
    """The reference's import_file() tail + csr_to_format() for COO input (base-0 host arrays in the reference's entry
    order, symmetry already expanded -- what mtx.read_mtx returns): COO -> CSR on the GPU (b200spmv_create_coo,
    csrc/coo.cu; reference coo_to_csr, lib/storage_formats/csr/csr_gen.c:178-213), then the format."""
    row = np.ascontiguousarray(row, np.int32)
    col = np.ascontiguousarray(col, np.int32)
    val = np.ascontiguousarray(val, np.float64)
    nnz = int(row.size)
    if col.size != nnz or val.size != nnz:
        raise capi.B200Error("COO array sizes differ")
    o = capi.default_opts(format=format, **opts)
    h = capi.Handle(row if nnz else None, col if nnz else None, val if nnz else None, m, n, nnz, o, coo=True)
    return MatrixFormat(h, format, cache_x=cache_x)


class MatrixFormatF32:
    """`struct Matrix_Format` of the reference's `_f` build (make.sh:215-216: ValueType = float): spmv(x, y) takes float32
    vectors; the matrix was handed to csr_to_format_f32 as double (ValueTypeReference) and narrowed on the GPU."""

    def __init__(self, handle):
        self._h = handle
        self.m, self.n, self.nnz = handle.m, handle.n, handle.nnz
        self.format_name = "B200_CSR_F32"
        self.mem_footprint = self.nnz * (4 + 4) + (self.m + 1) * 4          # K/csr.cpp:229 with sizeof(ValueType) = 4
        self.csr_mem_footprint = self.nnz * (4 + 4) + (self.m + 1) * 4
        self.gpu_launches_per_spmv = 2
        self.num_loops = 0

    def spmv(self, x, y):
        if x.dtype != np.float32 or y.dtype != np.float32:
            raise capi.B200Error("the _f build takes float32 vectors (ValueType = float)")
        self._h.spmv_host(x, y)
        self.num_loops += 1

    def spmv_device(self, dx, dy, stream=None):
        self._h.spmv_device(dx, dy, stream)

    def statistics_start(self):
        self.num_loops = 0

    def statistics_print_data(self):
        return ""

    @property
    def handle(self):
        return self._h

    def compulsory_bytes(self):
        """fp32 compulsory bytes: 4 nnz (values) + 4 nnz (col_ind) + 4 (m+1) (row_ptr) + 4 n (x once) + 4 m (y)."""
        return 8 * self.nnz + 4 * (self.m + 1) + 4 * self.n + 4 * self.m

    def close(self):
        self._h.close()


def csr_to_format_f32(row_ptr, col_ind, values, m, n, nnz, symmetric=0, symmetry_expanded=0, *, device=-1):
    """csr_to_format of the `_f` build: same arguments (values are DOUBLE, ValueTypeReference), CSR only."""
    if symmetric and not symmetry_expanded:
        raise capi.B200Error("symmetric matrices have to be expanded to be supported by this format")
    row_ptr = np.ascontiguousarray(row_ptr, np.int32)
    col_ind = np.ascontiguousarray(col_ind, np.int32)
    values = np.ascontiguousarray(values, np.float64)
    if row_ptr.size != m + 1 or col_ind.size < nnz or values.size < nnz:
        raise capi.B200Error("CSR array sizes do not match m / nnz")
    return MatrixFormatF32(capi.HandleF32(row_ptr, col_ind if nnz else None, values if nnz else None, m, n, nnz, device))

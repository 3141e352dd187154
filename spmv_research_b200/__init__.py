"""b200-spmv: B200-native fp64 SpMV engine behind the LiHaoxu/SpMV-Research plugin interface.

Product path = libb200spmv.so (hand-written sm_100a CUDA, C-ABI in include/b200spmv.h).
This package is the host-side mirror of the reference interface plus plumbing.
"""
from . import capi  # noqa: F401
from .formats import MatrixFormat, csr_to_format, csr_to_format_device, coo_to_format, csr_to_format_f32, statistics_print_labels  # noqa: F401
from .capi import CSR_ADAPTIVE, SELL_C_SIGMA, AUTO, B200Error  # noqa: F401

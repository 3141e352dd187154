"""Minimal re-hosting of the reference harness around the plugin boundary:
synthetic import -> csr_to_format -> warm-up -> timed loop -> validation -> CSV-style record.

Reference: benchmark_code/BENCH/src/bench.cpp:242-283,600-633 and bench_spmv.cpp:287-301,335-404.
"""
import os
import time

import numpy as np

from . import amg
from .formats import csr_to_format

_HERE = os.path.dirname(os.path.abspath(__file__))


def load_param_file(path):
    out = []
    for ln in open(path):
        ln = ln.strip()
        if ln and not ln.startswith("#"):
            out.append(amg.MatrixParams.parse(ln))
    return out


def medium_sample():
    """The committed stratified sample of synthetic_matrices_medium_dataset.txt (tools/make_medium_sample.py)."""
    return load_param_file(os.path.join(_HERE, "params", "medium_sample.txt"))


CONFIG1 = amg.MatrixParams.parse("1000000 1000000 10 5 normal random 0.3 0 0.5 0.5 14")
CONFIG4_G1 = amg.MatrixParams.parse("8388607 8388607 50 150 gamma random 0.6 0 0.05 0.05 14")
CONFIG4_G2 = amg.MatrixParams.parse("8388607 8388607 50 16.6667 gamma random 0.6 10000 0.05 0.05 14")


def config5(world, rows_per_gpu=4194304):
    """Weak-scaled series in the style of the large dataset; world = 8 gives the 32M-row, ~2e9-nnz
    matrix of BASELINE config 5 (33554432 rows x 60 nnz/row, nnz < 2^31)."""
    r = rows_per_gpu * world
    return amg.MatrixParams(r, r, 60, 20, "normal", "random", 0.3, 0, 0.5, 0.5, 14)


def compulsory_bytes(m, n, nnz):
    """B_c = 8 nnz (values) + 4 nnz (col_ind) + 4 (m+1) (row_ptr) + 8 n (x once) + 8 m (y)."""
    return 12 * nnz + 4 * (m + 1) + 8 * n + 8 * m


def timed_loop(MF, x, y, min_loops=64, min_runtime=2.0, warmup=1, max_loops=100000):
    """The reference's timing protocol: warm-up call(s), then per-call host-timed spmv until both
    >= min_loops iterations and >= min_runtime seconds; returns (min, median, max) seconds."""
    for _ in range(warmup):
        MF.spmv(x, y)
    times = []
    total = 0.0
    while (total < min_runtime or len(times) < min_loops) and len(times) < max_loops:
        t0 = time.clock_gettime(time.CLOCK_MONOTONIC_RAW)
        MF.spmv(x, y)
        dt = time.clock_gettime(time.CLOCK_MONOTONIC_RAW) - t0
        times.append(dt)
        total += dt
    times.sort()
    return times[0], times[len(times) // 2], times[-1], len(times)


def run_one(params, fmt, x=None, validate=None, **timing):
    """Generate -> convert -> time -> (optionally) validate one matrix; returns a record dict."""
    t0 = time.time()
    rp, col, val = amg.generate(params)
    t_gen = time.time() - t0
    m, n, nnz = params.rows, params.cols, int(rp[-1])
    t0 = time.time()
    MF = csr_to_format(rp, col, val, m, n, nnz, 0, 0, format=fmt, cache_x=True)
    t_conv = time.time() - t0
    if x is None:
        x = np.ones(n)                     # bench_spmv.cpp:601-605
    y = np.full(m + 64, 1.0)               # bench_spmv.cpp:606-609
    tmin, tmed, tmax, loops = timed_loop(MF, x, y, **timing)
    rec = dict(matrix=params.line(), m=m, n=n, nnz=nnz, format_name=MF.format_name, time_generate=t_gen, time_convert=t_conv,
               loops=loops, time_min=tmin, time_median=tmed, time_max=tmax, gflops=2.0 * nnz / tmed * 1e-9,
               eff_GBs=compulsory_bytes(m, n, nnz) / tmed * 1e-9, mem_footprint_MB=MF.mem_footprint / 2 ** 20)
    if validate is not None:
        rec.update(validate(rp, col, val, x, y[:m]))
    MF.close()
    return rec

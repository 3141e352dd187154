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


# synthetic_matrices_medium_dataset.txt (BASELINE config 2) is a grid: 15 row counts per avg-nnz class x bw x skew x
# neighbours x similarity (std = avg / 3 to 4 decimals, normal, random, seed 14), minus 10 lines the reference's file does
# not contain.  Stored as the grid, not as 16 190 lines; tests/test_amg.py checks it against the reference's file as a set
# when /root/reference is present.
_MEDIUM_ROWS = {
    5: [65535, 147455, 229375, 311295, 393215, 524287, 2097151, 3670015, 5242879, 6815743, 8388607, 13418495, 18448383, 23478271, 28508159],
    10: [33825, 76106, 118387, 160668, 202950, 270600, 1082401, 1894201, 2706002, 3517803, 4329604, 6925675, 9521746, 12117817, 14713889],
    20: [17189, 38676, 60164, 81651, 103138, 137518, 550072, 962627, 1375181, 1787736, 2200290, 3519605, 4838920, 6158235, 7477550],
    50: [6944, 15624, 24304, 32985, 41665, 55553, 222214, 388875, 555536, 722198, 888859, 1421827, 1954795, 2487763, 3020732],
    100: [3483, 7838, 12192, 16547, 20901, 27869, 111476, 195083, 278691, 362298, 445906, 713275, 980644, 1248014, 1515383],
    500: [698, 1571, 2445, 3318, 4191, 5588, 22354, 39120, 55886, 72652, 89418, 143035, 196651, 250268, 303884],
}
_MEDIUM_STD = {5: 1.6667, 10: 3.3333, 20: 6.6667, 50: 16.6667, 100: 33.3333, 500: 166.6667}
_MEDIUM_BW, _MEDIUM_SKEW = (0.05, 0.3, 0.6), (0, 100, 1000, 10000)
_MEDIUM_NEIGH, _MEDIUM_CRS = (0.05, 0.5, 0.95, 1.4, 1.9), (0.05, 0.5, 0.95)
_MEDIUM_ABSENT = {(4191, 1.9, 0.05), (4191, 1.9, 0.5), (4191, 1.9, 0.95), (72652, 1.9, 0.05), (72652, 1.9, 0.5), (72652, 1.9, 0.95),
                  (303884, 1.4, 0.95), (303884, 1.9, 0.05), (303884, 1.9, 0.5), (303884, 1.9, 0.95)}      # all at avg 500, bw 0.6, skew 10000


def medium_dataset():
    """All 16 190 parameter lines of the reference's synthetic_matrices_medium_dataset.txt, grouped so that lines sharing
    pass 1 of the generator (same rows / avg / skew) are adjacent."""
    out = []
    for avg, rows_list in _MEDIUM_ROWS.items():
        for rows in rows_list:
            for skew in _MEDIUM_SKEW:
                for bw in _MEDIUM_BW:
                    for neigh in _MEDIUM_NEIGH:
                        for crs in _MEDIUM_CRS:
                            if avg == 500 and bw == 0.6 and skew == 10000 and (rows, neigh, crs) in _MEDIUM_ABSENT:
                                continue
                            out.append(amg.MatrixParams(rows, rows, avg, _MEDIUM_STD[avg], "normal", "random", bw, skew, neigh, crs, 14))
    return out


def validation_twins_range(samples=10, percentage=30):
    """BASELINE config 3: the 3 637 parameter lines of the reference's validation_matrices_10_samples_30_range_twins.txt, in
    the file's order.  The file is an expansion (matrix_generation_parameters/create_validation_twins_param_file.py): for each of
    the 52 twins, `samples` scalings in [1 - percentage %, 1 + percentage %] applied to one feature at a time -- rows and cols
    together (truncated), then avg, std, bw, skew, neigh, crs (rounded to 5 decimals) -- with repeated lines dropped.  Stored as
    the 52 base lines (params/validation_twins_rounded.txt), not as the expansion; tests/test_amg.py checks it line by line
    against the reference's file when /root/reference is present.  Returns [(MatrixParams, text line)]."""
    base = load_param_file(os.path.join(_HERE, "params", "validation_twins_rounded.txt"))
    scale = np.linspace(1 - percentage * 0.01, 1 + percentage * 0.01, num=samples)
    out, seen = [], set()

    def emit(rows, cols, f, name):
        text = " ".join(str(v) for v in (rows, cols, f[0], f[1], "normal", "random", f[2], f[3], f[4], f[5], 14))
        if text not in seen:
            seen.add(text)
            out.append((amg.MatrixParams(rows, cols, float(f[0]), float(f[1]), "normal", "random", float(f[2]), float(f[3]), float(f[4]),
                                         float(f[5]), 14, name), text))

    for b in base:
        f = [np.float64(v) for v in (b.avg, b.std, b.bw, b.skew, b.neigh, b.crs)]
        for x in scale:
            emit(int(b.rows * x), int(b.cols * x), f, b.name)
        for k in range(6):
            for x in scale:
                g = list(f)
                g[k] = np.round(f[k] * x, 5)
                emit(b.rows, b.cols, g, b.name)
    return out


def footprint_class(p):
    """CSR footprint class of SURVEY 8d from the nominal nnz = rows x avg."""
    mb = (12 * p.rows * p.avg + 4 * (p.rows + 1)) / 2 ** 20
    return "<32MB" if mb < 32 else "32-128MB" if mb < 128 else "128-512MB" if mb < 512 else ">=512MB"


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

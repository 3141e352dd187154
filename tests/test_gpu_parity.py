"""Parity tests proper: the CUDA path, called through the C-ABI, against the CPU oracle.

Bar (BASELINE.json north_star): integer outputs bit-exact; y within
|y - y_ref| <= 1e-12 * sum_j |a_ij x_j| per row of the serial CSR kernel.
"""
import numpy as np
import pytest

import matrices as M
from tests.golden.make_golden import MERGE_T, PART_W, SELL_CFG

pytestmark = pytest.mark.gpu
REL = 1e-12


@pytest.fixture(scope="module")
def S(native):
    import spmv_research_b200 as S
    if S.capi.device_count() < 1:
        pytest.fail("no CUDA device visible: the gpu-marked tests must run on a B200")
    return S


def _spmv(S, fmt_kwargs, rp, col, val, n, x):
    m = len(rp) - 1
    MF = S.csr_to_format(rp, col, val, m, n, int(rp[-1]), 0, 0, **fmt_kwargs)
    y = np.full(m + 64, 1.0)               # the harness pre-fills y with 1.0 and over-allocates (bench_spmv.cpp:598-609)
    MF.spmv(x, y)
    assert np.all(y[m:] == 1.0), "wrote past y[m]"
    return MF, y[:m]


def _check(oracle, rp, col, val, x, y, what):
    yr = oracle.csr_spmv(rp, col, val, x)
    bound = oracle.row_abs_bound(rp, col, val, x)
    err = np.abs(y - yr)
    bad = np.nonzero(~(err <= REL * bound))[0]
    assert bad.size == 0, "%s: %d rows off, first row %d: got %r want %r bound %r" % (
        what, bad.size, bad[0], y[bad[0]], yr[bad[0]], bound[bad[0]])


CSR_VARIANTS = [dict(csr_nnz_per_thread=8), dict(csr_nnz_per_thread=16), dict(csr_nnz_per_thread=32),
                dict(csr_nnz_per_thread=8, csr_ext=64), dict(csr_nnz_per_thread=8, x_hint=1)]


# ------------------------------------------------------------------------------------------ CSR
def test_csr_generator_matrices(S, oracle, golden, golden_matrices):
    for li, (p, rp, col, val) in enumerate(golden_matrices):
        for x in (np.ones(p.cols), M.rand_x(p.cols)):
            for kw in CSR_VARIANTS:
                MF, y = _spmv(S, dict(format=S.CSR_ADAPTIVE, **kw), rp, col, val, p.cols, x)
                _check(oracle, rp, col, val, x, y, "%s %r" % (M.GOLDEN_LINES[li], kw))
                MF.close()
        # and against the reference's own y (golden fixture made from K/csr.cpp)
        x = M.rand_x(p.cols)
        MF, y = _spmv(S, dict(format=S.CSR_ADAPTIVE), rp, col, val, p.cols, x)
        bound = oracle.row_abs_bound(rp, col, val, x)
        assert np.all(np.abs(y[::7] - golden["g%d_y_sample" % li]) <= REL * bound[::7])


def test_csr_edge_cases(S, oracle):
    for name, rp, col, val, n in M.edge_cases():
        for x in (np.ones(n), M.rand_x(n, 3)):
            for kw in CSR_VARIANTS:
                MF, y = _spmv(S, dict(format=S.CSR_ADAPTIVE, **kw), rp, col, val, n, x)
                _check(oracle, rp, col, val, x, y, "%s %r" % (name, kw))
                lens = np.diff(rp)
                assert np.all(y[lens == 0] == 0.0), name          # empty rows are written 0
                MF.close()


def test_csr_split_rows_and_tile_plan(S, oracle):
    """The tile plan against a numpy model: first owned row per tile, split rows = rows leaving their
    owner window by more than ext; giant rows are completed by the last of their tiles to finish."""
    for name, rp, col, val, n in M.edge_cases():
        m, nnz = len(rp) - 1, int(rp[-1])
        for npt, ext in ((8, 64), (8, 1024), (16, 2048)):
            T = 256 * npt
            MF = S.csr_to_format(rp, col, val, m, n, nnz, format=S.CSR_ADAPTIVE, csr_nnz_per_thread=npt, csr_ext=ext, csr_compress_rows=1)
            ntiles = max(1, -(-nnz // T))
            want = np.minimum(np.searchsorted(rp, np.arange(ntiles) * T, side="left"), m).astype(np.int32)
            want[0] = 0
            want = np.append(want, m).astype(np.int32)
            got = MF.handle.export(S.capi.X_TILE_ROW)
            assert np.array_equal(got, want), (name, npt)
            split_rows, split_tiles = [], []
            for t in range(ntiles):
                r0, r1 = want[t], want[t + 1]
                if r1 > r0 and int(rp[r1]) - (t + 1) * T > ext:
                    split_rows.append(r1 - 1)
                    split_tiles.append(t)
            assert np.array_equal(MF.handle.export(S.capi.X_SPLIT_ROW), np.array(split_rows, np.int32)), (name, npt, ext)
            assert np.array_equal(MF.handle.export(S.capi.X_SPLIT_TILE), np.array(split_tiles, np.int32))
            # few split rows: finished inside the tile kernel (last-arriving CTA); many: the separate fix-up launch
            assert MF.gpu_launches_per_spmv == (2 if 16 * len(split_rows) > ntiles else 1)
            MF.close()


def test_csr_compressed_rows(S, oracle):
    """Row compression: the plan runs on the non-empty rows, the empty ones are zeroed by their own launch.  Forced on
    every edge case (same y, plan = the numpy model evaluated on the compressed row_ptr and mapped back), and switched on
    by itself when one tile would own a long run of empty rows (the mawi twin's shape: 99 % empty rows)."""
    for name, rp, col, val, n in M.edge_cases():
        m, nnz = len(rp) - 1, int(rp[-1])
        x = M.rand_x(n, 5)
        MF, y = _spmv(S, dict(format=S.CSR_ADAPTIVE, csr_compress_rows=2, csr_ext=64), rp, col, val, n, x)
        _check(oracle, rp, col, val, x, y, "compressed " + name)
        lens = np.diff(rp)
        assert np.all(y[lens == 0] == 0.0), name
        st = MF.handle.stats()
        assert st["compressed_rows"] == 1 and st["n_empty"] == int((lens == 0).sum())
        nz = np.nonzero(lens)[0]
        rp_nz = np.append(rp[nz], nnz)
        T, ext = 2048, 64
        ntiles = max(1, -(-nnz // T))
        k = np.minimum(np.searchsorted(rp_nz, np.arange(ntiles) * T, side="left"), len(nz))
        k[0] = 0
        k = np.append(k, len(nz))
        rows_nz = np.append(nz, m)
        assert np.array_equal(MF.handle.export(S.capi.X_TILE_ROW), rows_nz[k].astype(np.int32)), name
        split = [rows_nz[k[t + 1] - 1] for t in range(ntiles) if k[t + 1] > k[t] and int(rp_nz[k[t + 1]]) - (t + 1) * T > ext]
        assert np.array_equal(MF.handle.export(S.capi.X_SPLIT_ROW), np.array(split, np.int32)), name
        assert MF.gpu_launches_per_spmv == (2 if 16 * len(split) > ntiles else 1) + (1 if st["n_empty"] else 0)
        MF.close()
    # auto: 60 000 empty rows in one run -> one tile would own them all
    rng = np.random.default_rng(3)
    lens = np.concatenate([rng.integers(1, 9, 3000), np.zeros(60000, np.int64), rng.integers(0, 5, 5000), [9000], np.zeros(7000, np.int64)])
    rp, col, val = M.csr_from_lengths(lens, 20000, 4)
    x = M.rand_x(20000, 6)
    MF, y = _spmv(S, dict(format=S.CSR_ADAPTIVE), rp, col, val, 20000, x)
    assert MF.handle.stats()["compressed_rows"] == 1
    _check(oracle, rp, col, val, x, y, "auto-compressed")
    assert np.all(y[lens == 0] == 0.0)
    MF.close()
    MF, y2 = _spmv(S, dict(format=S.CSR_ADAPTIVE, csr_compress_rows=1), rp, col, val, 20000, x)
    assert MF.handle.stats()["compressed_rows"] == 0
    _check(oracle, rp, col, val, x, y2, "not compressed")
    MF.close()


def test_csr_deterministic_and_repeatable(S, golden_matrices):
    p, rp, col, val = golden_matrices[3]                       # giant-row head: exercises the split-row fold
    x = M.rand_x(p.cols)
    MF = S.csr_to_format(rp, col, val, p.rows, p.cols, int(rp[-1]), csr_ext=64)
    assert MF.gpu_launches_per_spmv == 1 and MF.handle.stats()["nsplit"] >= 1
    ys = []
    for _ in range(3):
        y = np.full(p.rows, 1.0)
        MF.spmv(x, y)
        ys.append(y)
    assert np.array_equal(ys[0], ys[1]) and np.array_equal(ys[0], ys[2])      # no atomics: bit-reproducible


def test_csr_short_rows_bit_exact_summation_order(S, oracle):
    """Thread-per-row bins add the products in ascending j like subkernel_csr_scalar; without FMA
    contraction on the CPU side that is bit-identical."""
    name, rp, col, val, n = M.edge_cases()[4]
    x = M.rand_x(n)
    MF, y = _spmv(S, dict(format=S.CSR_ADAPTIVE, csr_nnz_per_thread=8), rp, col, val, n, x)
    seq = np.zeros(len(rp) - 1)
    for i in range(len(rp) - 1):
        s = 0.0
        for j in range(rp[i], rp[i + 1]):
            s += val[j] * x[col[j]]
        seq[i] = s
    full = rp[:-1] < (int(rp[-1]) // 2048) * 2048          # rows owned by tiles that are full (> 128 rows each here)
    assert full.sum() > 4000 and np.array_equal(y[full], seq[full])


# ------------------------------------------------------------------------------------------ SELL-C-sigma
def test_sell_conversion_bit_exact(S, oracle, golden, golden_matrices):
    keys = ((S.capi.X_ROW_ORDER, "row_order"), (S.capi.X_SLICE_WIDTH, "slice_width"), (S.capi.X_SLICE_PTR, "slice_ptr"),
            (S.capi.X_SELL_COL, "col"), (S.capi.X_SELL_VAL, "val"))
    for li, (p, rp, col, val) in enumerate(golden_matrices):
        for (C_, sg) in SELL_CFG:
            MF = S.csr_to_format(rp, col, val, p.rows, p.cols, int(rp[-1]), format=S.SELL_C_SIGMA, sell_C=C_, sell_sigma=sg)
            O = oracle.sellcs_convert(rp, col, val, C_, sg)
            fnv = []
            for sel, k in keys:
                got = MF.handle.export(sel)
                assert np.array_equal(got, O[k]), (M.GOLDEN_LINES[li], C_, sg, k)
                fnv.append(oracle.fnv1a64(got))
            assert np.array_equal(np.array(fnv, np.uint64), golden["g%d_sell_%d_%d" % (li, C_, sg)])   # = the reference library's output
            MF.close()


def test_sell_conversion_edge_cases(S, oracle):
    for name, rp, col, val, n in M.edge_cases():
        m = len(rp) - 1
        for (C_, sg) in ((32, 16384), (2, 2), (8, 24), (256, 256), (64, 8192)):
            MF = S.csr_to_format(rp, col, val, m, n, int(rp[-1]), format=S.SELL_C_SIGMA, sell_C=C_, sell_sigma=sg)
            O = oracle.sellcs_convert(rp, col, val, C_, sg)
            for sel, k in ((S.capi.X_ROW_ORDER, "row_order"), (S.capi.X_SLICE_WIDTH, "slice_width"), (S.capi.X_SLICE_PTR, "slice_ptr"),
                           (S.capi.X_SELL_COL, "col"), (S.capi.X_SELL_VAL, "val")):
                assert np.array_equal(MF.handle.export(sel), O[k]), (name, C_, sg, k)
            for x in (np.ones(n), M.rand_x(n, 5)):
                y = np.full(m, 1.0)
                MF.spmv(x, y)
                _check(oracle, rp, col, val, x, y, "sell %s C=%d sigma=%d" % (name, C_, sg))
                if len(O["slice_width"]) == 0 or int(np.max(O["slice_width"])) <= 128:
                    # slices no wider than one k chunk (SELL_KCH): same fma chain per lane as the reference kernel -> bit-equal;
                    # wider slices are cut into chunks whose partials are added in chunk order (bounded by _check above)
                    assert np.array_equal(y, oracle.sellcs_spmv(O, x)), name
            MF.close()


def test_sell_spmv_generator_matrices(S, oracle, golden_matrices):
    for (p, rp, col, val) in golden_matrices:
        for x in (np.ones(p.cols), M.rand_x(p.cols)):
            for kw in (dict(), dict(sell_C=64, sell_sigma=4096), dict(sell_C=4, sell_sigma=64, x_hint=1)):
                MF, y = _spmv(S, dict(format=S.SELL_C_SIGMA, **kw), rp, col, val, p.cols, x)
                _check(oracle, rp, col, val, x, y, "sell %r" % kw)
                MF.close()


def test_sell_large_sigma_global_sort(S, oracle):
    """sigma larger than the shared-memory sort block (4096 keys) takes the global bitonic steps."""
    rng = np.random.default_rng(21)
    rp, col, val = M.csr_from_lengths(rng.integers(0, 40, 70000), 5000, 21)
    for sg in (8192, 16384, 65536, 1 << 20):
        MF = S.csr_to_format(rp, col, val, 70000, 5000, int(rp[-1]), format=S.SELL_C_SIGMA, sell_C=32, sell_sigma=sg)
        O = oracle.sellcs_convert(rp, col, val, 32, sg)
        assert np.array_equal(MF.handle.export(S.capi.X_ROW_ORDER), O["row_order"]), sg
        assert np.array_equal(MF.handle.export(S.capi.X_SLICE_PTR), O["slice_ptr"]), sg
        MF.close()


def test_col_range_and_band_upload(S, oracle, golden_matrices):
    """b200spmv_col_range is the exact min / max column (+1); the host entry point uploads only that band of x, so
    x entries outside it must not influence y (NaN there stays invisible) while entries inside do."""
    for (p, rp, col, val) in golden_matrices:
        for fmt in (S.CSR_ADAPTIVE, S.SELL_C_SIGMA):
            MF = S.csr_to_format(rp, col, val, p.rows, p.cols, int(rp[-1]), format=fmt)
            lo, hi = MF.handle.col_range()
            assert (lo, hi) == (int(col.min()), int(col.max()) + 1)
            x = M.rand_x(p.cols)
            xn = x.copy()
            xn[:lo] = np.nan
            xn[hi:] = np.nan
            if fmt == S.SELL_C_SIGMA:
                xn[0] = x[0]          # SELL padding entries are (column 0, 0.0) and multiply x[0], as in the reference
            y = np.full(p.rows, 1.0)
            MF.spmv(xn, y)
            _check(oracle, rp, col, val, x, y, "band upload")
            MF.close()
    # out-of-range column indices are refused at create time
    rp = np.array([0, 1], np.int32)
    with pytest.raises(S.B200Error):
        S.csr_to_format(rp, np.array([5], np.int32), np.ones(1), 1, 3, 1)


# ------------------------------------------------------------------------------------------ AUTO
def test_auto_format_resolves_and_matches(S, oracle, golden_matrices):
    """B200SPMV_AUTO builds both formats, times them on the device and keeps one: whatever it keeps must pass the
    same per-row bound, export the layout of the kept format bit-exactly, and report its trial times."""
    for (p, rp, col, val) in golden_matrices:
        x = M.rand_x(p.cols)
        MF, y = _spmv(S, dict(format=S.AUTO), rp, col, val, p.cols, x)
        assert MF.format in (S.CSR_ADAPTIVE, S.SELL_C_SIGMA) and MF.format_name.startswith("B200_AUTO:")
        assert "csr" in MF.trial_ms and MF.trial_ms["csr"] > 0
        _check(oracle, rp, col, val, x, y, "auto -> %s" % MF.format_name)
        if MF.format == S.SELL_C_SIGMA:
            O = oracle.sellcs_convert(rp, col, val, 32, MF.sigma)
            assert np.array_equal(MF.handle.export(S.capi.X_ROW_ORDER), O["row_order"])
            assert np.array_equal(MF.handle.export(S.capi.X_SLICE_PTR), O["slice_ptr"])
        # the reference GPU-plugin convention on an AUTO handle: the first call must upload x even when asked to skip
        y2 = np.full(p.rows, 7.0)
        MF.handle.spmv_host(x, y2, 0)
        assert np.array_equal(y2, y)
        MF.close()
    # empty matrix and a matrix of empty rows resolve without timing anything
    rp = np.zeros(6, np.int32)
    MF = S.csr_to_format(rp, np.zeros(0, np.int32), np.zeros(0), 5, 5, 0, format=S.AUTO)
    y = np.full(5, 3.0)
    MF.spmv(np.ones(5), y)
    assert np.all(y == 0.0)
    MF.close()


def test_auto_skip_flag_on_first_call(S, oracle, golden_matrices):
    p, rp, col, val = golden_matrices[0]
    x = M.rand_x(p.cols)
    MF = S.csr_to_format(rp, col, val, p.rows, p.cols, int(rp[-1]), format=S.AUTO)
    y = np.full(p.rows, 1.0)
    MF.handle.spmv_host(x, y, S.capi.SKIP_X_UPLOAD)          # nothing staged yet: must upload anyway
    _check(oracle, rp, col, val, x, y, "auto first call with SKIP_X_UPLOAD")
    MF.close()


# ------------------------------------------------------------------------------------------ partitioners
def test_partition_bit_exact(S, oracle, golden, golden_matrices):
    for li, (p, rp, col, val) in enumerate(golden_matrices):
        for W in PART_W:
            got = S.capi.partition(rp, p.rows, int(rp[-1]), W)
            assert np.array_equal(got, golden["g%d_part%d" % (li, W)]), (li, W)
    rng = np.random.default_rng(3)
    for _ in range(100):
        m = int(rng.integers(1, 500))
        rp = np.zeros(m + 1, np.int32)
        np.cumsum(rng.integers(0, 9, m) * (rng.random(m) < rng.random()), out=rp[1:])
        rp += int(rng.integers(0, 5))
        W = int(rng.integers(1, 40))
        assert np.array_equal(S.capi.partition(rp, m, int(rp[-1] - rp[0]), W), oracle.partition(W, rp, m, int(rp[-1] - rp[0])))
    with pytest.raises(S.B200Error, match="Empty Sums array"):
        S.capi.partition(np.zeros(1, np.int32), 0, 0, 2)


def test_merge_coords_bit_exact(S, oracle, golden, golden_matrices):
    for li, (p, rp, col, val) in enumerate(golden_matrices):
        MF = S.csr_to_format(rp, col, val, p.rows, p.cols, int(rp[-1]))
        for T in MERGE_T:
            assert np.array_equal(MF.handle.merge_coords(T), golden["g%d_merge%d" % (li, T)]), (li, T)
        T = 148 * 4
        _, want = oracle.merge_csrmv(T, rp, col, val, np.ones(p.cols))
        assert np.array_equal(MF.handle.merge_coords(T), want)
        MF.close()


# ------------------------------------------------------------------------------------------ device-resident path
def test_device_path_scale_norm_peers_shard(S, oracle, golden_matrices):
    import torch
    p, rp, col, val = golden_matrices[1]
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    x = M.rand_x(p.cols)
    dx = torch.from_numpy(x).to(dev)
    yr = oracle.csr_spmv(rp, col, val, x)
    bound = oracle.row_abs_bound(rp, col, val, x)
    for fmt in (S.CSR_ADAPTIVE, S.SELL_C_SIGMA):
        MF = S.csr_to_format(rp, col, val, p.rows, p.cols, int(rp[-1]), format=fmt)
        dy = torch.full((p.rows,), 1.0, dtype=torch.float64, device=dev)
        peers = [torch.zeros(p.rows, dtype=torch.float64, device=dev) for _ in range(3)]
        scale = torch.tensor([0.5], dtype=torch.float64, device=dev)
        norm2 = torch.zeros(1, dtype=torch.float64, device=dev)
        MF.spmv_device(dx, dy, st, scale=scale, norm2=norm2, peers=peers)
        torch.cuda.synchronize()
        y = dy.cpu().numpy()
        assert np.all(np.abs(y - 0.5 * yr) <= REL * bound)
        for pb in peers:
            assert torch.equal(pb, dy)
        assert abs(norm2.item() - float(np.sum(y * y))) <= 1e-12 * float(np.sum(y * y))
        MF.close()
    # a row shard built from device arrays with a non-zero row_ptr offset
    a, b = p.rows // 3, (2 * p.rows) // 3 + 7
    d_rp = torch.from_numpy(rp[a:b + 1].copy()).to(dev)
    d_col = torch.from_numpy(col[rp[a]:rp[b]].copy()).to(dev)
    d_val = torch.from_numpy(val[rp[a]:rp[b]].copy()).to(dev)
    for fmt in (S.CSR_ADAPTIVE, S.SELL_C_SIGMA):
        MF = S.csr_to_format_device(d_rp, d_col, d_val, b - a, p.cols, int(rp[b] - rp[a]), format=fmt)
        dy = torch.full((b - a,), 1.0, dtype=torch.float64, device=dev)
        MF.spmv_device(dx, dy, st)
        torch.cuda.synchronize()
        assert np.all(np.abs(dy.cpu().numpy() - yr[a:b]) <= REL * bound[a:b])
        MF.close()


def test_column_panels(S, oracle, golden_matrices):
    """Column panels (x band larger than L2): the handle is a shell around P sub-handles launched in ascending column
    order, y accumulated, the last panel finishing the rows.  Forced here on small matrices: same per-row bound for every
    format, every panel count, the edge cases (empty rows, giant rows, rectangular), the fused epilogue and PCG."""
    import torch
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    cases = [(p.cols, rp, col, val) for (p, rp, col, val) in golden_matrices] + [(n, rp, col, val) for _, rp, col, val, n in M.edge_cases()]
    for n, rp, col, val in cases:
        m, nnz = len(rp) - 1, int(rp[-1])
        x = M.rand_x(n, 11)
        for fmt in (S.CSR_ADAPTIVE, S.SELL_C_SIGMA, S.AUTO):
            for P in (2, 5):
                MF, y = _spmv(S, dict(format=fmt, col_panels=P), rp, col, val, n, x)
                _check(oracle, rp, col, val, x, y, "panels %d fmt %d" % (P, fmt))
                st_ = MF.handle.stats()
                assert st_["col_panels"] == (P if nnz > 0 and m > 0 else 1)
                if nnz > 0:
                    with pytest.raises(S.B200Error):
                        MF.handle.export(S.capi.X_TILE_ROW)
                MF.close()
    # fused epilogue through the last panel: y = s A x, norm, peer destinations limited to a row range
    p, rp, col, val = golden_matrices[1]
    x = M.rand_x(p.cols)
    dx = torch.from_numpy(x).to(dev)
    yr = oracle.csr_spmv(rp, col, val, x)
    bound = oracle.row_abs_bound(rp, col, val, x)
    for fmt in (S.CSR_ADAPTIVE, S.SELL_C_SIGMA):
        MF = S.csr_to_format(rp, col, val, p.rows, p.cols, int(rp[-1]), format=fmt, col_panels=3)
        dy = torch.full((p.rows,), 1.0, dtype=torch.float64, device=dev)
        peer = torch.zeros(p.rows, dtype=torch.float64, device=dev)
        scale = torch.tensor([0.25], dtype=torch.float64, device=dev)
        norm2 = torch.zeros(1, dtype=torch.float64, device=dev)
        lo, hi = p.rows // 4, p.rows // 2
        MF.spmv_device(dx, dy, st, scale=scale, norm2=norm2, peers=[peer], peer_rows=(lo, hi))
        torch.cuda.synchronize()
        y = dy.cpu().numpy()
        assert np.all(np.abs(y - 0.25 * yr) <= REL * bound)
        pb = peer.cpu().numpy()
        assert np.array_equal(pb[lo:hi], y[lo:hi]) and np.all(pb[:lo] == 0) and np.all(pb[hi:] == 0)
        assert abs(norm2.item() - float(np.sum(y * y))) <= 1e-12 * float(np.sum(y * y))
        MF.close()
    # auto rule: small bands never split
    MF = S.csr_to_format(rp, col, val, p.rows, p.cols, int(rp[-1]))
    assert MF.handle.stats()["col_panels"] == 1
    MF.close()
    # panels whose rows are compressed (rows empty IN A PANEL go through the zero-rows pass, which must keep the earlier panels'
    # partial, apply the scale, feed the peers and the norm): forced compression, plain and fused epilogue
    for n, rp2, col2, val2 in cases[:3] + cases[-4:]:
        m2, nnz2 = len(rp2) - 1, int(rp2[-1])
        if nnz2 == 0 or m2 == 0:
            continue
        x2 = M.rand_x(n, 12)
        MF, y = _spmv(S, dict(format=S.CSR_ADAPTIVE, col_panels=3, csr_compress_rows=2), rp2, col2, val2, n, x2)
        _check(oracle, rp2, col2, val2, x2, y, "compressed panels")
        dx2 = torch.from_numpy(x2).to(dev)
        dy = torch.full((m2,), 7.0, dtype=torch.float64, device=dev)
        peer = torch.zeros(m2, dtype=torch.float64, device=dev)
        scale = torch.tensor([-0.5], dtype=torch.float64, device=dev)
        norm2 = torch.zeros(1, dtype=torch.float64, device=dev)
        MF.spmv_device(dx2, dy, st, scale=scale, norm2=norm2, peers=[peer])
        torch.cuda.synchronize()
        y = dy.cpu().numpy()
        yr2 = oracle.csr_spmv(rp2, col2, val2, x2)
        assert np.all(np.abs(y + 0.5 * yr2) <= REL * oracle.row_abs_bound(rp2, col2, val2, x2))
        assert np.array_equal(peer.cpu().numpy(), y)
        assert abs(norm2.item() - float(np.sum(y * y))) <= 1e-12 * max(float(np.sum(y * y)), 1e-300)
        MF.close()
    # a malformed row_ptr is refused at create time (it would send the kernels out of bounds)
    bad = rp.copy()
    bad[5] = bad[6] + 3
    with pytest.raises(S.B200Error, match="malformed"):
        S.csr_to_format(bad, col, val, p.rows, p.cols, int(rp[-1]))
    # PCG over a panel shell
    rps, cols, vals = M.spd_matrix(20000, 6, 8)
    b = np.random.default_rng(9).uniform(-1, 1, 20000)
    x_ref, ref = oracle.pcg(rps, cols, vals, b, 500)
    MF = S.csr_to_format(rps, cols, vals, 20000, 20000, int(rps[-1]), keep_diag=1, col_panels=2)
    xs, res = MF.handle.pcg(b, 500)
    assert res["converged"] == 1 and np.linalg.norm(xs - x_ref) <= 1e-11 * np.linalg.norm(x_ref)
    MF.close()


def test_host_path_ranges_with_split_rows_and_cut_slices(S, oracle):
    """The host-vector path launches matrices of >= 262 144 rows as 4 unit ranges so that the D2H copy of a range overlaps the
    next range's kernel.  Since round 2 that also holds with split rows (finished by the last of their tiles -- possibly in a
    LATER range: their y entry is copied once more at the end) and with cut SELL slices (a range holds whole slices).  Giant
    rows are planted exactly where the range boundaries fall in nnz space; rows of 300 make every SELL slice a cut one."""
    m, n = 300000, 300000
    rng = np.random.default_rng(21)
    for base_len, giants in ((10, {74000: 60000, 150100: 9000, 226000: 2500}), (300, {75000: 200000, 149990: 5000})):
        lens = np.full(m, base_len, np.int64)
        for r, l in giants.items():
            lens[r] = l
        rp = np.zeros(m + 1, np.int32)
        np.cumsum(lens, out=rp[1:])
        nnz = int(rp[-1])
        col = ((np.arange(nnz, dtype=np.int64) * 2654435761) % n).astype(np.int32)
        val = rng.uniform(0.0, 1.0, nnz)
        x = M.rand_x(n, 6)
        for fmt in (S.CSR_ADAPTIVE, S.SELL_C_SIGMA):
            if fmt == S.SELL_C_SIGMA and base_len == 10:
                continue                              # a 60 000-wide slice pads 32 x: nothing to learn
            MF, y = _spmv(S, dict(format=fmt), rp, col, val, n, x)
            _check(oracle, rp, col, val, x, y, "ranges fmt %d base %d" % (fmt, base_len))
            if fmt == S.CSR_ADAPTIVE:
                assert MF.handle.stats()["nsplit"] >= len(giants)
            y2 = np.full(m + 64, 1.0)
            MF.spmv(x, y2)                            # counters reset by the finisher: a second product gives the same bits
            assert np.array_equal(y2[:m], y)
            MF.close()


def test_host_flags_reference_convention(S, oracle, golden_matrices):
    """cache_x=True follows the reference GPU plugins: x is uploaded on the first call only
    (GPU_clean/cuda_csr.cu:353-382), so a changed x at the same address is NOT seen."""
    p, rp, col, val = golden_matrices[0]
    x = M.rand_x(p.cols)
    MF = S.csr_to_format(rp, col, val, p.rows, p.cols, int(rp[-1]), cache_x=True)
    y1, y2 = np.full(p.rows, 1.0), np.full(p.rows, 1.0)
    MF.spmv(x, y1)
    x *= 2.0
    MF.spmv(x, y2)
    assert np.array_equal(y1, y2)
    MF2 = S.csr_to_format(rp, col, val, p.rows, p.cols, int(rp[-1]))
    MF2.spmv(x, y2)
    assert np.array_equal(y2, 2.0 * y1)


def test_pin_host_vectors(S, oracle, golden_matrices):
    """pin_host_vectors page-locks the caller's pageable x / y on first sight and releases them at destroy: same y, the
    buffers stay usable by the host and by a second handle afterwards."""
    p, rp, col, val = golden_matrices[1]
    x = M.rand_x(p.cols)
    y = np.full(p.rows + 64, 1.0)
    for rep in range(2):
        MF = S.csr_to_format(rp, col, val, p.rows, p.cols, int(rp[-1]), format=S.CSR_ADAPTIVE, pin_host_vectors=1)
        for _ in range(3):
            y[:] = 1.0
            MF.spmv(x, y)
            _check(oracle, rp, col, val, x, y[:p.rows], "pinned host vectors")
            assert np.all(y[p.rows:] == 1.0)
        MF.close()
        x = x * 1.0 if rep == 0 else x          # a fresh (different) buffer for the second handle


def test_bad_options_fail_loudly(S, golden_matrices):
    p, rp, col, val = golden_matrices[0]
    SELL = dict(format=S.SELL_C_SIGMA)
    for kw in (dict(SELL, sell_C=24), dict(SELL, sell_C=512), dict(SELL, sell_C=32, sell_sigma=48), dict(csr_nnz_per_thread=12),
               dict(csr_ext=100), dict(format=7), dict(device=99)):
        with pytest.raises(S.B200Error):
            S.csr_to_format(rp, col, val, p.rows, p.cols, int(rp[-1]), **kw)


# ------------------------------------------------------------------------------------------ full sizes
def test_config1_full_size(S, oracle):
    """BASELINE config 1 (1M x 1M, nnz 10 038 177): full comparison with the oracle, CSR and SELL,
    x = 1 (the harness's vector) and random x; plus linearity A(ax + bz) = aAx + bAz."""
    from spmv_research_b200 import amg
    p = amg.MatrixParams.parse("1000000 1000000 10 5 normal random 0.3 0 0.5 0.5 14")
    rp, col, val = amg.generate(p)
    assert int(rp[-1]) == 10038177
    x, z = M.rand_x(p.cols, 1), M.rand_x(p.cols, 2)
    for fmt in (S.CSR_ADAPTIVE, S.SELL_C_SIGMA):
        MF = S.csr_to_format(rp, col, val, p.rows, p.cols, int(rp[-1]), format=fmt)
        ys = {}
        for name, v in (("ones", np.ones(p.cols)), ("x", x), ("z", z), ("lin", 0.75 * x - 1.5 * z)):
            y = np.full(p.rows, 1.0)
            MF.spmv(v, y)
            _check(oracle, rp, col, val, v, y, "config1 %s fmt %d" % (name, fmt))
            ys[name] = y
        babs = oracle.row_abs_bound(rp, col, val, np.abs(x) + np.abs(z))
        assert np.all(np.abs(ys["lin"] - (0.75 * ys["x"] - 1.5 * ys["z"])) <= 4e-12 * babs)
        MF.close()


def test_giant_row_full_size(S, oracle):
    """G2-style skew at 200k rows (row 0 has 50 100 non-zeros; SURVEY App. C fingerprint line 2)."""
    from spmv_research_b200 import amg
    p = amg.MatrixParams.parse("200000 200000 50 16.6667 gamma random 0.6 1000 0.05 0.05 14")
    rp, col, val = amg.generate(p)
    assert int(rp[-1]) == 10050050 and int(np.max(np.diff(rp))) == 50100
    x = M.rand_x(p.cols, 4)
    for kw in (dict(format=S.CSR_ADAPTIVE), dict(format=S.CSR_ADAPTIVE, csr_nnz_per_thread=32), dict(format=S.SELL_C_SIGMA)):
        MF, y = _spmv(S, kw, rp, col, val, p.cols, x)
        _check(oracle, rp, col, val, x, y, "giant %r" % kw)
        if kw["format"] == S.CSR_ADAPTIVE:
            assert MF.gpu_launches_per_spmv == 1 and MF.handle.stats()["nsplit"] >= 1
        MF.close()


@pytest.mark.gpu
def test_reference_harness_drives_the_plugin(S, oracle, tmp_path):
    """Drop-in proof: the reference's unmodified harness (bench.cpp + bench_spmv.cpp, built by oracle/Makefile) linked with
    host/b200_plugin.cpp loads a Matrix Market file, calls csr_to_format + spmv(x, y) in its own timing loop and validates
    y against its _Float128 Kahan gold (bench_spmv.cpp:108-235): no "Test failed", tiny mae, our format name in its CSV row."""
    from spmv_research_b200 import amg
    import refharness as RH
    if not RH.available("spmv_b200_csr_d.exe"):
        pytest.skip("oracle/_ref harness executables not built (needs /root/reference at build time)")
    p = amg.MatrixParams.parse("30000 30000 12 6 normal random 0.3 50 0.5 0.5 14")
    rp, col, val = amg.generate(p)
    path = str(tmp_path / "amg_30000.mtx")
    RH.write_matrix_market(path, rp, col, val, p.cols)
    for exe, name in (("spmv_b200_csr_d.exe", "B200_CSR_ADAPTIVE"), ("spmv_b200_sell_d.exe", "B200_SELL_C_SIGMA"),
                      ("spmv_b200_auto_d.exe", "B200_AUTO_")):
        r = RH.run(exe, [path], gpu_kernel=True, threads=4)
        assert r["returncode"] == 0, r["stderr"][-2000:]
        assert not r["test_failed"], r["stdout"][-2000:]
        assert any(name in f for f in r["csv"]), r["csv"]
        assert r["errors"]["max_ae"] < 1e-11 and r["errors"]["mae"] < 1e-12, r["errors"]
        assert r["loops"] >= 64                       # the reference's >= 64 iterations / >= 2 s rule (bench_spmv.cpp:335-382)

"""The product's host generator (csrc/amg.c) against the reference generator: golden fingerprints
made from oracle/_ref/libref_amg.so, shard-wise generation, parameter parsing, error behaviour."""
import numpy as np
import pytest

import matrices as M
from spmv_research_b200 import amg


def test_generator_fingerprints_vs_golden(golden, golden_matrices):
    for li, (p, rp, col, val) in enumerate(golden_matrices):
        assert int(rp[-1]) == int(golden["g%d_nnz" % li])
        got = np.array([amg.fnv1a64(rp), amg.fnv1a64(col), amg.fnv1a64(val)], np.uint64)
        assert np.array_equal(got, golden["g%d_fnv" % li]), M.GOLDEN_LINES[li]


def test_config1_shape(native):
    """BASELINE config 1 (1M x 1M, avg 10, std 5, normal, random, seed 14): nnz and max row of SURVEY App. C."""
    p = amg.MatrixParams.parse("1000000 1000000 10 5 normal random 0.3 0 0.5 0.5 14")
    rp, bw = amg.row_ptr(p)
    assert int(rp[-1]) == 10038177 and int(np.max(np.diff(rp))) == 34 and int(np.sum(np.diff(rp) == 0)) == 28580


def test_shard_generation_matches_whole(golden_matrices):
    """Rows [i_s, i_e) generated on their own equal the slice of the whole matrix (any cut points,
    not only anchor boundaries): what each rank of a row-sharded run relies on."""
    for (p, rp, col, val) in golden_matrices[:5]:
        _, bw = amg.row_ptr(p)
        cuts = [0, 1, p.rows // 3 + 5, p.rows // 2, p.rows - 1, p.rows]
        for a, b in zip(cuts[:-1], cuts[1:]):
            c, v = amg.fill(p, rp, bw, a, b)
            assert np.array_equal(c, col[rp[a]:rp[b]]) and np.array_equal(v, val[rp[a]:rp[b]])


def test_csr_wellformed(golden_matrices):
    for (p, rp, col, val) in golden_matrices:
        assert rp[0] == 0 and np.all(np.diff(rp) >= 0)
        assert col.min() >= 0 and col.max() < p.cols and np.all((val >= 0) & (val < 1))
        for i in np.random.default_rng(1).integers(0, p.rows, 200):
            assert np.all(np.diff(col[rp[i]:rp[i + 1]]) > 0)           # sorted, unique columns per row


def test_vs_reference_generator_direct(oracle):
    if not oracle.ref_available():
        pytest.skip("oracle/_ref not built")
    for line in ["1000 1000 5 1.6667 normal random 0.001 0 0.05 0.5 14", "5000 5000 20 6 normal simple 0.05 0 1.9 0.95 3",
                 "2500 4000 8 30 gamma diagonal 0.9 3 0.0 0.0 99", "700 700 3 0 normal random 2.0 0 0 0 1"]:
        p = amg.MatrixParams.parse(line)
        a = amg.generate(p)
        b = oracle.ref_amg_generate(p.rows, p.cols, p.avg, p.std, p.distribution, p.placement, p.bw, p.skew, p.neigh, p.crs, p.seed)
        assert all(np.array_equal(x, y) for x, y in zip(a, b)), line


def test_errors_and_parsing(native):
    with pytest.raises(RuntimeError):
        amg.row_ptr(amg.MatrixParams(50, 50, 5, 1))                      # reference divides by nr_rows/100
    with pytest.raises(ValueError):
        amg.MatrixParams.parse("1 2 3")
    p = amg.MatrixParams.parse("17099 17099 5.6 4.39 normal random 0.29 61.9 0.8 0.63 14 scircuit")
    assert p.name == "scircuit" and p.seed == 14 and amg.MatrixParams.parse(p.line()).rows == 17099


def test_bw_corrected_alone_equals_pass1(native):
    """b200amg_bw_corrected(p, nnz) is pass 1's second output (a sweep shares row_ptr between lines that differ in bw /
    neighbours / similarity only)."""
    for line in M.GOLDEN_LINES[:4]:
        p = amg.MatrixParams.parse(line)
        rp, bw = amg.row_ptr(p)
        assert amg.bw_corrected(p, int(rp[-1])) == bw


def test_medium_dataset_grid_is_the_reference_file():
    """harness.medium_dataset() (a grid description) is, as a set, the reference's synthetic_matrices_medium_dataset.txt."""
    import collections
    import os
    from spmv_research_b200 import harness as H
    ds = H.medium_dataset()
    assert len(ds) == 16190 and len(set(p.line() for p in ds)) == 16190
    cls = collections.Counter(H.footprint_class(p) for p in ds)
    assert (cls["<32MB"], cls["32-128MB"], cls["128-512MB"], cls[">=512MB"]) == (6477, 900, 4497, 4316)
    ref = "/root/reference/matrix_generation_parameters/synthetic_matrices_medium_dataset.txt"
    if not os.path.exists(ref):
        pytest.skip("reference not mounted (authoring container only)")
    want = set(amg.MatrixParams.parse(l).line() for l in open(ref) if l.strip())
    assert want == set(p.line() for p in ds)


TWINS_RANGE_SHA16 = "77a0d5396005bfcd"


def test_validation_twins_range_is_the_reference_file():
    """harness.validation_twins_range() (52 base lines expanded the way create_validation_twins_param_file.py does) is the
    reference's validation_matrices_10_samples_30_range_twins.txt (BASELINE config 3), line by line and in order."""
    import hashlib
    import os
    from spmv_research_b200 import harness as H
    ds = H.validation_twins_range()
    text = [t for _, t in ds]
    assert len(ds) == 3637 and len(set(text)) == 3637                      # 52 x 7 x 10 minus the 3 repeated lines
    assert max(p.rows for p, _ in ds) > 70_000_000 and all(p.seed == 14 and p.rows >= 100 for p, _ in ds)
    # fingerprint of the expansion (holds where the reference is not mounted, e.g. on the GPU box)
    assert hashlib.sha256("\n".join(text).encode()).hexdigest()[:16] == TWINS_RANGE_SHA16
    for p, t in ds[::97]:                                                   # the parsed parameters are the text's
        q = amg.MatrixParams.parse(t)
        assert (q.rows, q.cols, q.avg, q.std, q.bw, q.skew, q.neigh, q.crs) == (p.rows, p.cols, p.avg, p.std, p.bw, p.skew, p.neigh, p.crs)
    ref = "/root/reference/matrix_generation_parameters/validation_matrices_10_samples_30_range_twins.txt"
    if not os.path.exists(ref):
        pytest.skip("reference not mounted (authoring container only)")
    assert [" ".join(l.split()) for l in open(ref) if l.strip()] == text

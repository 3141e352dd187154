"""b200spmv_partition_device_noalloc: the reference partitioner with its scratch in a static device array (no cudaMalloc
inside the call) -- an entry point without a counterpart in the reference, added for
callers that cannot allocate (CUDA graph capture, serving loops).  Called through the C-ABI and compared with the existing
entry point b200spmv_partition_device: integer row bounds, so BIT-EXACT (the bar of every partition test here), on calls
in a row that share the one scratch array (a large W first, so that the later, smaller calls find stale contents in it).
Collected last on purpose (file name)."""
import numpy as np
import pytest

pytestmark = [pytest.mark.gpu]


def test_partition_noalloc_matches_existing_entry_point(native, oracle):
    import torch
    from spmv_research_b200 import capi
    if capi.device_count() < 1:
        pytest.fail("no CUDA device visible: the gpu-marked tests must run on a B200")
    dev = torch.device("cuda:0")
    torch.cuda.set_device(0)
    rng = np.random.default_rng(14)
    cases = []
    for m, hi in ((1, 5), (37, 3), (5000, 40), (300000, 12)):
        lens = rng.integers(0, hi, m)
        if m >= 5000:
            lens[rng.integers(0, m, 5)] = 20000                  # skewed rows: uneven nnz-balanced ranges
        cases.append(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32))
    cases.append(np.zeros(101, np.int32))                        # all rows empty: total = 0
    stream = torch.cuda.current_stream().cuda_stream
    for rp in cases:
        m, nnz = len(rp) - 1, int(rp[-1])
        d_rp = torch.from_numpy(rp).to(dev)
        torch.cuda.synchronize()
        # two (and more) calls in a row on the shared scratch, largest W first
        for W in (capi.NOALLOC_MAX_WORKERS, 1000, 8, 3, 2, 1, 8):
            want = capi.partition(d_rp, m, nnz, W, device=True)                    # existing entry point (allocates its scratch)
            got = capi.partition_noalloc(d_rp, m, nnz, W, stream)
            assert got.dtype == want.dtype and np.array_equal(got, want), (m, W)
            again = capi.partition_noalloc(d_rp, m, nnz, W, None)                 # legacy default stream, same bits
            assert np.array_equal(again, want), (m, W)
            ref = np.asarray(oracle.partition(W, rp, m, nnz), np.int64).reshape(W, 2)
            assert np.array_equal(got, ref), (m, W)                                # and the oracle (pinned on the reference macro)
        del d_rp
    # beyond the static scratch: refused with a clear error, nothing launched
    d_rp = torch.from_numpy(cases[2]).to(dev)
    with pytest.raises(capi.B200Error, match="NOALLOC_MAX_WORKERS"):
        capi.partition_noalloc(d_rp, len(cases[2]) - 1, int(cases[2][-1]), capi.NOALLOC_MAX_WORKERS + 1)
    with pytest.raises(capi.B200Error):
        capi.partition_noalloc(d_rp, 0, 0, 2)                                      # "Empty Sums array." like the reference
    with pytest.raises(capi.B200Error):
        capi.partition_noalloc(d_rp, 10, 10, 0)

"""world_size-2 (and 3) gloo runs of the multi-GPU host logic on CPU: shard bounds from the reference
partitioner, per-rank shard generation, unequal-count all-gather, folded normalisation.  The local
product is supplied by the CPU oracle here (test infrastructure); on a GPU box it is libb200spmv."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LINE = "6000 6000 12 4 normal random 0.3 0 0.5 0.5 14"
ITERS = 6


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle as O
    from spmv_research_b200 import amg, power
    p = amg.MatrixParams.parse(LINE)
    rp, bw = amg.row_ptr(p)
    m, nnz = p.rows, int(rp[-1])
    bounds = power.shard_bounds(rp, m, nnz, world, lambda s, N, t, W: O.partition(W, s, N, t))
    i_s, i_e = bounds[rank]
    col, val = amg.fill(p, rp, bw, i_s, i_e)              # this rank generates only its rows
    rp_loc = (rp[i_s:i_e + 1] - rp[i_s]).astype(np.int32)

    def local_spmv(cur, slot, scale, norm2, parity):
        y = O.csr_spmv(rp_loc, col, val, cur.numpy()) * scale.item()
        slot.copy_(torch.from_numpy(y))
        norm2.fill_(float(np.sum(y * y)))

    # band-limited exchange: the union of the shards' column bands, as local rows of this rank
    band, loc = power.exchange_rows((int(col.min()), int(col.max()) + 1) if col.size else (0, 0), bounds, rank)
    it = power.ShardedPowerIteration(p.cols, bounds, rank, world, local_spmv, torch.device("cpu"))
    it.set_start(np.ones(p.cols))
    for _ in range(ITERS):
        it.step()
    q.put((rank, it.eigenvalue(), it.vector(), bounds.tolist(), band, loc))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_power_iteration_matches_serial(world, oracle, native):
    from spmv_research_b200 import amg
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + world
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for pr in procs:
        pr.start()
    res = [q.get(timeout=180) for _ in range(world)]
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    # serial run of the same iteration with the oracle
    p = amg.MatrixParams.parse(LINE)
    rp, col, val = amg.generate(p)
    x = np.ones(p.cols) / np.sqrt(p.cols)
    lam = 0.0
    for _ in range(ITERS):
        y = oracle.csr_spmv(rp, col, val, x)
        lam = float(np.linalg.norm(y))
        x = y / lam
    want_bounds = oracle.partition(world, rp, p.rows, int(rp[-1])).tolist()
    gband = (int(col.min()), int(col.max()) + 1)
    for rank, eig, vec, bounds, band, loc in res:
        assert bounds == want_bounds
        i_s, i_e = want_bounds[rank]
        assert tuple(band) == gband
        assert tuple(loc) == (min(max(gband[0] - i_s, 0), i_e - i_s), min(max(gband[1] - i_s, 0), i_e - i_s))
        assert abs(eig - lam) <= 1e-9 * lam                 # Rayleigh-type estimate within 1e-9 relative
        assert np.allclose(vec, x, rtol=0, atol=1e-12)

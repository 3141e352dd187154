"""Multi-GPU iterated SpMV (power iteration) over a row-sharded matrix -- one process per GPU.

Net-new layer (the reference has no multi-GPU code); what it reuses from the reference
is the row partition: rank p of P owns the rows [i_s, i_e) that loop_partitioner_balance_prefix_sums
(lib/parallel_util.h:156-184) gives worker p of P over row_ptr -- evaluated on the GPU, bit-exact.

Per iteration and rank:  u' = s * (A_p u)  written straight into its slot [i_s, i_e) of the next-x
buffer (the scale s = 1/||u|| is folded into the kernel: A(s u) = s A u), the local sum of squares
comes out of the same kernel, then the slots are exchanged (all-gather with unequal counts =
one broadcast per owner, in place) and one scalar all-reduce gives the next norm.
`exchange="p2p"` instead lets the SpMV epilogue store every finished row into all peers' next-x
buffers over NVLink (CUDA IPC mappings), so the only collective left is the scalar all-reduce,
which also orders the iterations.
torch.distributed is plumbing here; the arithmetic is in libb200spmv.so.
"""
import numpy as np
import torch
import torch.distributed as dist


def shard_bounds(row_ptr, m, nnz, world, partition_fn):
    """[world, 2] row ranges; partition_fn is capi.partition on a GPU box (tests may pass the oracle)."""
    b = np.asarray(partition_fn(row_ptr, m, nnz, world), np.int64).reshape(world, 2)
    assert b[0, 0] == 0 and b[-1, 1] == m and np.all(b[1:, 0] == b[:-1, 1]), "partition does not tile [0, m)"
    return b


def allgather_rows(vec, bounds, group=None):
    """In-place all-gather with unequal counts: slot p of `vec` is broadcast from rank p."""
    works = []
    for p, (s, e) in enumerate(bounds):
        if e > s:
            works.append(dist.broadcast(vec[s:e], src=p, group=group, async_op=True))
    for w in works:
        w.wait()


class ShardedPowerIteration:
    """local_spmv(x_full, y_slot, scale, norm2, parity) computes y_slot = scale * (A_p x_full) and
    norm2 = sum(y_slot^2); parity (0/1) says which of the two x buffers is being written (p2p stores)."""

    def __init__(self, n, bounds, rank, world, local_spmv, device, exchange="nccl", dtype=torch.float64):
        self.n, self.bounds, self.rank, self.world = n, bounds, rank, world
        self.i_s, self.i_e = int(bounds[rank, 0]), int(bounds[rank, 1])
        self.local_spmv = local_spmv
        self.device = device
        self.exchange = exchange
        self.cur = torch.zeros(n, dtype=dtype, device=device)
        self.nxt = torch.zeros(n, dtype=dtype, device=device)
        self.scale = torch.ones(1, dtype=dtype, device=device)
        self.norm2 = torch.zeros(1, dtype=dtype, device=device)
        self.eig = None
        self.parity = 0

    def set_start(self, x0):
        """x0: full-length start vector (same on every rank)."""
        self.cur.copy_(torch.as_tensor(x0, dtype=self.cur.dtype))
        self.scale.fill_(1.0 / float(np.linalg.norm(np.asarray(x0))))

    def step(self):
        slot = self.nxt[self.i_s:self.i_e]
        self.local_spmv(self.cur, slot, self.scale, self.norm2, self.parity)
        if self.exchange == "nccl" and self.world > 1:
            allgather_rows(self.nxt, self.bounds)
        if self.world > 1:
            dist.all_reduce(self.norm2)          # also orders iteration k+1 after every rank's stores of iteration k
        # ||u'|| = ||A x|| with ||x|| = 1: the dominant-eigenvalue estimate; next scale = 1 / ||u'||
        self.eig = torch.sqrt(self.norm2)
        torch.reciprocal(self.eig, out=self.scale)
        self.cur, self.nxt = self.nxt, self.cur
        self.parity ^= 1

    def eigenvalue(self):
        return float(self.eig.item())

    def vector(self):
        """The current normalised iterate (host copy)."""
        return (self.cur * self.scale).cpu().numpy()

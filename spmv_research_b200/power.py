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
which also orders the iterations.  `exchange="mc"` stores each finished row ONCE to an NVSwitch
multicast address (NVLS) that the switch replicates into every GPU's next-x buffer, so a rank's
NVLink egress is 8 bytes per row instead of 8 * (P - 1).
torch.distributed is plumbing here; the arithmetic is in libb200spmv.so.

Since round 2 the loop itself lives under the C-ABI (csrc/power.cu: b200spmv_power_*, NCCL bound inside the library, CUDA
graph captured in C, scalar arithmetic in a one-thread kernel); `CPowerIteration` below is the thin caller and the default
of bench.py.  `ShardedPowerIteration` (the round-1 Python loop over torch.distributed) is kept as a second, independent
implementation of the same exchange for cross-checks and for the multicast experiment.
"""
import numpy as np
import torch
import torch.distributed as dist


def shard_bounds(row_ptr, m, nnz, world, partition_fn):
    """[world, 2] row ranges; partition_fn is capi.partition on a GPU box (tests may pass the oracle)."""
    b = np.asarray(partition_fn(row_ptr, m, nnz, world), np.int64).reshape(world, 2)
    assert b[0, 0] == 0 and b[-1, 1] == m and np.all(b[1:, 0] == b[:-1, 1]), "partition does not tile [0, m)"
    return b


def exchange_rows(col_range, bounds, rank, group=None):
    """Rows of this rank's shard that some rank reads in the next product: the union over ranks of the shards' referenced
    column bands [lo, hi) (b200spmv_col_range), intersected with the rows [i_s, i_e) this rank owns, as LOCAL row indices.
    Returns ((band_lo, band_hi), (local_lo, local_hi)); an empty shard band (lo == hi) does not widen the union."""
    lo, hi = int(col_range[0]), int(col_range[1])
    big = 1 << 62
    t = torch.tensor([-(lo if hi > lo else big), (hi if hi > lo else -big)], dtype=torch.int64)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
        t = t.to(dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        t = t.cpu()
    glo, ghi = int(-t[0].item()), int(t[1].item())
    if ghi <= glo:
        glo = ghi = 0
    i_s, i_e = int(bounds[rank][0]), int(bounds[rank][1])
    rows = i_e - i_s
    return (glo, ghi), (min(max(glo - i_s, 0), rows), min(max(ghi - i_s, 0), rows))


def allgather_rows(vec, bounds, group=None):
    """In-place all-gather with unequal counts: slot p of `vec` is broadcast from rank p."""
    works = []
    for p, (s, e) in enumerate(bounds):
        if e > s:
            works.append(dist.broadcast(vec[s:e], src=p, group=group, async_op=True))
    for w in works:
        w.wait()


class _RawCuda:
    """__cuda_array_interface__ view of n doubles at a raw device address."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 3, "strides": None}


class PeerVectors:
    """Two full-length x buffers per rank in plain cudaMalloc memory, each mapped into every peer of the box with
    CUDA IPC (b200spmv_ipc_*): peer[b][q] is the address, valid in THIS process, of rank q's buffer b."""

    def __init__(self, n, rank, world, device):
        from . import capi
        self.n = n
        self.ptr = [capi.malloc(8 * n) for _ in range(2)]
        self._views = [_RawCuda(p, n) for p in self.ptr]
        self.tensors = [torch.as_tensor(v, device=device) for v in self._views]
        for t in self.tensors:
            t.zero_()
        torch.cuda.synchronize()
        mine = [capi.ipc_export(p) for p in self.ptr]
        every = [None] * world
        dist.all_gather_object(every, mine)
        self.peer = [[self.ptr[b] if q == rank else capi.ipc_import(every[q][b]) for q in range(world)] for b in range(2)]


def multicast_supported(device_index):
    """CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED of one device (driver query through cuda-python); False on any failure."""
    try:
        from cuda.bindings import driver
        err, = driver.cuInit(0)
        err, dev = driver.cuDeviceGet(int(device_index))
        if int(err) != 0:
            return False
        err, v = driver.cuDeviceGetAttribute(driver.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev)
        return int(err) == 0 and int(v) != 0
    except Exception:
        return False


class MulticastVectors:
    """Two full-length x buffers per rank in symmetric memory bound to an NVSwitch multicast object.
    torch.distributed._symmetric_memory is the plumbing (cuMemCreate / cuMulticastCreate / handle exchange);
    the stores are issued by the SpMV epilogue of libb200spmv.  mc[b] is the multicast address of buffer b
    (a store to mc[b] + off lands at offset off of EVERY rank's buffer b, this rank's included);
    peer[b][q] is the unicast address of rank q's buffer b.  Raises when the box has no multicast support."""

    def __init__(self, n, rank, world, device):
        import os
        os.environ.setdefault("TORCH_SYMMMEM_IMPLICIT_POOL", "0")      # one allocation per tensor: offset 0 in its block
        import torch.distributed._symmetric_memory as symm_mem
        try:
            symm_mem.set_backend("CUDA")          # cuMem* + multicast objects (not NVSHMEM / NCCL windows)
        except Exception:
            pass
        self.n = n
        self.tensors, self.hdl, self.mc, self.peer = [], [], [], []
        for _ in range(2):
            t = symm_mem.empty(n, dtype=torch.float64, device=device)
            h = symm_mem.rendezvous(t, dist.group.WORLD)
            off = int(getattr(h, "offset", 0) or 0)
            mc = int(h.multicast_ptr or 0)
            if mc == 0:
                raise RuntimeError("symmetric memory has no multicast mapping on this box")
            ptrs = [int(q) + off for q in h.buffer_ptrs]
            if ptrs[rank] != t.data_ptr():
                raise RuntimeError("symmetric memory: local buffer address does not match the tensor")
            t.zero_()
            self.tensors.append(t)
            self.hdl.append(h)
            self.mc.append(mc + off)
            self.peer.append(ptrs)
        torch.cuda.synchronize()
        dist.barrier()


class ShardedPowerIteration:
    """local_spmv(x_full, y_slot, scale, norm2, parity) computes y_slot = scale * (A_p x_full) and
    norm2 = sum(y_slot^2); parity (0/1) says which of the two x buffers is being written (p2p stores)."""

    def __init__(self, n, bounds, rank, world, local_spmv, device, exchange="nccl", dtype=torch.float64, buffers=None):
        self.n, self.bounds, self.rank, self.world = n, bounds, rank, world
        self.i_s, self.i_e = int(bounds[rank, 0]), int(bounds[rank, 1])
        self.local_spmv = local_spmv
        self.device = device
        self.exchange = exchange
        if buffers is not None:
            self.cur, self.nxt = buffers          # parity 0 writes buffers[1], parity 1 writes buffers[0]
        else:
            self.cur = torch.zeros(n, dtype=dtype, device=device)
            self.nxt = torch.zeros(n, dtype=dtype, device=device)
        self.scale = torch.ones(1, dtype=dtype, device=device)
        self.norm2 = torch.zeros(1, dtype=dtype, device=device)
        self.eig = torch.zeros(1, dtype=dtype, device=device)
        self.parity = 0
        self._graph = None

    def set_start(self, x0):
        """x0: full-length start vector (same on every rank)."""
        if self.parity:                       # restart from buffer 0 so that peer-store targets stay in step on every rank
            self.cur, self.nxt = self.nxt, self.cur
            self.parity = 0
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
        torch.sqrt(self.norm2, out=self.eig)
        torch.reciprocal(self.eig, out=self.scale)
        self.cur, self.nxt = self.nxt, self.cur
        self.parity ^= 1

    def capture_pair(self):
        """Captures two consecutive iterations (one of each buffer parity) into a CUDA graph; replay() then runs
        2 iterations per launch with no host work in between.  local_spmv must launch on torch's CURRENT stream."""
        self.step()                       # warm both parities outside capture (lazy init inside kernels / NCCL)
        self.step()
        if self.parity:                   # capture from an even iteration so that python-side state survives replays
            self.step()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self.step()
            self.step()
        self._graph = g
        return g

    def replay(self, pairs):
        for _ in range(pairs):
            self._graph.replay()

    def eigenvalue(self):
        return float(self.eig.item())

    def vector(self):
        """The current normalised iterate (host copy)."""
        return (self.cur * self.scale).cpu().numpy()


class CPowerIteration:
    """Thin caller of b200spmv_power_* (include/b200spmv.h, csrc/power.cu).  The only thing Python contributes is carrying
    the 128-byte NCCL id from rank 0 to the other ranks (torch.distributed when initialised, else `uid`)."""

    EXCHANGE = {"nccl": 0, "fused": 1, "p2p": 1}

    def __init__(self, MF, bounds, n, rank, world, device_index, exchange="fused", chunks=0, use_graph=True, scalar_sync=0,
                 full_exchange=False, uid=None):
        from . import capi
        self.rank, self.world = rank, world
        self.comm = None
        if world > 1:
            if uid is None:
                box = [capi.comm_unique_id() if rank == 0 else None]
                dist.broadcast_object_list(box, src=0)
                uid = box[0]
            self.comm = capi.Comm(uid, rank, world, device_index)
        self.pw = capi.Power(MF.handle, self.comm, bounds, n, exchange=self.EXCHANGE[exchange], chunks=chunks,
                             full_exchange=full_exchange, use_graph=use_graph, scalar_sync=scalar_sync)
        self._eig = float("nan")

    def set_start(self, x0):
        self.pw.set_start(x0)

    def iterate(self, iters):
        """Runs `iters` iterations; returns their device time in ms."""
        self._eig, ms = self.pw.iterate(iters)
        return ms

    def step(self):
        self.iterate(1)

    def eigenvalue(self):
        return self._eig

    def get_x(self):
        return self.pw.get_x()

    def info(self):
        return self.pw.info()

    def close(self):
        self.pw.close()
        if self.comm is not None:
            self.comm.close()

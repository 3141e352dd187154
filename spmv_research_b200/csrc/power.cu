// power.cu -- multi-GPU iterated SpMV (power iteration) over a row-sharded matrix, under the C-ABI (include/b200spmv.h).
//
// Net-new layer (the reference has no multi-GPU code).  What comes from the reference is the row partition
// (loop_partitioner_balance_prefix_sums, lib/parallel_util.h:156-184 -> b200spmv_partition*) and the product itself
// (subkernel_csr_scalar, K/csr.cpp:334-350 -> csr_tile_kernel).  One process per GPU; per iteration and rank
//     u' = s * (A_p u)  written into the rank's slot of the next-x buffer,  s' = 1 / ||u'||
// with the scale folded into the SpMV epilogue (A (s u) = s A u) and the sum of squares coming out of the same kernel.
// Two exchanges of u':
//   NCCL  : grouped ncclSend / ncclRecv between every pair (an all-gather with unequal counts, in place), band-limited, issued per
//           row chunk on a second stream so that chunk c travels while chunk c + 1 is computed; ncclAllReduce of the norm;
//   FUSED : the epilogue stores each finished row of the band into every peer's next-x buffer (NVLink P2P stores through CUDA
//           IPC mappings); what is left per iteration is one scalar synchronisation -- ncclAllReduce, or a mailbox barrier
//           written by a one-warp kernel (each rank posts its sum of squares + a sequence number into every peer's mailbox,
//           then waits for all P entries of its own; fixed summation order, so every rank computes the same bits).
// A column-panel shard is launched interleaved: chunk c of panel 0, chunk c of panel 1, ..., chunk c + 1 of panel 0, ...
// (b200spmv_chunk_schedule) -- rows become final progressively instead of only in the last panel's launch.
// NCCL is bound at run time with dlopen, so libb200spmv.so itself has no link-time dependency on it (the reference-harness
// plugin links the same library and never touches this file's entry points).
#include "common.cuh"
#include <dlfcn.h>
#include <math.h>
#include <stdlib.h>
#include <nccl.h>          // types and prototypes only; the functions are resolved below

constexpr int PW_MAXW = 8;          // ranks (YOut carries up to 8 peer destinations)
constexpr int PW_MAXK = 16;         // row chunks
constexpr int PW_MAXP = 8;          // column panels

// ---------------------------------------------------------------------------------------------
// NCCL, bound at run time
// ---------------------------------------------------------------------------------------------
struct NcclApi {
	void * lib;
	decltype(&ncclGetUniqueId) GetUniqueId;
	decltype(&ncclCommInitRank) CommInitRank;
	decltype(&ncclCommDestroy) CommDestroy;
	decltype(&ncclAllReduce) AllReduce;
	decltype(&ncclBroadcast) Broadcast;
	decltype(&ncclAllGather) AllGather;
	decltype(&ncclSend) Send;
	decltype(&ncclRecv) Recv;
	decltype(&ncclGroupStart) GroupStart;
	decltype(&ncclGroupEnd) GroupEnd;
	decltype(&ncclGetErrorString) GetErrorString;
	decltype(&ncclGetVersion) GetVersion;
};
static NcclApi g_nccl = {};

static int nccl_load()
{
	if (g_nccl.lib)
		return 0;
	const char * cand[] = { getenv("B200SPMV_NCCL_LIB"), "libnccl.so.2", "libnccl.so" };
	void * lib = nullptr;
	for (const char * c : cand)
		if (c && *c && (lib = dlopen(c, RTLD_NOW | RTLD_GLOBAL)))
			break;
	if (!lib)
		B200_FAIL("NCCL not found (dlopen libnccl.so.2: %s); set B200SPMV_NCCL_LIB", dlerror());
	NcclApi a = {};
	a.lib = lib;
#define NCCL_SYM(field, name)                                                         \
	a.field = reinterpret_cast<decltype(a.field)>(dlsym(lib, name));              \
	if (!a.field)                                                                 \
		B200_FAIL("NCCL symbol %s missing", name);
	NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
	NCCL_SYM(CommInitRank, "ncclCommInitRank")
	NCCL_SYM(CommDestroy, "ncclCommDestroy")
	NCCL_SYM(AllReduce, "ncclAllReduce")
	NCCL_SYM(Broadcast, "ncclBroadcast")
	NCCL_SYM(AllGather, "ncclAllGather")
	NCCL_SYM(Send, "ncclSend")
	NCCL_SYM(Recv, "ncclRecv")
	NCCL_SYM(GroupStart, "ncclGroupStart")
	NCCL_SYM(GroupEnd, "ncclGroupEnd")
	NCCL_SYM(GetErrorString, "ncclGetErrorString")
	NCCL_SYM(GetVersion, "ncclGetVersion")
#undef NCCL_SYM
	g_nccl = a;
	return 0;
}

#define NC(call)                                                                                                  \
	do {                                                                                                      \
		ncclResult_t r_ = (call);                                                                         \
		if (r_ != ncclSuccess)                                                                            \
			B200_FAIL("%s:%d %s -> NCCL: %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(r_));  \
	} while (0)

struct b200spmv_comm {
	ncclComm_t comm;
	int rank, world, device;
};

extern "C" int b200spmv_comm_unique_id(unsigned char id_out[128])
{
	static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
	if (!id_out)
		B200_FAIL("NULL argument");
	if (nccl_load())
		return 1;
	ncclUniqueId id;
	NC(g_nccl.GetUniqueId(&id));
	memcpy(id_out, &id, 128);
	return 0;
}

extern "C" int b200spmv_comm_create(b200spmv_comm_t ** out, const unsigned char id[128], int32_t rank, int32_t world, int32_t device)
{
	if (!out || !id)
		B200_FAIL("NULL argument");
	*out = nullptr;
	if (world < 1 || world > PW_MAXW || rank < 0 || rank >= world)
		B200_FAIL("rank %d / world %d out of range (1..%d ranks)", rank, world, PW_MAXW);
	if (nccl_load())
		return 1;
	if (device >= 0)
		CU(cudaSetDevice(device));
	b200spmv_comm * c = (b200spmv_comm *) calloc(1, sizeof(b200spmv_comm));
	if (!c)
		B200_FAIL("out of host memory");
	CU(cudaGetDevice(&c->device));
	c->rank = rank;
	c->world = world;
	ncclUniqueId uid;
	memcpy(&uid, id, 128);
	ncclResult_t r = g_nccl.CommInitRank(&c->comm, world, uid, rank);
	if (r != ncclSuccess) {
		free(c);
		B200_FAIL("ncclCommInitRank(rank %d of %d) -> %s", rank, world, g_nccl.GetErrorString(r));
	}
	*out = c;
	return 0;
}

extern "C" void b200spmv_comm_destroy(b200spmv_comm_t * c)
{
	if (!c)
		return;
	cudaSetDevice(c->device);
	if (c->comm && g_nccl.CommDestroy)
		g_nccl.CommDestroy(c->comm);
	free(c);
}

// all-gather of a few host bytes per rank (IPC handles, chunk bounds, column bands) through the communicator
static int comm_allgather_host(b200spmv_comm * c, const void * send, void * recv, size_t bytes, cudaStream_t st)
{
	unsigned char * d = nullptr;
	CU(cudaMalloc(&d, bytes * (size_t) (c->world + 1)));
	CU(cudaMemcpyAsync(d, send, bytes, cudaMemcpyHostToDevice, st));
	NC(g_nccl.AllGather(d, d + bytes, bytes, ncclUint8, c->comm, st));
	CU(cudaMemcpyAsync(recv, d + bytes, bytes * (size_t) c->world, cudaMemcpyDeviceToHost, st));
	CU(cudaStreamSynchronize(st));
	CU(cudaFree(d));
	return 0;
}

// ---------------------------------------------------------------------------------------------
// row-chunk x column-panel launch order (pure host function)
// ---------------------------------------------------------------------------------------------
extern "C" int b200spmv_chunk_schedule(int32_t P, int32_t K, int64_t m, const int32_t * const * tile_row, const int64_t * ntiles,
                                       int64_t * unit_out, int64_t * final_row_out)
{
	if (P < 1 || P > PW_MAXP || K < 1 || K > PW_MAXK || !tile_row || !ntiles || !unit_out || !final_row_out)
		B200_FAIL("chunk schedule: bad arguments");
	for (int q = 0; q < P; q++)
		if (!tile_row[q] || ntiles[q] < 1 || tile_row[q][0] != 0 || tile_row[q][ntiles[q]] != m)
			B200_FAIL("chunk schedule: panel %d tile rows must start at 0 and end at m", q);
	for (int c = 0; c <= K; c++) {
		// panel 0 sets the pace (equal tile counts per chunk); each later panel stops at the last tile boundary that does not
		// pass what the panel before it has finished, so its read-modify-write of y never meets an unwritten partial
		int64_t t = (c == K) ? ntiles[0] : ntiles[0] * c / K;
		int64_t bound = (c == K) ? m : tile_row[0][t];
		unit_out[c] = t;
		for (int q = 1; q < P; q++) {
			if (c == K) {
				t = ntiles[q];
			} else if (c == 0) {
				t = 0;                          // every tile is launched: chunk 0 starts at the first one
			} else {
				// largest t with tile_row[q][t] <= bound (tile_row is non-decreasing, [0] = 0)
				int64_t lo = 0, hi = ntiles[q];
				while (lo < hi) {
					const int64_t mid = (lo + hi + 1) / 2;
					if (tile_row[q][mid] <= bound)
						lo = mid;
					else
						hi = mid - 1;
				}
				t = lo;
				bound = tile_row[q][t];
			}
			unit_out[(int64_t) q * (K + 1) + c] = t;
		}
		final_row_out[c] = (c == K) ? m : bound;
	}
	return 0;
}

// ---------------------------------------------------------------------------------------------
// the scalar step
// ---------------------------------------------------------------------------------------------
struct PowerBox {
	double val;
	unsigned long long seq;
};

struct PowerBoxPtrs {
	PowerBox * p[PW_MAXW];
};

// ||u'||^2 -> eigenvalue estimate ||A x|| (with ||x|| = 1) and the next scale
__global__ void power_scalar_kernel(const double * __restrict__ norm2, double * __restrict__ eig, double * __restrict__ scale)
{
	const double e = sqrt(*norm2);
	*eig = e;
	*scale = 1.0 / e;
}

__device__ __forceinline__ unsigned long long ld_acquire_sys_u64(const unsigned long long * p)
{
	unsigned long long v;
	asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
	return v;
}

__device__ __forceinline__ void st_release_sys_u64(unsigned long long * p, unsigned long long v)
{
	asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}

// Mailbox barrier + all-reduce of one double, one warp.  Entry [slot][r] of a rank's mailbox is written by rank r only.
// Launched after the product on the same stream: the product's peer stores are complete device-side when this kernel
// starts, the system-scope fence + release store order them before the sequence number every peer waits for.  Two slots
// (sequence parity) suffice: a rank can post iteration k + 2 only after every peer has posted k + 1, i.e. after every peer
// has finished reading iteration k.  A rank that waits longer than `spin_limit` clocks gives up and raises *err.
__global__ void power_mailbox_kernel(const double * __restrict__ norm2_local, PowerBoxPtrs peer_box, PowerBox * my_box, int rank, int world,
                                     unsigned long long * seq_counter, double * norm2, double * eig, double * scale, int * err,
                                     long long spin_limit)
{
	__shared__ unsigned long long s_seq;
	const int q = threadIdx.x;
	if (q == 0) {
		s_seq = *seq_counter + 1;
		*seq_counter = s_seq;
	}
	__syncthreads();
	const unsigned long long seq = s_seq;
	const int slot = (int) (seq & 1);
	__threadfence_system();
	if (q < world) {
		PowerBox * dst = peer_box.p[q] + slot * world + rank;
		*reinterpret_cast<volatile double *>(&dst->val) = *norm2_local;
		__threadfence_system();
		st_release_sys_u64(&dst->seq, seq);
	}
	double v = 0.0;
	if (q < world) {
		const PowerBox * src = my_box + slot * world + q;
		const long long t0 = clock64();
		bool ok = true;
		while (ld_acquire_sys_u64(&src->seq) != seq) {
			if (clock64() - t0 > spin_limit) {
				ok = false;
				break;
			}
		}
		if (ok)
			v = *reinterpret_cast<const volatile double *>(&src->val);
		else
			atomicExch(err, 1);
	}
	double tot = 0.0;
	for (int r = 0; r < world; r++)
		tot += __shfl_sync(0xffffffffu, v, r);      // rank order: the same bits on every rank
	if (q == 0) {
		const double e = sqrt(tot);
		*norm2 = tot;
		*eig = e;
		*scale = 1.0 / e;
	}
}

// ---------------------------------------------------------------------------------------------
// the iteration object
// ---------------------------------------------------------------------------------------------
struct b200spmv_power {
	b200spmv * h;
	b200spmv_comm * comm;
	int rank, world, device;
	int64_t n, i_s, i_e;
	int64_t bounds[2 * PW_MAXW];
	int exchange, scalar_sync, use_graph, full_exchange;

	double * buf[2];                     // the two full-length x buffers of this rank
	double * peer[2][PW_MAXW];           // peer[b][q]: rank q's buffer b as mapped into this process (own: buf[b])
	double * d_scal;                     // [0] scale, [1] norm2 (global), [2] eigenvalue, [3] norm2 (local)
	PowerBox * box;                      // [2][world]
	PowerBoxPtrs peer_box;
	unsigned long long * d_seq;
	int * d_err;

	int P, K;                            // column panels of the shard (1 = plain handle), row chunks
	bool ranged;                         // launched by tile ranges (CSR without split rows / row compression)
	int64_t unit[PW_MAXP][PW_MAXK + 1];
	int64_t final_row[PW_MAXK + 1];      // local rows [0, final_row[c]) are final once chunks < c have run
	int64_t all_final[PW_MAXW][PW_MAXK + 1];     // every rank's chunk bounds as GLOBAL rows
	int64_t band_lo, band_hi;            // global rows some rank reads (the exchanged rows)
	int64_t peer_lo, peer_hi;            // ... intersected with this rank's slot, as local rows

	cudaStream_t st, cst;
	cudaEvent_t ev_chunk[PW_MAXK], ev_norm, ev_join, t0, t1;
	cudaGraphExec_t gexec;
	int parity;
	int launches_per_iter, nccl_per_iter;
};

static b200spmv * last_panel(b200spmv * h) { return h->npanels > 1 ? h->panel[h->npanels - 1] : h; }

static void fill_launch(const b200spmv_power * pw, int parity, b200spmv_launch * l)
{
	memset(l, 0, sizeof(*l));
	l->scale = pw->d_scal + 0;
	l->norm2 = pw->d_scal + ((pw->world > 1) ? 3 : 1);
	if (pw->exchange == B200SPMV_X_FUSED && pw->world > 1) {
		int k = 0;
		for (int q = 0; q < pw->world; q++)
			if (q != pw->rank)
				l->y_peer[k++] = pw->peer[1 - parity][q] + pw->i_s;
		l->n_peer = k;
		l->peer_row_lo = pw->peer_lo;
		l->peer_row_hi = pw->peer_hi;
		if (pw->peer_hi <= pw->peer_lo) {
			// nothing of this slot is read by anyone: (0, 0) would mean "every row", so send an empty range explicitly
			l->peer_row_lo = pw->i_e - pw->i_s;
			l->peer_row_hi = pw->i_e - pw->i_s + 1;
		}
	}
}

// chunk c of every panel, ascending panels; the last panel carries the epilogue
static int launch_chunk(b200spmv_power * pw, int c, const double * cur, double * slot, const b200spmv_launch * l)
{
	b200spmv * h = pw->h;
	for (int q = 0; q < pw->P; q++) {
		b200spmv * s = pw->P > 1 ? h->panel[q] : h;
		const int64_t first = pw->unit[q][c], count = pw->unit[q][c + 1] - first;
		if (count <= 0)
			continue;
		if (launch_csr_tiles_range(s, cur, slot, pw->st, (q + 1 == pw->P) ? l : nullptr, first, count))
			return 1;
	}
	return 0;
}

static int finish_norm(b200spmv_power * pw, const b200spmv_launch * l)
{
	b200spmv * s = last_panel(pw->h);
	return launch_norm_reduce(s, s->ntiles + s->nsplit, l->norm2, pw->st);
}

static int one_iteration(b200spmv_power * pw)
{
	const int p = pw->parity;
	const double * cur = pw->buf[p];
	double * nxt = pw->buf[1 - p];
	double * slot = nxt + pw->i_s;
	b200spmv_launch l;
	fill_launch(pw, p, &l);
	const bool nccl_x = (pw->exchange == B200SPMV_X_NCCL && pw->world > 1);
	if (!nccl_x) {
		if (pw->ranged) {
			for (int c = 0; c < pw->K; c++)
				if (launch_chunk(pw, c, cur, slot, &l))
					return 1;
			if (finish_norm(pw, &l))
				return 1;
		} else if (launch_product(pw->h, cur, slot, pw->st, &l)) {
			return 1;
		}
		if (pw->world > 1) {
			if (pw->scalar_sync == 1) {
				power_mailbox_kernel<<<1, 32, 0, pw->st>>>(pw->d_scal + 3, pw->peer_box, pw->box, pw->rank, pw->world, pw->d_seq, pw->d_scal + 1,
				                                          pw->d_scal + 2, pw->d_scal + 0, pw->d_err, 20000000000LL);
				CU_LAUNCH();
			} else {
				NC(g_nccl.AllReduce(pw->d_scal + 3, pw->d_scal + 1, 1, ncclDouble, ncclSum, pw->comm->comm, pw->st));
				power_scalar_kernel<<<1, 1, 0, pw->st>>>(pw->d_scal + 1, pw->d_scal + 2, pw->d_scal + 0);
				CU_LAUNCH();
			}
		} else {
			power_scalar_kernel<<<1, 1, 0, pw->st>>>(pw->d_scal + 1, pw->d_scal + 2, pw->d_scal + 0);
			CU_LAUNCH();
		}
	} else {
		// NCCL exchange: chunk c is gathered on the second stream while chunk c + 1 is computed
		for (int c = 0; c < pw->K; c++) {
			if (pw->ranged) {
				if (launch_chunk(pw, c, cur, slot, &l))
					return 1;
			} else if (c == pw->K - 1) {
				b200spmv_launch l2 = l;
				if (launch_product(pw->h, cur, slot, pw->st, &l2))      // whole shard (SELL / split rows / compressed rows), incl. norm
					return 1;
			}
			CU(cudaEventRecord(pw->ev_chunk[c], pw->st));
			CU(cudaStreamWaitEvent(pw->cst, pw->ev_chunk[c], 0));
			// all-gather with unequal counts, in place: every rank sends the band rows of its chunk c to every peer and receives
			// theirs, one grouped ncclSend / ncclRecv exchange (direct peer-to-peer over NVSwitch; measured at N = 8: 8 grouped ring
			// broadcasts per chunk cost 2.2 ms per iteration, profiles/r02n8_bench_n8.json)
			int64_t lo[PW_MAXW], hi[PW_MAXW];
			for (int q = 0; q < pw->world; q++) {
				lo[q] = pw->all_final[q][c];
				hi[q] = pw->all_final[q][c + 1];
				if (!pw->full_exchange) {
					lo[q] = lo[q] > pw->band_lo ? lo[q] : pw->band_lo;
					hi[q] = hi[q] < pw->band_hi ? hi[q] : pw->band_hi;
				}
			}
			NC(g_nccl.GroupStart());
			for (int q = 0; q < pw->world; q++) {
				if (q == pw->rank)
					continue;
				if (hi[pw->rank] > lo[pw->rank])
					NC(g_nccl.Send(nxt + lo[pw->rank], (size_t) (hi[pw->rank] - lo[pw->rank]), ncclDouble, q, pw->comm->comm, pw->cst));
				if (hi[q] > lo[q])
					NC(g_nccl.Recv(nxt + lo[q], (size_t) (hi[q] - lo[q]), ncclDouble, q, pw->comm->comm, pw->cst));
			}
			NC(g_nccl.GroupEnd());
		}
		if (pw->ranged && finish_norm(pw, &l))
			return 1;
		CU(cudaEventRecord(pw->ev_norm, pw->st));
		CU(cudaStreamWaitEvent(pw->cst, pw->ev_norm, 0));
		NC(g_nccl.AllReduce(pw->d_scal + 3, pw->d_scal + 1, 1, ncclDouble, ncclSum, pw->comm->comm, pw->cst));
		power_scalar_kernel<<<1, 1, 0, pw->cst>>>(pw->d_scal + 1, pw->d_scal + 2, pw->d_scal + 0);
		CU_LAUNCH();
		CU(cudaEventRecord(pw->ev_join, pw->cst));
		CU(cudaStreamWaitEvent(pw->st, pw->ev_join, 0));
	}
	pw->parity ^= 1;
	return 0;
}

static void power_free(b200spmv_power * pw)
{
	if (!pw)
		return;
	cudaSetDevice(pw->device);
	if (pw->st)
		cudaStreamSynchronize(pw->st);
	if (pw->cst)
		cudaStreamSynchronize(pw->cst);
	if (pw->gexec)
		cudaGraphExecDestroy(pw->gexec);
	for (int b = 0; b < 2; b++)
		for (int q = 0; q < pw->world; q++)
			if (q != pw->rank && pw->peer[b][q])
				cudaIpcCloseMemHandle(pw->peer[b][q]);
	for (int q = 0; q < pw->world; q++)
		if (q != pw->rank && pw->peer_box.p[q])
			cudaIpcCloseMemHandle(pw->peer_box.p[q]);
	void * ptrs[] = { pw->buf[0], pw->buf[1], pw->d_scal, pw->box, pw->d_seq, pw->d_err };
	for (void * p : ptrs)
		if (p)
			cudaFree(p);
	for (cudaEvent_t e : pw->ev_chunk)
		if (e)
			cudaEventDestroy(e);
	cudaEvent_t evs[] = { pw->ev_norm, pw->ev_join, pw->t0, pw->t1 };
	for (cudaEvent_t e : evs)
		if (e)
			cudaEventDestroy(e);
	if (pw->st)
		cudaStreamDestroy(pw->st);
	if (pw->cst)
		cudaStreamDestroy(pw->cst);
	free(pw);
}

// tile rows of a plain CSR sub-handle on the host
static int fetch_tile_rows(const b200spmv * s, int32_t ** out)
{
	int64_t cnt = 0;
	if (b200spmv_export(s, B200SPMV_X_TILE_ROW, nullptr, 0, &cnt))
		return 1;
	int32_t * a = (int32_t *) malloc((size_t) (cnt > 0 ? cnt : 1) * sizeof(int32_t));
	if (!a)
		B200_FAIL("out of host memory");
	if (b200spmv_export(s, B200SPMV_X_TILE_ROW, a, cnt * (int64_t) sizeof(int32_t), &cnt)) {
		free(a);
		return 1;
	}
	*out = a;
	return 0;
}

static int power_setup(b200spmv_power * pw, const b200spmv_power_opts * o)
{
	b200spmv * h = pw->h;
	const int W = pw->world;
	CU(cudaStreamCreateWithFlags(&pw->st, cudaStreamNonBlocking));
	CU(cudaStreamCreateWithFlags(&pw->cst, cudaStreamNonBlocking));
	for (int c = 0; c < PW_MAXK; c++)
		CU(cudaEventCreateWithFlags(&pw->ev_chunk[c], cudaEventDisableTiming));
	CU(cudaEventCreateWithFlags(&pw->ev_norm, cudaEventDisableTiming));
	CU(cudaEventCreateWithFlags(&pw->ev_join, cudaEventDisableTiming));
	CU(cudaEventCreate(&pw->t0));
	CU(cudaEventCreate(&pw->t1));
	for (int b = 0; b < 2; b++) {
		CU(cudaMalloc(&pw->buf[b], (size_t) (pw->n > 0 ? pw->n : 1) * sizeof(double)));
		CU(cudaMemsetAsync(pw->buf[b], 0, (size_t) (pw->n > 0 ? pw->n : 1) * sizeof(double), pw->st));
		pw->peer[b][pw->rank] = pw->buf[b];
	}
	CU(cudaMalloc(&pw->d_scal, 4 * sizeof(double)));
	CU(cudaMemsetAsync(pw->d_scal, 0, 4 * sizeof(double), pw->st));
	// (2 MB: an allocation of its own, so that the IPC mapping of the mailbox exposes nothing else of this process)
	CU(cudaMalloc(&pw->box, 2u << 20));
	CU(cudaMemsetAsync(pw->box, 0, 2u << 20, pw->st));
	CU(cudaMalloc(&pw->d_seq, sizeof(unsigned long long)));
	CU(cudaMemsetAsync(pw->d_seq, 0, sizeof(unsigned long long), pw->st));
	CU(cudaMalloc(&pw->d_err, 2 * sizeof(int)));                 // [0] mailbox time-out flag, [1] barrier scratch
	CU(cudaMemsetAsync(pw->d_err, 0, 2 * sizeof(int), pw->st));
	pw->peer_box.p[pw->rank] = pw->box;
	CU(cudaStreamSynchronize(pw->st));

	// 1. launch order: tile ranges when every row is finished by the tile kernel
	const int P = h->npanels > 1 ? h->npanels : 1;
	bool ranged = true;
	for (int q = 0; q < P && ranged; q++) {
		const b200spmv * s = P > 1 ? h->panel[q] : h;
		ranged = (s->format == B200SPMV_CSR_ADAPTIVE && s->nsplit == 0 && !s->compressed && s->npanels <= 1);
	}
	// chunks pay when rows must become final early: the NCCL gather overlapping the next chunk's kernel, or a panel shard
	// under the fused exchange; a plain CSR shard under the fused exchange finishes its rows progressively anyway
	int K = o->chunks;
	if (K == 0)
		K = (W > 1 && (pw->exchange == B200SPMV_X_NCCL || P > 1)) ? 4 : 1;
	if (!ranged && !(pw->exchange == B200SPMV_X_NCCL && W > 1))
		K = 1;
	if (ranged) {
		const b200spmv * s0 = P > 1 ? h->panel[0] : h;
		if (s0->ntiles < 8 * (int64_t) K)
			K = 1;                          // too small to be worth cutting
	}
	// every rank issues the same K grouped collectives: agree on the largest request
	int64_t kk[PW_MAXW] = { 0 }, mine = K;
	if (W > 1) {
		if (comm_allgather_host(pw->comm, &mine, kk, sizeof(int64_t), pw->st))
			return 1;
		for (int q = 0; q < W; q++)
			K = kk[q] > K ? (int) kk[q] : K;
	}
	if (K == 1)
		ranged = false;                     // one launch of the whole shard: the epilogue instantiation, not the tile-range one
	pw->P = P;
	pw->K = K;
	pw->ranged = ranged;
	if (ranged) {
		int32_t * rows[PW_MAXP] = { nullptr };
		int64_t nt[PW_MAXP];
		int rc = 0;
		for (int q = 0; q < P && !rc; q++) {
			const b200spmv * s = P > 1 ? h->panel[q] : h;
			nt[q] = s->ntiles;
			rc = fetch_tile_rows(s, &rows[q]);
		}
		int64_t unit[PW_MAXP * (PW_MAXK + 1)];
		if (!rc)
			rc = b200spmv_chunk_schedule(P, K, h->m, rows, nt, unit, pw->final_row);
		for (int q = 0; q < P; q++) {
			free(rows[q]);
			for (int c = 0; c <= K && !rc; c++)
				pw->unit[q][c] = unit[q * (K + 1) + c];
		}
		if (rc)
			return 1;
	} else {
		// one launch of the whole shard, placed in the LAST chunk
		for (int c = 0; c <= K; c++)
			pw->final_row[c] = (c == K) ? h->m : 0;
	}

	// 2. what travels: every rank's chunk bounds (global rows) and the union of the referenced column bands
	int64_t send[PW_MAXK + 3], recv[PW_MAXW * (PW_MAXK + 3)];
	for (int c = 0; c <= PW_MAXK; c++)
		send[c] = pw->i_s + pw->final_row[c <= K ? c : K];
	send[PW_MAXK + 1] = h->col_hi > h->col_lo ? h->col_lo : INT64_MAX;
	send[PW_MAXK + 2] = h->col_hi > h->col_lo ? h->col_hi : INT64_MIN;
	if (W > 1) {
		if (comm_allgather_host(pw->comm, send, recv, sizeof(send), pw->st))
			return 1;
	} else {
		memcpy(recv, send, sizeof(send));
	}
	int64_t lo = INT64_MAX, hi = INT64_MIN;
	for (int q = 0; q < W; q++) {
		const int64_t * r = recv + (size_t) q * (PW_MAXK + 3);
		for (int c = 0; c <= K; c++)
			pw->all_final[q][c] = r[c];
		lo = r[PW_MAXK + 1] < lo ? r[PW_MAXK + 1] : lo;
		hi = r[PW_MAXK + 2] > hi ? r[PW_MAXK + 2] : hi;
	}
	if (hi <= lo)
		lo = hi = 0;
	if (pw->full_exchange) {
		lo = 0;
		hi = pw->n;
	}
	pw->band_lo = lo;
	pw->band_hi = hi;
	const int64_t rows = pw->i_e - pw->i_s;
	int64_t a = lo - pw->i_s, b = hi - pw->i_s;
	a = a < 0 ? 0 : (a > rows ? rows : a);
	b = b < 0 ? 0 : (b > rows ? rows : b);
	pw->peer_lo = a;
	pw->peer_hi = b;

	// 3. fused exchange: map every peer's buffers (and mailbox) into this process
	if (pw->exchange == B200SPMV_X_FUSED && W > 1) {
		cudaIpcMemHandle_t mine_h[3], all_h[3 * PW_MAXW];
		CU(cudaIpcGetMemHandle(&mine_h[0], pw->buf[0]));
		CU(cudaIpcGetMemHandle(&mine_h[1], pw->buf[1]));
		CU(cudaIpcGetMemHandle(&mine_h[2], pw->box));
		if (comm_allgather_host(pw->comm, mine_h, all_h, sizeof(mine_h), pw->st))
			return 1;
		for (int q = 0; q < W; q++) {
			if (q == pw->rank)
				continue;
			void * p0 = nullptr, * p1 = nullptr, * p2 = nullptr;
			CU(cudaIpcOpenMemHandle(&p0, all_h[3 * q + 0], cudaIpcMemLazyEnablePeerAccess));
			CU(cudaIpcOpenMemHandle(&p1, all_h[3 * q + 1], cudaIpcMemLazyEnablePeerAccess));
			CU(cudaIpcOpenMemHandle(&p2, all_h[3 * q + 2], cudaIpcMemLazyEnablePeerAccess));
			pw->peer[0][q] = (double *) p0;
			pw->peer[1][q] = (double *) p1;
			pw->peer_box.p[q] = (PowerBox *) p2;
		}
	}
	// launches / NCCL calls per iteration (reporting)
	int lp = 0;
	if (ranged) {
		for (int c = 0; c < K; c++)
			for (int q = 0; q < P; q++)
				lp += (pw->unit[q][c + 1] > pw->unit[q][c]) ? 1 : 0;
		lp += 1;                                  // norm reduce
	} else {
		lp = b200spmv_launches_per_spmv(h) + 1;
	}
	lp += 1;                                      // scalar / mailbox kernel
	pw->launches_per_iter = lp;
	pw->nccl_per_iter = (W > 1) ? ((pw->exchange == B200SPMV_X_NCCL) ? K + 1 : (pw->scalar_sync == 1 ? 0 : 1)) : 0;
	return 0;
}

extern "C" int b200spmv_power_create(b200spmv_power_t ** out, b200spmv_t * shard, b200spmv_comm_t * comm, const int64_t * bounds, int64_t n,
                                     const b200spmv_power_opts * opts)
{
	B200Range nvtx_range_("b200spmv_power_create");
	if (!out || !shard || !bounds)
		B200_FAIL("NULL argument");
	*out = nullptr;
	b200spmv_power_opts o;
	memset(&o, 0, sizeof(o));
	if (opts)
		o = *opts;
	const int W = comm ? comm->world : 1, rank = comm ? comm->rank : 0;
	if (o.exchange != B200SPMV_X_NCCL && o.exchange != B200SPMV_X_FUSED)
		B200_FAIL("unknown exchange %d", o.exchange);
	if (o.chunks < 0 || o.chunks > PW_MAXK)
		B200_FAIL("chunks must be in [0, %d]", PW_MAXK);
	if (W > 1 && nccl_load())
		return 1;
	if (bounds[0] != 0 || bounds[2 * W - 1] != n)
		B200_FAIL("bounds must tile [0, n)");
	for (int q = 0; q < W; q++)
		if (bounds[2 * q + 1] < bounds[2 * q] || (q > 0 && bounds[2 * q] != bounds[2 * q - 1]))
			B200_FAIL("bounds must tile [0, n)");
	if (shard->m != bounds[2 * rank + 1] - bounds[2 * rank] || shard->n != n)
		B200_FAIL("shard is %lld x %lld, expected %lld x %lld", (long long) shard->m, (long long) shard->n,
		          (long long) (bounds[2 * rank + 1] - bounds[2 * rank]), (long long) n);
	if (comm && comm->device != shard->device)
		B200_FAIL("communicator on device %d, shard on device %d", comm->device, shard->device);
	CU(cudaSetDevice(shard->device));
	b200spmv_power * pw = (b200spmv_power *) calloc(1, sizeof(b200spmv_power));
	if (!pw)
		B200_FAIL("out of host memory");
	pw->h = shard;
	pw->comm = comm;
	pw->rank = rank;
	pw->world = W;
	pw->device = shard->device;
	pw->n = n;
	pw->i_s = bounds[2 * rank];
	pw->i_e = bounds[2 * rank + 1];
	memcpy(pw->bounds, bounds, 2 * (size_t) W * sizeof(int64_t));
	pw->exchange = o.exchange;
	pw->scalar_sync = o.scalar_sync ? 1 : 0;
	pw->use_graph = o.use_graph ? 1 : 0;
	pw->full_exchange = o.full_exchange ? 1 : 0;
	if (power_setup(pw, &o)) {
		power_free(pw);
		return 1;
	}
	*out = pw;
	return 0;
}

extern "C" void b200spmv_power_destroy(b200spmv_power_t * pw) { power_free(pw); }

extern "C" int b200spmv_power_set_start(b200spmv_power_t * pw, const double * x0)
{
	if (!pw || !x0)
		B200_FAIL("NULL argument");
	CU(cudaSetDevice(pw->device));
	CU(cudaStreamSynchronize(pw->st));
	CU(cudaStreamSynchronize(pw->cst));
	if (pw->parity) {
		// restart from buffer 0 on every rank; an odd number of iterations so far means one more keeps the peers' store
		// targets in step -- simpler: flip back (all ranks have run the same count)
		pw->parity = 0;
	}
	double s2 = 0.0;
	for (int64_t i = 0; i < pw->n; i++)
		s2 += x0[i] * x0[i];
	const double sc[4] = { 1.0 / sqrt(s2), s2, sqrt(s2), 0.0 };
	CU(cudaMemcpyAsync(pw->buf[0], x0, (size_t) pw->n * sizeof(double), cudaMemcpyHostToDevice, pw->st));
	CU(cudaMemcpyAsync(pw->d_scal, sc, sizeof(sc), cudaMemcpyHostToDevice, pw->st));
	CU(cudaStreamSynchronize(pw->st));
	return 0;
}

static int capture_pair(b200spmv_power * pw)
{
	cudaGraph_t g = nullptr;
	CU(cudaStreamBeginCapture(pw->st, cudaStreamCaptureModeThreadLocal));
	int rc = one_iteration(pw) || one_iteration(pw);
	cudaError_t e = cudaStreamEndCapture(pw->st, &g);
	if (rc || e != cudaSuccess || !g) {
		if (g)
			cudaGraphDestroy(g);
		cudaGetLastError();
		if (!rc)
			B200_FAIL("CUDA graph capture of the iteration failed: %s", cudaGetErrorString(e));
		return 1;
	}
	e = cudaGraphInstantiate(&pw->gexec, g, 0);
	cudaGraphDestroy(g);
	if (e != cudaSuccess) {
		pw->gexec = nullptr;
		B200_FAIL("cudaGraphInstantiate: %s", cudaGetErrorString(e));
	}
	return 0;
}

extern "C" int b200spmv_power_iterate(b200spmv_power_t * pw, int32_t iters, double * eigenvalue_out, float * ms_out)
{
	B200Range nvtx_range_("b200spmv_power_iterate");
	if (!pw || iters < 0)
		B200_FAIL("bad arguments");
	CU(cudaSetDevice(pw->device));
	int left = iters;
	// line the ranks up on the device before the first iteration: a rank still busy elsewhere (building its shard) must not
	// run its peers into the mailbox time-out, and the timed region below should not include waiting for a late starter
	if (pw->world > 1)
		NC(g_nccl.AllReduce(pw->d_err + 1, pw->d_err + 1, 1, ncclInt, ncclSum, pw->comm->comm, pw->st));
	if (pw->use_graph && !pw->gexec && left >= 4) {
		// warm both parities eagerly (lazy initialisation inside NCCL), then capture from an even iteration
		if (pw->parity) {
			if (one_iteration(pw))
				return 1;
			left--;
		}
		if (one_iteration(pw) || one_iteration(pw))
			return 1;
		left -= 2;
		CU(cudaStreamSynchronize(pw->st));
		CU(cudaStreamSynchronize(pw->cst));
		if (capture_pair(pw))
			return 1;
		// capture does not execute: parity is back where it was after the two captured (not run) iterations
	}
	CU(cudaEventRecord(pw->t0, pw->st));
	if (pw->gexec) {
		if (pw->parity && left > 0) {
			if (one_iteration(pw))
				return 1;
			left--;
		}
		for (; left >= 2; left -= 2)
			CU(cudaGraphLaunch(pw->gexec, pw->st));
	}
	for (; left > 0; left--)
		if (one_iteration(pw))
			return 1;
	CU(cudaEventRecord(pw->t1, pw->st));
	CU(cudaStreamSynchronize(pw->st));
	CU(cudaStreamSynchronize(pw->cst));
	int err = 0;
	CU(cudaMemcpy(&err, pw->d_err, sizeof(int), cudaMemcpyDeviceToHost));
	if (err)
		B200_FAIL("power iteration: a peer did not reach the mailbox barrier in time");
	if (ms_out)
		CU(cudaEventElapsedTime(ms_out, pw->t0, pw->t1));
	if (eigenvalue_out)
		CU(cudaMemcpy(eigenvalue_out, pw->d_scal + 2, sizeof(double), cudaMemcpyDeviceToHost));
	return 0;
}

extern "C" int b200spmv_power_get_x(b200spmv_power_t * pw, double * x_host, double * scale_out)
{
	if (!pw)
		B200_FAIL("NULL argument");
	CU(cudaSetDevice(pw->device));
	CU(cudaStreamSynchronize(pw->st));
	CU(cudaStreamSynchronize(pw->cst));
	if (x_host)
		CU(cudaMemcpy(x_host, pw->buf[pw->parity], (size_t) pw->n * sizeof(double), cudaMemcpyDeviceToHost));
	if (scale_out)
		CU(cudaMemcpy(scale_out, pw->d_scal, sizeof(double), cudaMemcpyDeviceToHost));
	return 0;
}

extern "C" int b200spmv_power_info(const b200spmv_power_t * pw, int64_t * out, int32_t n_out)
{
	if (!pw || !out)
		B200_FAIL("NULL argument");
	const int64_t v[7] = { pw->K, pw->P, pw->band_lo, pw->band_hi, pw->gexec ? 1 : 0, pw->launches_per_iter, pw->nccl_per_iter };
	for (int i = 0; i < n_out && i < 7; i++)
		out[i] = v[i];
	return 0;
}

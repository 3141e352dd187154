// partition.cu -- the reference's integer work partitioners, evaluated on the GPU, bit-exact.
//
//  * nnz-balanced row ranges: loop_partitioner_balance_prefix_sums (lib/parallel_util.h:156-184)
//    with macros_binary_search's closest-index rule (lib/macros/macrolib.h:537-581): exact hit ->
//    leftmost index holding the value, otherwise the nearer neighbour, ties to the upper one.
//    Used to row-shard a matrix over GPUs (W = number of ranks).
//  * merge-path diagonal search: MergePathSearch (K/merge.cpp:230-251).
#include "common.cuh"

__device__ int64_t closest_index_dev(const int32_t * A, int64_t lo, int64_t hi, int64_t target)
{
	int64_t s = lo, e = hi;
	if (target < A[s])
		return s;
	if (target > A[e])
		return e;
	for (;;) {
		const int64_t mid = (s + e) / 2;
		if (mid == s || mid == e)
			break;
		if (target > A[mid])
			s = mid;
		else
			e = mid;
	}
	if (target == A[s])
		return s;
	if (target == A[e])
		return e;
	const int64_t ds = llabs(target - (int64_t) A[s]);
	const int64_t de = llabs(target - (int64_t) A[e]);
	return (ds < de) ? s : e;
}

__global__ void partition_kernel(const int32_t * __restrict__ sums, int64_t N, int64_t total, int W, int64_t * __restrict__ bounds)
{
	const int p = blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= W)
		return;
	const int64_t target = sums[0] + (total * p) / W;
	const int64_t target_next = sums[0] + (total * (p + 1)) / W;
	bounds[2 * p] = (p == 0) ? 0 : closest_index_dev(sums, 0, N - 1, target);
	bounds[2 * p + 1] = (p == W - 1) ? N : closest_index_dev(sums, 0, N - 1, target_next);
}

int b200spmv_partition_device(const int32_t * d_sums, int64_t N, int64_t total, int32_t W, int64_t * bounds_out_host)
{
	if (N < 1)
		B200_FAIL("Empty Sums array.");          // the reference's own error text (parallel_util.h:166)
	if (W < 1)
		B200_FAIL("num_workers must be >= 1");
	int64_t * d_b;
	CU(cudaMalloc(&d_b, 2 * (size_t) W * sizeof(int64_t)));
	partition_kernel<<<(W + 127) / 128, 128>>>(d_sums, N, total, W, d_b);
	CU_LAUNCH();
	CU(cudaMemcpy(bounds_out_host, d_b, 2 * (size_t) W * sizeof(int64_t), cudaMemcpyDeviceToHost));
	CU(cudaFree(d_b));
	return 0;
}

int b200spmv_partition(const int32_t * sums_host, int64_t N, int64_t total, int32_t W, int64_t * bounds_out_host)
{
	if (N < 1)
		B200_FAIL("Empty Sums array.");
	int32_t * d_s;
	CU(cudaMalloc(&d_s, (size_t) N * sizeof(int32_t)));
	CU(cudaMemcpy(d_s, sums_host, (size_t) N * sizeof(int32_t), cudaMemcpyHostToDevice));
	const int rc = b200spmv_partition_device(d_s, N, total, W, bounds_out_host);
	cudaFree(d_s);
	return rc;
}

__global__ void merge_coords_kernel(const int32_t * __restrict__ row_end, int64_t m, int64_t nnz, int W, int64_t * __restrict__ coords)
{
	const int t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t > W)
		return;
	const int64_t items = m + nnz;
	const int64_t per = (items + W - 1) / W;
	int64_t d = per * t;
	if (d > items)
		d = items;
	int64_t x_min = d - nnz > 0 ? d - nnz : 0;
	int64_t x_max = d < m ? d : m;
	while (x_min < x_max) {
		const int64_t pivot = (x_min + x_max) >> 1;
		if (row_end[pivot] <= d - pivot - 1)
			x_min = pivot + 1;
		else
			x_max = pivot;
	}
	coords[2 * t] = x_min < m ? x_min : m;
	coords[2 * t + 1] = d - x_min;
}

int b200spmv_merge_coords(const b200spmv_t * h, int32_t W, int64_t * coords_out_host)
{
	if (!h || W < 1)
		B200_FAIL("bad arguments");
	int64_t * d_c;
	CU(cudaMalloc(&d_c, 2 * (size_t) (W + 1) * sizeof(int64_t)));
	merge_coords_kernel<<<(W + 1 + 127) / 128, 128, 0, h->stream>>>(h->d_row_ptr + 1, h->m, h->nnz, W, d_c);
	CU_LAUNCH();
	CU(cudaStreamSynchronize(h->stream));
	CU(cudaMemcpy(coords_out_host, d_c, 2 * (size_t) (W + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost));
	CU(cudaFree(d_c));
	return 0;
}

// ---------------------------------------------------------------------------------------------
// b200spmv_partition_device_noalloc -- the same partition for callers that cannot call cudaMalloc (code under CUDA graph
// capture, tight serving loops, harnesses that forbid it).  No counterpart in the reference.
// The 2 W row bounds that b200spmv_partition_device puts into a cudaMalloc'd scratch buffer go into a
// file-scope device array instead: B200SPMV_NOALLOC_MAX_WORKERS (4096) workers -> 64 KB that every GPU which loads the library
// holds until exit.  Nothing here allocates: no cudaMalloc / cudaFree / cudaMallocHost, no allocating library routine.
//
// The array is ONE per device and shared by every call: it keeps the last call's contents (nothing has to start at zero -- the
// kernel overwrites all 2 W entries the copy reads, and only those are copied), and two calls that overlap on the GPU
// (different streams, or threads racing on one stream's tail) race on it.  Calls on one stream are ordered and safe.
//
// g_partition_scratch is a HOST-SIDE STAND-IN in host code, not a device address: handing it to a kernel faults (or silently
// uses the wrong memory where the GPU can reach host memory).  The device address comes from cudaGetSymbolAddress, once per
// device, with that device current.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) int64_t g_partition_scratch[2 * B200SPMV_NOALLOC_MAX_WORKERS];

static int64_t * g_partition_scratch_addr[64] = { nullptr };      // per device ordinal; filled at first use on that device

static int partition_scratch_address(int64_t ** out)
{
	int dev = -1;
	CU(cudaGetDevice(&dev));
	if (dev < 0 || dev >= 64)
		B200_FAIL("device ordinal %d outside [0, 64)", dev);
	if (!g_partition_scratch_addr[dev]) {
		void * p = nullptr;
		const cudaError_t e = cudaGetSymbolAddress(&p, g_partition_scratch);      // the current device's copy of the array
		if (e != cudaSuccess || !p)
			B200_FAIL("cudaGetSymbolAddress(g_partition_scratch) on device %d -> %s", dev, cudaGetErrorString(e));
		g_partition_scratch_addr[dev] = (int64_t *) p;
	}
	*out = g_partition_scratch_addr[dev];
	return 0;
}

int b200spmv_partition_device_noalloc(const int32_t * d_sums, int64_t N, int64_t total, int32_t W, int64_t * bounds_out_host, void * stream)
{
	if (!d_sums || !bounds_out_host)
		B200_FAIL("NULL argument");
	if (N < 1)
		B200_FAIL("Empty Sums array.");          // the reference's own error text (parallel_util.h:166)
	if (W < 1)
		B200_FAIL("num_workers must be >= 1");
	if (W > B200SPMV_NOALLOC_MAX_WORKERS)
		B200_FAIL("b200spmv_partition_device_noalloc: %d workers exceed the static scratch (B200SPMV_NOALLOC_MAX_WORKERS = %d); "
		          "use b200spmv_partition_device", W, B200SPMV_NOALLOC_MAX_WORKERS);
	int64_t * d_b = nullptr;
	if (partition_scratch_address(&d_b))
		return 1;
	cudaStream_t st = (cudaStream_t) stream;
	partition_kernel<<<(W + 127) / 128, 128, 0, st>>>(d_sums, N, total, W, d_b);
	CU_LAUNCH();
	CU(cudaMemcpyAsync(bounds_out_host, d_b, 2 * (size_t) W * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
	CU(cudaStreamSynchronize(st));
	return 0;
}

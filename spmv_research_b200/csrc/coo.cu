// coo.cu -- COO -> CSR on the GPU: the step between the Matrix Market reader (libb200mtx.so,
// include/b200mtx.h) and csr_to_format when the input is a real matrix file.
//
// Replaces the reference's coo_to_csr() (lib/storage_formats/csr/csr_gen.c:178-213, called from import_file(),
// benchmark_code/BENCH/src/bench.cpp:208-229): a bucket sort of the entries by row (row_ptr = bucket offsets) followed by an
// ascending column sort inside every row.  Here: count per row (atomics) -> exclusive scan -> scatter of
// (column << 32 | entry index) keys into the row's bucket in arbitrary order -> one warp sorts each row's keys -> columns and
// values are emitted in key order.  The keys are unique, so the result does not depend on the scatter order: columns
// ascending, equal columns in input order -- integer outputs and value placement bit-identical to the oracle's restatement
// (oracle_coo_to_csr), which is pinned on the reference's own import path.  HBM-bound integer work, once per matrix.
#include "common.cuh"
#include "coo_sort.h"

constexpr int COO_THREADS = 256;

__global__ void __launch_bounds__(COO_THREADS) coo_count_kernel(const int32_t * __restrict__ row, const int32_t * __restrict__ col, int64_t m,
                                                                 int64_t n, int64_t nnz, unsigned long long * __restrict__ cnt,
                                                                 unsigned long long * __restrict__ bad)
{
	for (int64_t e = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += (int64_t) gridDim.x * blockDim.x) {
		const int64_t r = row[e], c = col[e];
		if (r < 0 || r >= m || c < 0 || c >= n) {
			atomicMin(bad, (unsigned long long) e);       // first offending entry
			continue;
		}
		atomicAdd(&cnt[r], 1ull);
	}
}

__global__ void __launch_bounds__(COO_THREADS) coo_rowptr_kernel(const int64_t * __restrict__ off, int64_t m, int32_t * __restrict__ row_ptr)
{
	for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i <= m; i += (int64_t) gridDim.x * blockDim.x)
		row_ptr[i] = (int32_t) off[i];
}

__global__ void __launch_bounds__(COO_THREADS) coo_scatter_kernel(const int32_t * __restrict__ row, const int32_t * __restrict__ col, int64_t nnz,
                                                                   const int64_t * __restrict__ off, unsigned int * __restrict__ cursor,
                                                                   uint64_t * __restrict__ keys)
{
	for (int64_t e = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += (int64_t) gridDim.x * blockDim.x) {
		const int32_t r = row[e];
		const int64_t pos = off[r] + atomicAdd(&cursor[r], 1u);
		keys[pos] = ((uint64_t) (uint32_t) col[e] << 32) | (uint32_t) e;
	}
}

// One warp per row (grid-stride over rows).  <= 32 keys: bitonic sort in registers (shuffles); longer rows: the
// same-direction bitonic network of coo_sort.h in global memory.  Then the row is emitted: col_out / val_out in key order,
// adjacent equal columns counted as duplicates.
__global__ void __launch_bounds__(COO_THREADS) coo_sort_rows_kernel(const int32_t * __restrict__ row_ptr, int64_t m, uint64_t * keys,
                                                                     const double * __restrict__ val, int32_t * __restrict__ col_out,
                                                                     double * __restrict__ val_out, unsigned long long * __restrict__ dups)
{
	const int lane = threadIdx.x & 31;
	const int64_t warp = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int64_t nwarps = ((int64_t) gridDim.x * blockDim.x) >> 5;
	unsigned long long my_dups = 0;
	for (int64_t i = warp; i < m; i += nwarps) {
		const int64_t a = row_ptr[i], n = row_ptr[i + 1] - a;
		if (n == 0)
			continue;
		uint64_t * k = keys + a;
		if (n <= 32) {
			uint64_t v = lane < n ? k[lane] : ~0ull;
			for (int kk = 2; kk <= 32; kk <<= 1)
				for (int j = kk >> 1; j > 0; j >>= 1) {
					const uint64_t o = __shfl_xor_sync(0xffffffffu, v, j);
					v = coo_bitonic32_step(v, o, lane, kk, j);
				}
			const uint64_t prev = __shfl_up_sync(0xffffffffu, v, 1);
			if (lane < n) {
				col_out[a + lane] = (int32_t) (v >> 32);
				val_out[a + lane] = val[(uint32_t) v];
				if (lane > 0 && (v >> 32) == (prev >> 32))
					my_dups++;
			}
			continue;
		}
		int lp = 0;
		while (((int64_t) 1 << lp) < n)
			lp++;
		const int64_t half = ((int64_t) 1 << lp) >> 1;
		for (int lk = 1; lk <= lp; lk++) {
			for (int64_t t = lane; t < half; t += 32) {
				int64_t lo, hi;
				coo_flip_pair(t, lk, &lo, &hi);
				coo_cmpx(k, lo, hi, n);
			}
			__syncwarp();
			for (int ld = lk - 2; ld >= 0; ld--) {
				for (int64_t t = lane; t < half; t += 32) {
					int64_t lo, hi;
					coo_half_pair(t, ld, &lo, &hi);
					coo_cmpx(k, lo, hi, n);
				}
				__syncwarp();
			}
		}
		for (int64_t t = lane; t < n; t += 32) {
			const uint64_t v = k[t];
			col_out[a + t] = (int32_t) (v >> 32);
			val_out[a + t] = val[(uint32_t) v];
			if (t > 0 && (v >> 32) == (k[t - 1] >> 32))
				my_dups++;
		}
	}
	if (my_dups)
		atomicAdd(dups, my_dups);
}

namespace {
struct scratch {
	void * p[8] = {};
	int n = 0;
	~scratch()
	{
		for (int i = 0; i < n; i++)
			cudaFree(p[i]);
	}
	template <class T> cudaError_t get(T ** out, size_t bytes)
	{
		void * q = nullptr;
		cudaError_t e = cudaMalloc(&q, bytes ? bytes : 1);
		if (e == cudaSuccess)
			p[n++] = q;
		*out = (T *) q;
		return e;
	}
};

int blocks_for(int64_t items)
{
	int64_t b = (items + COO_THREADS - 1) / COO_THREADS;
	const int64_t cap = 148 * 8 * 4;                  // a few waves of 8 resident CTAs on each of the 148 SMs
	return (int) (b < 1 ? 1 : (b > cap ? cap : b));
}
}        // namespace

extern "C" int b200spmv_coo_to_csr(const int32_t * d_row, const int32_t * d_col, const double * d_val, int64_t m, int64_t n, int64_t nnz,
                                   int32_t * d_row_ptr, int32_t * d_col_out, double * d_val_out, int64_t * duplicates_out, void * stream)
{
	if (m < 0 || n < 0 || nnz < 0 || nnz > 0x7FFFFFFFLL)
		B200_FAIL("coo_to_csr: bad sizes (nnz must fit int32, like the reference's row_ptr)");
	if (!d_row_ptr || (nnz > 0 && (!d_row || !d_col || !d_val || !d_col_out || !d_val_out)))
		B200_FAIL("coo_to_csr: null argument");
	{
		int ndev = 0, cur = 0, major = 0;
		cudaError_t e = cudaGetDeviceCount(&ndev);
		if (e != cudaSuccess || ndev < 1)
			B200_FAIL("no CUDA device: libb200spmv has no CPU fallback (%s)", cudaGetErrorString(e));
		CU(cudaGetDevice(&cur));
		CU(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, cur));
		if (major != 10)
			B200_FAIL("device %d has compute capability %d.x; this library carries sm_100a code only", cur, major);
	}
	cudaStream_t st = (cudaStream_t) stream;
	scratch S;
	int64_t * d_off;
	unsigned int * d_cursor;
	uint64_t * d_keys;
	unsigned long long * d_flags;                       // [0] first bad entry, [1] duplicates
	CU(S.get(&d_off, (size_t) (m + 1) * sizeof(int64_t)));
	CU(S.get(&d_cursor, (size_t) (m > 0 ? m : 1) * sizeof(unsigned int)));
	CU(S.get(&d_keys, (size_t) (nnz > 0 ? nnz : 1) * sizeof(uint64_t)));
	CU(S.get(&d_flags, 2 * sizeof(unsigned long long)));
	CU(cudaMemsetAsync(d_off, 0, (size_t) (m + 1) * sizeof(int64_t), st));
	CU(cudaMemsetAsync(d_cursor, 0, (size_t) (m > 0 ? m : 1) * sizeof(unsigned int), st));
	const unsigned long long init[2] = {~0ull, 0ull};
	CU(cudaMemcpyAsync(d_flags, init, sizeof(init), cudaMemcpyHostToDevice, st));
	if (nnz > 0) {
		coo_count_kernel<<<blocks_for(nnz), COO_THREADS, 0, st>>>(d_row, d_col, m, n, nnz, (unsigned long long *) d_off, d_flags);
		CU_LAUNCH();
	}
	unsigned long long flags[2];
	CU(cudaMemcpyAsync(flags, d_flags, sizeof(flags), cudaMemcpyDeviceToHost, st));
	CU(cudaStreamSynchronize(st));
	if (flags[0] != ~0ull)
		B200_FAIL("coo_to_csr: entry %llu has an index outside [0, m) x [0, n)", flags[0]);
	if (device_exclusive_scan_i64(d_off, m + 1, st))
		return 1;
	coo_rowptr_kernel<<<blocks_for(m + 1), COO_THREADS, 0, st>>>(d_off, m, d_row_ptr);
	CU_LAUNCH();
	if (nnz > 0) {
		coo_scatter_kernel<<<blocks_for(nnz), COO_THREADS, 0, st>>>(d_row, d_col, nnz, d_off, d_cursor, d_keys);
		CU_LAUNCH();
		coo_sort_rows_kernel<<<blocks_for(m * 32), COO_THREADS, 0, st>>>(d_row_ptr, m, d_keys, d_val, d_col_out, d_val_out, d_flags + 1);
		CU_LAUNCH();
	}
	CU(cudaMemcpyAsync(flags, d_flags, sizeof(flags), cudaMemcpyDeviceToHost, st));
	CU(cudaStreamSynchronize(st));
	if (duplicates_out)
		*duplicates_out = (int64_t) flags[1];
	return 0;
}

// import_file() + csr_to_format() in one call: HOST COO arrays (e.g. from b200mtx_read) -> upload -> CSR on the GPU ->
// format handle.  Duplicate (row, col) pairs are refused: the reference orders them with an unstable sort, so no
// bit-exact behaviour exists to reproduce.
extern "C" int b200spmv_create_coo(b200spmv_t ** out, const int32_t * row, const int32_t * col, const double * val, int64_t m, int64_t n,
                                   int64_t nnz, const b200spmv_opts * opts)
{
	if (!out)
		B200_FAIL("create_coo: null argument");
	*out = nullptr;
	if (m < 0 || n < 0 || nnz < 0 || nnz > 0x7FFFFFFFLL)
		B200_FAIL("create_coo: bad sizes");
	if (nnz > 0 && (!row || !col || !val))
		B200_FAIL("create_coo: null argument");
	{
		int ndev = 0;
		cudaError_t e = cudaGetDeviceCount(&ndev);
		if (e != cudaSuccess || ndev < 1)
			B200_FAIL("no CUDA device: libb200spmv has no CPU fallback (%s)", cudaGetErrorString(e));
		if (opts && opts->device >= ndev)
			B200_FAIL("device %d out of range (%d devices)", opts->device, ndev);
		if (opts && opts->device >= 0)
			CU(cudaSetDevice(opts->device));
	}
	scratch S;
	int32_t *d_r, *d_c, *d_rp, *d_co;
	double *d_v, *d_vo;
	const size_t k = (size_t) (nnz > 0 ? nnz : 1);
	CU(S.get(&d_r, k * sizeof(int32_t)));
	CU(S.get(&d_c, k * sizeof(int32_t)));
	CU(S.get(&d_v, k * sizeof(double)));
	CU(S.get(&d_rp, (size_t) (m + 1) * sizeof(int32_t)));
	CU(S.get(&d_co, k * sizeof(int32_t)));
	CU(S.get(&d_vo, k * sizeof(double)));
	if (nnz > 0) {
		CU(cudaMemcpy(d_r, row, (size_t) nnz * sizeof(int32_t), cudaMemcpyHostToDevice));
		CU(cudaMemcpy(d_c, col, (size_t) nnz * sizeof(int32_t), cudaMemcpyHostToDevice));
		CU(cudaMemcpy(d_v, val, (size_t) nnz * sizeof(double), cudaMemcpyHostToDevice));
	}
	int64_t dups = 0;
	if (b200spmv_coo_to_csr(d_r, d_c, d_v, m, n, nnz, d_rp, d_co, d_vo, &dups, nullptr))
		return 1;
	if (dups)
		B200_FAIL("create_coo: %lld duplicate (row, col) entries", (long long) dups);
	return create_from_device(out, d_rp, d_co, d_vo, m, n, nnz, opts);
}

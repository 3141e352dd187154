// f32.cu -- the fp32 (`_f` build) CSR product.
//
// The reference harness has a single-precision build axis (benchmark_code/BENCH/src/make.sh:215-216: ValueType = float,
// ValueTypeReference = double): the matrix still arrives as double, the format's constructor narrows the values with a C
// cast (K/csr.cpp:58-167), x and y are float, and the kernel (K/csr.cpp:334-350) accumulates in float.  This file is that
// build of the B200 plugin's CSR format: b200spmv_f32_create narrows on the GPU (cvt.rn.f32.f64 = the C cast),
// b200spmv_f32_spmv_host / _device multiply.  8 bytes per non-zero instead of 12, float gathers (32 x values per 128-byte
// line instead of 16).  Separate from the fp64 handle on purpose: nothing in the measured fp64 path changes.
//
// Kernel: nnz-tiled like the fp64 one (one CTA per window of 2 048 non-zeros, lane-consecutive coalesced streams with
// L1-no-allocate / L2-evict-first loads, products through shared memory), but with the simpler row ownership of f32_tile.h:
// a window finishes the rows that lie inside it and writes one carry and one tail partial; f32_fixup_kernel completes the
// rows that cross windows in window order.  Deterministic; y needs no memset.
#include "common.cuh"
#include "f32_tile.h"

constexpr int F32_THREADS = 256;
constexpr int F32_NPT = 8;
constexpr int F32_T = F32_THREADS * F32_NPT;

struct b200spmv_f32 {
	int device;
	int64_t m, n, nnz, ntiles;
	int32_t * d_row_ptr;
	int32_t * d_col;          // padded to ntiles * T (zero-filled)
	float * d_val;            // padded likewise
	int32_t * d_tile_row;     // ntiles
	float * d_carry;          // ntiles
	float * d_tail;           // ntiles
	float * d_x;
	float * d_y;              // staging for the host entry point
	cudaStream_t stream;
};

__device__ __forceinline__ float ld_stream_f32(const float * p, uint64_t pol)
{
	float r;
	asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(r) : "l"(p), "l"(pol));
	return r;
}

// x gathers like the fp64 kernel's: L1-allocating, L2 evict_last (x is the only data with reuse)
__device__ __forceinline__ float ld_x_f32(const float * p, uint64_t pol)
{
	float r;
	asm volatile("ld.global.nc.L1::evict_last.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(r) : "l"(p), "l"(pol));
	return r;
}

__global__ void __launch_bounds__(256) f32_narrow_kernel(const double * __restrict__ in, float * __restrict__ out, int64_t n)
{
	for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x)
		out[i] = (float) in[i];                           // cvt.rn.f32.f64: round to nearest even, the C cast
}

// tile_row[t] = last r with row_ptr[r] <= t * T  (upper_bound - 1)
__global__ void __launch_bounds__(256) f32_tile_row_kernel(const int32_t * __restrict__ row_ptr, int64_t m, int64_t ntiles, int32_t * __restrict__ tile_row)
{
	for (int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; t < ntiles; t += (int64_t) gridDim.x * blockDim.x) {
		tile_row[t] = f32_tile_row_of(row_ptr, m, t * (int64_t) F32_T);
	}
}

__global__ void __launch_bounds__(F32_THREADS, 6)
f32_tile_kernel(const int32_t * __restrict__ row_ptr, const int32_t * __restrict__ col, const float * __restrict__ val,
                const int32_t * __restrict__ tile_row, const float * __restrict__ x, float * __restrict__ y,
                float * __restrict__ carry, float * __restrict__ tail, int64_t m, int64_t nnz, int64_t ntiles)
{
	__shared__ float s_prod[F32_T];
	const int tid = threadIdx.x;
	const int64_t tile = blockIdx.x;
	const uint64_t pol_stream = make_policy_evict_first();
	const uint64_t pol_x = make_policy_evict_last();
	const int64_t base0 = tile * (int64_t) F32_T;

	int c[F32_NPT];
	float v[F32_NPT];
	#pragma unroll
	for (int k = 0; k < F32_NPT; k++) {
		const int64_t idx = base0 + k * F32_THREADS + tid;
		c[k] = ld_stream_i32(col + idx, pol_stream);
		v[k] = ld_stream_f32(val + idx, pol_stream);
	}
	f32_geom g;
	f32_geometry(row_ptr, tile_row, tile, ntiles, m, nnz, F32_T, &g);
	#pragma unroll
	for (int k = 0; k < F32_NPT; k++)
		v[k] = __fmul_rn(v[k], ld_x_f32(x + c[k], pol_x));
	#pragma unroll
	for (int k = 0; k < F32_NPT; k++)
		s_prod[k * F32_THREADS + tid] = v[k];
	__syncthreads();

	if (tile == 0)
		for (int64_t r = tid; r < g.r_first; r += F32_THREADS)      // leading empty rows
			y[r] = 0.0f;

	const int64_t nrows = g.r_last - g.r_first + 1;
	if (nrows >= F32_WARP_ROWS_BELOW) {
		// thread per row: sequential ascending-j sum (the oracle's order, products rounded before the add)
		for (int64_t r = g.r_first + tid; r <= g.r_last; r += F32_THREADS) {
			int a, b;
			const int kind = f32_row_piece(row_ptr, &g, r, &a, &b);
			float acc = 0.0f;
			for (int j = a; j < b; j++)
				acc = __fadd_rn(acc, s_prod[j]);
			if (kind == 1)
				y[r] = acc;
			else if (kind == 2)
				carry[tile] = acc;
			else
				tail[tile] = acc;
		}
	} else {
		// LPR lanes per row (f32_tile.h): lanes stride the piece, butterfly reduction inside the group
		const int lpr = f32_lanes_per_row(nrows);
		const int groups = F32_THREADS / lpr, grp = tid / lpr, lane = tid % lpr;
		for (int64_t rb = g.r_first; rb <= g.r_last; rb += groups) {
			const int64_t r = rb + grp;
			const bool active = (r <= g.r_last);
			int a = 0, b = 0, kind = 0;
			if (active)
				kind = f32_row_piece(row_ptr, &g, r, &a, &b);
			float acc = 0.0f;
			for (int j = a + lane; j < b; j += lpr)
				acc = __fadd_rn(acc, s_prod[j]);
			for (int o = lpr >> 1; o > 0; o >>= 1)
				acc = __fadd_rn(acc, __shfl_xor_sync(0xffffffffu, acc, o));
			if (active && lane == 0) {
				if (kind == 1)
					y[r] = acc;
				else if (kind == 2)
					carry[tile] = acc;
				else
					tail[tile] = acc;
			}
		}
	}
}

__global__ void __launch_bounds__(256) f32_fixup_kernel(const int32_t * __restrict__ row_ptr, const int32_t * __restrict__ tile_row,
                                                         const float * __restrict__ carry, const float * __restrict__ tail, float * __restrict__ y,
                                                         int64_t m, int64_t nnz, int64_t ntiles)
{
	for (int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; t < ntiles; t += (int64_t) gridDim.x * blockDim.x) {
		int64_t r;
		float s;
		if (f32_fixup(row_ptr, tile_row, carry, tail, t, ntiles, m, nnz, F32_T, &r, &s))
			y[r] = s;
	}
}

__global__ void __launch_bounds__(256) f32_zero_kernel(float * __restrict__ y, int64_t m)
{
	for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t) gridDim.x * blockDim.x)
		y[i] = 0.0f;
}

static int grid_for(int64_t items)
{
	int64_t b = (items + 255) / 256;
	return (int) (b < 1 ? 1 : (b > 148 * 32 ? 148 * 32 : b));
}

static void f32_free(b200spmv_f32 * h)
{
	if (!h)
		return;
	cudaFree(h->d_row_ptr);
	cudaFree(h->d_col);
	cudaFree(h->d_val);
	cudaFree(h->d_tile_row);
	cudaFree(h->d_carry);
	cudaFree(h->d_tail);
	cudaFree(h->d_x);
	cudaFree(h->d_y);
	if (h->stream)
		cudaStreamDestroy(h->stream);
	free(h);
}

extern "C" void b200spmv_f32_destroy(b200spmv_f32_t * h) { f32_free(h); }

static int f32_create_impl(b200spmv_f32 * h, const int32_t * row_ptr, const int32_t * col_ind, const double * values)
{
	const int64_t m = h->m, nnz = h->nnz;
	h->ntiles = (nnz + F32_T - 1) / F32_T;
	const size_t padded = (size_t) (h->ntiles > 0 ? h->ntiles : 1) * F32_T;
	CU(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
	CU(cudaMalloc(&h->d_row_ptr, (size_t) (m + 1) * sizeof(int32_t)));
	CU(cudaMalloc(&h->d_col, padded * sizeof(int32_t)));
	CU(cudaMalloc(&h->d_val, padded * sizeof(float)));
	CU(cudaMalloc(&h->d_tile_row, (size_t) (h->ntiles > 0 ? h->ntiles : 1) * sizeof(int32_t)));
	CU(cudaMalloc(&h->d_carry, (size_t) (h->ntiles > 0 ? h->ntiles : 1) * sizeof(float)));
	CU(cudaMalloc(&h->d_tail, (size_t) (h->ntiles > 0 ? h->ntiles : 1) * sizeof(float)));
	CU(cudaMalloc(&h->d_x, (size_t) (h->n > 0 ? h->n : 1) * sizeof(float)));
	CU(cudaMalloc(&h->d_y, (size_t) (m > 0 ? m : 1) * sizeof(float)));
	CU(cudaMemsetAsync(h->d_col, 0, padded * sizeof(int32_t), h->stream));
	CU(cudaMemsetAsync(h->d_val, 0, padded * sizeof(float), h->stream));
	CU(cudaMemcpyAsync(h->d_row_ptr, row_ptr, (size_t) (m + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
	if (nnz > 0) {
		double * d_tmp;
		CU(cudaMalloc(&d_tmp, (size_t) nnz * sizeof(double)));
		cudaError_t e = cudaMemcpyAsync(h->d_col, col_ind, (size_t) nnz * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream);
		if (e == cudaSuccess)
			e = cudaMemcpyAsync(d_tmp, values, (size_t) nnz * sizeof(double), cudaMemcpyHostToDevice, h->stream);
		if (e == cudaSuccess) {
			f32_narrow_kernel<<<grid_for(nnz), 256, 0, h->stream>>>(d_tmp, h->d_val, nnz);
			f32_tile_row_kernel<<<grid_for(h->ntiles), 256, 0, h->stream>>>(h->d_row_ptr, m, h->ntiles, h->d_tile_row);
			e = cudaPeekAtLastError();
		}
		if (e == cudaSuccess)
			e = cudaStreamSynchronize(h->stream);
		cudaFree(d_tmp);
		if (e != cudaSuccess)
			B200_FAIL("f32 create: %s", cudaGetErrorString(e));
	} else
		CU(cudaStreamSynchronize(h->stream));
	return 0;
}

extern "C" int b200spmv_f32_create(b200spmv_f32_t ** out, const int32_t * row_ptr, const int32_t * col_ind, const double * values, int64_t m,
                                   int64_t n, int64_t nnz, int32_t device)
{
	if (!out)
		B200_FAIL("out is NULL");
	*out = nullptr;
	if (!row_ptr || (nnz > 0 && (!col_ind || !values)))
		B200_FAIL("NULL CSR array");
	if (m < 0 || n < 0 || nnz < 0 || m >= 0x7fffffffLL || n > 0x7fffffffLL || nnz > 0x7fffffffLL)
		B200_FAIL("m, n and nnz must be non-negative and fit int32 indices (INT_T = int32_t)");
	int ndev = 0, cur = 0, major = 0;
	cudaError_t e = cudaGetDeviceCount(&ndev);
	if (e != cudaSuccess || ndev < 1)
		B200_FAIL("no CUDA device: libb200spmv has no CPU fallback (%s)", cudaGetErrorString(e));
	if (device >= ndev)
		B200_FAIL("device %d out of range (%d devices)", device, ndev);
	if (device >= 0)
		CU(cudaSetDevice(device));
	CU(cudaGetDevice(&cur));
	CU(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, cur));
	if (major != 10)
		B200_FAIL("device %d has compute capability %d.x; this library carries sm_100a code only", cur, major);
	b200spmv_f32 * h = (b200spmv_f32 *) calloc(1, sizeof(b200spmv_f32));
	if (!h)
		B200_FAIL("out of host memory");
	h->device = cur;
	h->m = m;
	h->n = n;
	h->nnz = nnz;
	if (f32_create_impl(h, row_ptr, col_ind, values)) {
		f32_free(h);
		return 1;
	}
	*out = h;
	return 0;
}

static int f32_launch(b200spmv_f32 * h, const float * dx, float * dy, cudaStream_t st)
{
	if (h->m == 0)
		return 0;
	if (h->ntiles == 0) {
		f32_zero_kernel<<<grid_for(h->m), 256, 0, st>>>(dy, h->m);
		CU_LAUNCH();
		return 0;
	}
	f32_tile_kernel<<<(unsigned) h->ntiles, F32_THREADS, 0, st>>>(h->d_row_ptr, h->d_col, h->d_val, h->d_tile_row, dx, dy, h->d_carry, h->d_tail,
	                                                              h->m, h->nnz, h->ntiles);
	CU_LAUNCH();
	f32_fixup_kernel<<<grid_for(h->ntiles), 256, 0, st>>>(h->d_row_ptr, h->d_tile_row, h->d_carry, h->d_tail, dy, h->m, h->nnz, h->ntiles);
	CU_LAUNCH();
	return 0;
}

extern "C" int b200spmv_f32_spmv_device(b200spmv_f32_t * h, const float * dx, float * dy, void * stream)
{
	if (!h || !dx || !dy)
		B200_FAIL("f32 spmv_device: null argument");
	return f32_launch(h, dx, dy, (cudaStream_t) stream);
}

// Matrix_Format::spmv(ValueType * x, ValueType * y) of the `_f` build: host vectors, stream synchronised on return
extern "C" int b200spmv_f32_spmv_host(b200spmv_f32_t * h, const float * x, float * y)
{
	if (!h || !x || !y)
		B200_FAIL("f32 spmv_host: null argument");
	CU(cudaSetDevice(h->device));
	CU(cudaMemcpyAsync(h->d_x, x, (size_t) h->n * sizeof(float), cudaMemcpyHostToDevice, h->stream));
	if (f32_launch(h, h->d_x, h->d_y, h->stream))
		return 1;
	CU(cudaMemcpyAsync(y, h->d_y, (size_t) h->m * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
	CU(cudaStreamSynchronize(h->stream));
	return 0;
}

// device timing of `iters` products (CUDA events on the handle's stream), optional L2 flush before each
extern "C" int b200spmv_f32_time_device(b200spmv_f32_t * h, const float * dx, float * dy, int32_t iters, int64_t flush_bytes, float * ms_out)
{
	if (!h || !dx || !dy || !ms_out || iters < 1)
		B200_FAIL("f32 time_device: bad argument");
	CU(cudaSetDevice(h->device));
	cudaEvent_t e0, e1;
	CU(cudaEventCreate(&e0));
	CU(cudaEventCreate(&e1));
	void * d_flush = nullptr;
	int rc = 0;
	if (flush_bytes > 0 && cudaMalloc(&d_flush, (size_t) flush_bytes) != cudaSuccess)
		rc = 1;
	for (int i = 0; i < iters && !rc; i++) {
		if (d_flush)
			cudaMemsetAsync(d_flush, i & 0xff, (size_t) flush_bytes, h->stream);
		cudaEventRecord(e0, h->stream);
		rc = f32_launch(h, dx, dy, h->stream);
		cudaEventRecord(e1, h->stream);
		if (cudaEventSynchronize(e1) != cudaSuccess)
			rc = 1;
		if (!rc)
			cudaEventElapsedTime(&ms_out[i], e0, e1);
	}
	cudaFree(d_flush);
	cudaEventDestroy(e0);
	cudaEventDestroy(e1);
	if (rc)
		B200_FAIL("f32 time_device failed: %s", cudaGetErrorString(cudaGetLastError()));
	return 0;
}

// capi.cu -- the extern "C" boundary of libb200spmv (see include/b200spmv.h for the contract).
#include "common.cuh"

thread_local char g_b200_err[512] = "";

extern "C" const char * b200spmv_last_error(void) { return g_b200_err; }
extern "C" const char * b200spmv_version(void) { return "b200spmv 0.1 (sm_100a)"; }

extern "C" void b200spmv_default_opts(b200spmv_opts * o)
{
	memset(o, 0, sizeof(*o));
	o->format = B200SPMV_CSR_ADAPTIVE;
	o->device = -1;
	o->sell_C = 32;
	o->sell_sigma = 16384;
}

static int check_device(int device)
{
	int ndev = 0;
	cudaError_t e = cudaGetDeviceCount(&ndev);
	if (e != cudaSuccess || ndev < 1)
		B200_FAIL("no CUDA device: libb200spmv has no CPU fallback (%s)", cudaGetErrorString(e));
	if (device >= ndev)
		B200_FAIL("device %d out of range (%d devices)", device, ndev);
	if (device >= 0)
		CU(cudaSetDevice(device));
	int cur = 0, major = 0;
	CU(cudaGetDevice(&cur));
	CU(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, cur));
	if (major != 10)
		B200_FAIL("device %d has compute capability %d.x; this library carries sm_100a code only", cur, major);
	return 0;
}

static bool is_pow2(int64_t v) { return v > 0 && (v & (v - 1)) == 0; }

static int validate_opts(const b200spmv_opts * in, b200spmv_opts * o, int64_t m, int64_t n, int64_t nnz)
{
	if (in)
		*o = *in;
	else
		b200spmv_default_opts(o);
	if (m < 0 || n < 0 || nnz < 0)
		B200_FAIL("negative dimension");
	if (m >= 0x7fffffffLL || n > 0x7fffffffLL || nnz > 0x7fffffffLL)
		B200_FAIL("m, n and nnz must fit int32 indices (INT_T = int32_t)");
	if (o->format != B200SPMV_CSR_ADAPTIVE && o->format != B200SPMV_SELL_C_SIGMA && o->format != B200SPMV_AUTO)
		B200_FAIL("unknown format %d", o->format);
	if (o->sell_sigma_alt == 0)
		o->sell_sigma_alt = 256;
	if (o->col_panels < 0 || o->col_panels > 8)
		B200_FAIL("col_panels must be in [0, 8]");
	if (o->csr_compress_rows < 0 || o->csr_compress_rows > 2)
		B200_FAIL("csr_compress_rows must be 0 (auto), 1 (never) or 2 (always)");
	if (o->sell_variant < 0 || o->sell_variant > 4)
		B200_FAIL("sell_variant must be in [0, 4]");
	if (o->sell_C == 0)
		o->sell_C = 32;
	if (o->sell_sigma == 0)
		o->sell_sigma = 16384;
	if (!is_pow2(o->sell_C) || o->sell_C < 2 || o->sell_C > 256)
		B200_FAIL("sell_C must be a power of two in [2, 256]");
	if (o->sell_sigma < o->sell_C || o->sell_sigma % o->sell_C)
		B200_FAIL("sell_sigma must be a positive multiple of sell_C");
	if (o->csr_nnz_per_thread == 0) {
		// auto: larger windows for longer rows keep the tail extension a small fraction of the window
		// measured on B200 (profiles/r01e_sweep.txt): 2 048-nnz tiles win for every row length, also avg 500
		o->csr_nnz_per_thread = 8;
	}
	if (o->csr_nnz_per_thread != 4 && o->csr_nnz_per_thread != 8 && o->csr_nnz_per_thread != 16 && o->csr_nnz_per_thread != 32)
		B200_FAIL("csr_nnz_per_thread must be 4, 8, 16 or 32");
	if (o->csr_ext == 0)
		o->csr_ext = (o->csr_nnz_per_thread <= 8) ? 256 : 2048;     // shared memory comes out of the L1 the x gathers live in
	if (o->csr_ext < 64 || o->csr_ext % 64 || o->csr_ext > 8192)
		B200_FAIL("csr_ext must be a multiple of 64 in [64, 8192]");
	return 0;
}

// Smallest / largest column index present: x entries outside [lo, hi] are never read by any product of this matrix.
__global__ void col_range_kernel(const int32_t * __restrict__ col, int64_t nnz, int * __restrict__ lo_hi)
{
	int lo = 0x7fffffff, hi = -1;
	for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (int64_t) gridDim.x * blockDim.x) {
		const int c = col[i];
		lo = c < lo ? c : lo;
		hi = c > hi ? c : hi;
	}
	for (int o = 16; o > 0; o >>= 1) {
		const int l2 = __shfl_xor_sync(0xffffffffu, lo, o), h2 = __shfl_xor_sync(0xffffffffu, hi, o);
		lo = l2 < lo ? l2 : lo;
		hi = h2 > hi ? h2 : hi;
	}
	if ((threadIdx.x & 31) == 0) {
		atomicMin(&lo_hi[0], lo);
		atomicMax(&lo_hi[1], hi);
	}
}

static int compute_col_range(b200spmv * h)
{
	h->col_lo = 0;
	h->col_hi = 0;
	if (h->nnz == 0)
		return 0;
	int * d = nullptr;
	int init[2] = { 0x7fffffff, -1 }, got[2];
	CU(cudaMalloc(&d, 2 * sizeof(int)));
	CU(cudaMemcpyAsync(d, init, sizeof(init), cudaMemcpyHostToDevice, h->stream));
	col_range_kernel<<<148 * 8, 256, 0, h->stream>>>(h->d_col, h->nnz, d);
	CU_LAUNCH();
	CU(cudaMemcpyAsync(got, d, sizeof(got), cudaMemcpyDeviceToHost, h->stream));
	CU(cudaStreamSynchronize(h->stream));
	CU(cudaFree(d));
	if (got[0] < 0 || got[1] >= h->n)
		B200_FAIL("col_ind out of range: found [%d, %d], n = %lld", got[0], got[1], (long long) h->n);
	h->col_lo = got[0];
	h->col_hi = (int64_t) got[1] + 1;
	return 0;
}

// host_pipeline: column range of each block of blk_rows consecutive rows (one CTA per block), so that the host entry point
// knows which part of x a unit range of rows needs.  lo_hi[2 b] = smallest, lo_hi[2 b + 1] = largest column (lo > hi: none).
__global__ void __launch_bounds__(256)
block_col_range_kernel(const int32_t * __restrict__ row_ptr, const int32_t * __restrict__ col, int64_t m, int64_t blk_rows, int * __restrict__ lo_hi)
{
	__shared__ int s_lo[8], s_hi[8];
	const int64_t r0 = (int64_t) blockIdx.x * blk_rows, r1 = r0 + blk_rows < m ? r0 + blk_rows : m;
	int lo = 0x7fffffff, hi = -1;
	if (r0 < m)
		for (int64_t j = (int64_t) row_ptr[r0] + threadIdx.x; j < row_ptr[r1]; j += 256) {
			const int c = col[j];
			lo = c < lo ? c : lo;
			hi = c > hi ? c : hi;
		}
	for (int o = 16; o > 0; o >>= 1) {
		const int l2 = __shfl_xor_sync(0xffffffffu, lo, o), h2 = __shfl_xor_sync(0xffffffffu, hi, o);
		lo = l2 < lo ? l2 : lo;
		hi = h2 > hi ? h2 : hi;
	}
	if ((threadIdx.x & 31) == 0) {
		s_lo[threadIdx.x >> 5] = lo;
		s_hi[threadIdx.x >> 5] = hi;
	}
	__syncthreads();
	if (threadIdx.x == 0) {
		for (int w = 1; w < 8; w++) {
			lo = s_lo[w] < lo ? s_lo[w] : lo;
			hi = s_hi[w] > hi ? s_hi[w] : hi;
		}
		lo_hi[2 * blockIdx.x] = lo;
		lo_hi[2 * blockIdx.x + 1] = hi;
	}
}

static int compute_block_col_ranges(b200spmv * h)
{
	const int NB = 256;
	h->blk_rows = 0;
	if (!h->pipeline || h->m < 262144 || h->nnz == 0)
		return 0;
	const int64_t blk_rows = (h->m + NB - 1) / NB;
	const int nb = (int) ((h->m + blk_rows - 1) / blk_rows);
	int * d = nullptr;
	int got[2 * NB];
	CU(cudaMalloc(&d, 2 * NB * sizeof(int)));
	block_col_range_kernel<<<nb, 256, 0, h->stream>>>(h->d_row_ptr, h->d_col, h->m, blk_rows, d);
	CU_LAUNCH();
	CU(cudaMemcpyAsync(got, d, 2 * nb * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
	CU(cudaStreamSynchronize(h->stream));
	CU(cudaFree(d));
	for (int b = 0; b < NB; b++) {
		h->blk_lo[b] = b < nb ? got[2 * b] : 0x7fffffff;
		h->blk_hi[b] = b < nb ? got[2 * b + 1] : -1;
	}
	h->blk_rows = blk_rows;
	return 0;
}

// row_ptr must start at 0, never decrease and end at nnz: a malformed array would send the kernels out of bounds
__global__ void validate_row_ptr_kernel(const int32_t * __restrict__ rp, int64_t m, int64_t nnz, int * __restrict__ bad)
{
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i > m)
		return;
	bool b = (i < m) ? (rp[i] > rp[i + 1]) : ((int64_t) rp[m] != nnz);
	if (i == 0)
		b = b || rp[0] != 0;
	if (b)
		atomicExch(bad, 1);
}

static int validate_row_ptr(b200spmv * h)
{
	int * d = nullptr, bad = 0;
	CU(cudaMalloc(&d, sizeof(int)));
	CU(cudaMemsetAsync(d, 0, sizeof(int), h->stream));
	validate_row_ptr_kernel<<<(unsigned) ((h->m + 1 + 255) / 256), 256, 0, h->stream>>>(h->d_row_ptr, h->m, h->nnz, d);
	CU_LAUNCH();
	CU(cudaMemcpyAsync(&bad, d, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
	CU(cudaStreamSynchronize(h->stream));
	CU(cudaFree(d));
	if (bad)
		B200_FAIL("row_ptr is malformed: it must start at 0, be non-decreasing and end at nnz = %lld", (long long) h->nnz);
	return 0;
}

__global__ void rebase_row_ptr_kernel(const int32_t * __restrict__ in, int32_t * __restrict__ out, int64_t m1)
{
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < m1)
		out[i] = in[i] - in[0];
}

static void free_sell(b200spmv * h)
{
	void ** ptrs[] = { (void **) &h->d_row_order, (void **) &h->d_slice_width, (void **) &h->d_slice_ptr, (void **) &h->d_scol,
	                   (void **) &h->d_sval, (void **) &h->d_sell_items, (void **) &h->d_sell_cuts, (void **) &h->d_sell_partial,
	                   (void **) &h->d_sell_slot_cut, (void **) &h->d_sell_cut_cnt };
	for (void ** p : ptrs) {
		if (*p)
			cudaFree(*p);
		*p = nullptr;
	}
	h->sell_nitems = h->sell_ncuts = h->sell_nslots = 0;
	h->nslices = h->padded = 0;
}

static void free_csr_plan(b200spmv * h, bool arrays)
{
	void ** ptrs[] = { (void **) &h->d_tile_row, (void **) &h->d_split_row, (void **) &h->d_split_tile, (void **) &h->d_tile_head,
	                   (void **) &h->d_tile_tail, (void **) &h->d_tile_split, (void **) &h->d_split_info, (void **) &h->d_split_cnt };
	for (void ** p : ptrs) {
		if (*p)
			cudaFree(*p);
		*p = nullptr;
	}
	void ** more[] = { (void **) &h->d_rp_nz, (void **) &h->d_rows_nz, (void **) &h->d_empty_rows };
	for (void ** p : more) {
		if (*p)
			cudaFree(*p);
		*p = nullptr;
	}
	h->d_rp_plan = nullptr;
	h->compressed = 0;
	h->n_empty = 0;
	if (arrays) {
		if (h->d_col) cudaFree(h->d_col);
		if (h->d_val) cudaFree(h->d_val);
		h->d_col = nullptr;
		h->d_val = nullptr;
	}
}

constexpr size_t TRIAL_FLUSH_BYTES = 192u << 20;

// AUTO trial: median of 5 device-timed products (after 2 untimed) of the format h->format currently names.
static int trial_ms(b200spmv * h, cudaEvent_t e0, cudaEvent_t e1, float * out)
{
	float t[5];
	for (int i = 0; i < 7; i++) {
		// a matrix larger than L2 is timed cold (what a sweep over many matrices / a long vector pipeline sees): rewrite a
		// scratch buffer larger than the 126 MB L2 before every product, outside the event pair
		if (h->d_trial_flush)
			CU(cudaMemsetAsync(h->d_trial_flush, i, TRIAL_FLUSH_BYTES, h->stream));
		if (i >= 2)
			CU(cudaEventRecord(e0, h->stream));
		const int rc = (h->format == B200SPMV_CSR_ADAPTIVE) ? launch_csr_spmv(h, h->d_x, h->d_y, h->stream, nullptr)
		                                                    : launch_sell_spmv(h, h->d_x, h->d_y, h->stream, nullptr);
		if (rc)
			return 1;
		if (i >= 2) {
			CU(cudaEventRecord(e1, h->stream));
			CU(cudaEventSynchronize(e1));
			CU(cudaEventElapsedTime(&t[i - 2], e0, e1));
		}
	}
	for (int i = 1; i < 5; i++)
		for (int j = i; j > 0 && t[j] < t[j - 1]; j--) {
			const float x = t[j]; t[j] = t[j - 1]; t[j - 1] = x;
		}
	*out = t[2];
	return 0;
}

// B200SPMV_AUTO: plan CSR, convert SELL-C-sigma for each candidate window, time each on the device, keep the fastest.
// The gather pattern depends on col_ind only, so a zero x vector times the same memory traffic as a real one.
static int auto_select_impl(b200spmv * h, const b200spmv_opts & o, cudaEvent_t e0, cudaEvent_t e1);

static int auto_select(b200spmv * h, const b200spmv_opts & o)
{
	cudaEvent_t e0 = nullptr, e1 = nullptr;
	CU(cudaEventCreate(&e0));
	CU(cudaEventCreate(&e1));
	const int rc = auto_select_impl(h, o, e0, e1);
	if (h->d_trial_flush)
		cudaFree(h->d_trial_flush);
	h->d_trial_flush = nullptr;
	cudaEventDestroy(e0);
	cudaEventDestroy(e1);
	return rc;
}

static int auto_select_impl(b200spmv * h, const b200spmv_opts & o, cudaEvent_t e0, cudaEvent_t e1)
{
	B200Range nvtx_range_("b200spmv AUTO trial");
	CU(cudaMalloc(&h->d_x, (h->n > 0 ? h->n : 1) * sizeof(double)));
	CU(cudaMalloc(&h->d_y, (h->m > 0 ? h->m : 1) * sizeof(double)));
	CU(cudaMemsetAsync(h->d_x, 0, (h->n > 0 ? h->n : 1) * sizeof(double), h->stream));
	if (12 * h->nnz > (64LL << 20))
		CU(cudaMalloc(&h->d_trial_flush, TRIAL_FLUSH_BYTES));
	h->format = B200SPMV_CSR_ADAPTIVE;
	if (csr_plan_tiles(h))
		return 1;
	if (h->nnz == 0 || h->m == 0)
		return 0;                                   // nothing to time: CSR
	if (trial_ms(h, e0, e1, &h->trial_ms[0]))
		return 1;
	float best = h->trial_ms[0];
	int64_t best_sigma = 0;                         // 0 = CSR
	int best_variant = o.sell_variant;
	int64_t cand[2] = { o.sell_sigma, o.sell_sigma_alt > 0 ? (int64_t) o.sell_sigma_alt : 0 };
	if (cand[1] == cand[0] || (cand[1] > 0 && (cand[1] < h->C || cand[1] % h->C)))
		cand[1] = 0;
	h->sell_pad_limit = h->nnz + h->nnz / 2 + 64 * (int64_t) h->C;      // > 50 % padding cannot win: skip the candidate
	h->format = B200SPMV_SELL_C_SIGMA;
	for (int c = 0; c < 2; c++) {
		if (cand[c] <= 0)
			continue;
		h->sigma = cand[c];
		const int rc = sell_convert(h);
		if (rc == 1)
			return 1;
		if (rc == 0) {
			// both kernel shapes that win somewhere (8 k-steps in flight x 32 or 48 resident warps per SM) unless one is forced
			float tv[2] = { 0.f, 0.f };
			const int forced = o.sell_variant;
			for (int v = 0; v < 2; v++) {
				if (forced != 0 && v > 0)
					break;
				h->sell_variant = forced != 0 ? forced : 1 + v;
				if (trial_ms(h, e0, e1, &tv[v]))
					return 1;
			}
			const int bv = (forced == 0 && tv[1] < tv[0]) ? 1 : 0;
			h->trial_ms[1 + c] = tv[bv];
			if (tv[bv] < best) {
				best = tv[bv];
				best_sigma = cand[c];
				best_variant = forced != 0 ? forced : 1 + bv;
			}
		}
		// keep the arrays only when this is the last candidate and it is the winner; otherwise rebuild the winner below
		if (!(best_sigma == cand[c] && (c == 1 || cand[1] <= 0)))
			free_sell(h);
	}
	h->sell_pad_limit = 0;
	if (best_sigma == 0) {
		free_sell(h);
		h->format = B200SPMV_CSR_ADAPTIVE;
		h->sigma = o.sell_sigma;
		return 0;
	}
	h->sigma = best_sigma;
	h->sell_variant = best_variant;
	if (!h->d_scol && sell_convert(h))
		return 1;
	free_csr_plan(h, true);
	h->ntiles = h->nsplit = 0;
	return 0;
}

// Ranges of the host-vector path (see common.cuh): only when every row is finished by the main kernel alone.
static int plan_chunks(b200spmv * h)
{
	h->nchunks = 0;
	h->range_count = -1;
	const int NCH = 4;
	if (h->m < 262144 || h->npanels > 1)
		return 0;
	h->n_chunk_fix = 0;
	if (h->format == B200SPMV_CSR_ADAPTIVE) {
		if (h->compressed || h->ntiles < 64 * NCH || (h->nsplit > 0 && !h->fold_splits))
			return 0;
		for (int c = 0; c <= NCH; c++) {
			h->chunk_unit[c] = h->ntiles * c / NCH;
			int2 ti;
			CU(cudaMemcpy(&ti, reinterpret_cast<const int2 *>(h->d_tile_row) + h->chunk_unit[c], sizeof(int2), cudaMemcpyDeviceToHost));
			h->chunk_row[c] = ti.x & 0x7fffffff;
		}
		h->chunk_row[0] = 0;
		h->chunk_row[NCH] = h->m;
		if (h->nsplit > 0) {
			// a split row is finished by the LAST of its tiles (csr_fold_split_rows); when those tiles straddle a range boundary
			// the row is final only after a later range's launch: at most one such row per boundary, copied again at the end
			int4 * info = (int4 *) malloc((size_t) h->nsplit * sizeof(int4));
			if (!info)
				B200_FAIL("out of host memory");
			if (cudaMemcpy(info, h->d_split_info, (size_t) h->nsplit * sizeof(int4), cudaMemcpyDeviceToHost) != cudaSuccess) {
				free(info);
				B200_FAIL("plan_chunks: copy of the split-row table failed");
			}
			for (int64_t s = 0; s < h->nsplit; s++) {
				const int64_t k = info[s].y, last = k + info[s].z - 1;
				for (int c = 1; c < NCH; c++)
					if (k < h->chunk_unit[c] && last >= h->chunk_unit[c] && h->n_chunk_fix < 4) {
						h->chunk_fix_row[h->n_chunk_fix++] = info[s].x;
						break;
					}
			}
			free(info);
		}
	} else {
		const int64_t spw = h->sigma / h->C, nwin = (h->m + h->sigma - 1) / h->sigma;
		if (nwin < 2 * NCH)
			return 0;
		for (int c = 0; c <= NCH; c++) {
			const int64_t s = (nwin * c / NCH) * spw;
			h->chunk_unit[c] = h->sell_chunk_item[c];
			h->chunk_row[c] = s * h->C < h->m ? s * h->C : h->m;
		}
		h->chunk_unit[NCH] = h->sell_nitems;
		h->chunk_row[NCH] = h->m;
	}
	CU(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
	for (int c = 0; c < NCH; c++)
		CU(cudaEventCreateWithFlags(&h->chunk_ev[c], cudaEventDisableTiming));
	h->nchunks = NCH;
	if (h->pipeline && h->blk_rows > 0) {
		// columns referenced by the rows of each unit range (SELL ranges end on sort-window boundaries, so a range
		// holds the same set of rows before and after sorting): union of the row blocks it touches
		for (int c = 0; c < NCH; c++) {
			int lo = 0x7fffffff, hi = -1;
			const int64_t r0 = h->chunk_row[c], r1 = h->chunk_row[c + 1];
			if (r1 > r0)
				for (int64_t b = r0 / h->blk_rows; b <= (r1 - 1) / h->blk_rows && b < 256; b++) {
					lo = h->blk_lo[b] < lo ? h->blk_lo[b] : lo;
					hi = h->blk_hi[b] > hi ? h->blk_hi[b] : hi;
				}
			h->chunk_col_lo[c] = lo <= hi ? lo : 0;
			h->chunk_col_hi[c] = lo <= hi ? (int64_t) hi + 1 : 0;
		}
		CU(cudaStreamCreateWithFlags(&h->up_stream, cudaStreamNonBlocking));
		for (int c = 0; c < NCH; c++)
			CU(cudaEventCreateWithFlags(&h->up_ev[c], cudaEventDisableTiming));
	} else {
		h->pipeline = 0;
	}
	return 0;
}

static void free_handle(b200spmv * h)
{
	if (!h)
		return;
	for (int q = 0; q < h->npanels; q++)
		free_handle(h->panel[q]);
	for (int i = 0; i < h->n_pinned; i++)
		cudaHostUnregister(h->pinned[i]);
	if (h->copy_stream)
		cudaStreamDestroy(h->copy_stream);
	if (h->up_stream)
		cudaStreamDestroy(h->up_stream);
	for (cudaEvent_t e : h->chunk_ev)
		if (e)
			cudaEventDestroy(e);
	for (cudaEvent_t e : h->up_ev)
		if (e)
			cudaEventDestroy(e);
	void * ptrs[] = { h->d_row_ptr, h->d_col, h->d_val, h->d_tile_row, h->d_split_row, h->d_split_tile, h->d_tile_head, h->d_tile_tail, h->d_perm, h->d_perm_off, h->d_rot, h->d_tile_split, h->d_split_info, h->d_split_cnt,
	                  h->d_row_order, h->d_slice_width, h->d_slice_ptr, h->d_scol, h->d_sval, h->d_x, h->d_y, h->d_norm_partial,
	                  h->d_sell_items, h->d_sell_cuts, h->d_sell_partial, h->d_sell_slot_cut, h->d_sell_cut_cnt, h->d_rp_nz, h->d_rows_nz, h->d_empty_rows, h->d_diag };
	for (void * p : ptrs)
		if (p)
			cudaFree(p);
	if (h->stream)
		cudaStreamDestroy(h->stream);
	free(h);
}

// Builds a handle from CSR arrays; `kind` says where the inputs live.
static int create_common(b200spmv ** out, const int32_t * row_ptr, const int32_t * col_ind, const double * values,
                         int64_t m, int64_t n, int64_t nnz, const b200spmv_opts * opts_in, cudaMemcpyKind kind)
{
	B200Range nvtx_range_("b200spmv_create");
	if (!out)
		B200_FAIL("out is NULL");
	*out = nullptr;
	if (!row_ptr || (nnz > 0 && (!col_ind || !values)))
		B200_FAIL("NULL CSR array");
	b200spmv_opts o;
	if (validate_opts(opts_in, &o, m, n, nnz))
		return 1;
	if (check_device(o.device))
		return 1;
	b200spmv * h = (b200spmv *) calloc(1, sizeof(b200spmv));
	if (!h)
		B200_FAIL("out of host memory");
	h->format = o.format;
	h->requested_format = o.format;
	h->range_count = -1;
	CU(cudaGetDevice(&h->device));
	h->m = m;
	h->n = n;
	h->nnz = nnz;
	h->x_hint = o.x_hint;
	h->npt = o.csr_nnz_per_thread;
	h->tile_nnz = CSR_THREADS * h->npt;
	h->ext = o.csr_ext;
	h->C = o.sell_C;
	h->sigma = o.sell_sigma;
	h->sell_variant = o.sell_variant;
	h->compress_opt = o.csr_compress_rows;
	h->pin_host = o.pin_host_vectors;
	h->pipeline = o.host_pipeline ? 1 : 0;
	int rc = 1;
	cudaEvent_t e0 = nullptr, e1 = nullptr;
	do {
		if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess)
			break;
		cudaEventCreate(&e0);
		cudaEventCreate(&e1);
		cudaEventRecord(e0, h->stream);
		// CSR arrays, zero-padded so that whole windows (+ extension) can be loaded unconditionally
		const int64_t tiles = (nnz + h->tile_nnz - 1) / h->tile_nnz;
		const int64_t padded = (tiles > 0 ? tiles : 1) * h->tile_nnz + h->ext + 64;
		if (cudaMalloc(&h->d_row_ptr, (m + 1) * sizeof(int32_t)) != cudaSuccess) break;
		if (cudaMalloc(&h->d_col, padded * sizeof(int32_t)) != cudaSuccess) break;
		if (cudaMalloc(&h->d_val, padded * sizeof(double)) != cudaSuccess) break;
		cudaMemsetAsync(h->d_col + nnz, 0, (padded - nnz) * sizeof(int32_t), h->stream);
		cudaMemsetAsync(h->d_val + nnz, 0, (padded - nnz) * sizeof(double), h->stream);
		if (kind == cudaMemcpyHostToDevice) {
			if (cudaMemcpyAsync(h->d_row_ptr, row_ptr, (m + 1) * sizeof(int32_t), kind, h->stream) != cudaSuccess) break;
			if (m >= 0 && row_ptr[0] != 0) {
				snprintf(g_b200_err, sizeof(g_b200_err), "row_ptr[0] must be 0 for host input");
				goto fail;
			}
		} else {
			rebase_row_ptr_kernel<<<(unsigned) ((m + 1 + 255) / 256), 256, 0, h->stream>>>(row_ptr, h->d_row_ptr, m + 1);
		}
		if (nnz > 0) {
			if (cudaMemcpyAsync(h->d_col, col_ind, nnz * sizeof(int32_t), kind, h->stream) != cudaSuccess) break;
			if (cudaMemcpyAsync(h->d_val, values, nnz * sizeof(double), kind, h->stream) != cudaSuccess) break;
		}
		if (cudaStreamSynchronize(h->stream) != cudaSuccess) break;
		if (validate_row_ptr(h)) goto fail;
		if (compute_col_range(h)) goto fail;
		if (compute_block_col_ranges(h)) goto fail;
		if (o.keep_diag && pcg_extract_diag(h)) goto fail;
		int64_t norm_slots;
		const int P = panels_auto_count(h, o.col_panels);
		if (P > 1) {
			// column panels: this handle becomes a shell around P sub-handles (panels.cu)
			b200spmv_opts sub = o;
			sub.col_panels = 1;
			sub.keep_diag = 0;                  // (a panel may compress its rows: csr_zero_rows_kernel passes the rows that are empty
			                                    // in the panel through y_finish like the tile kernel does -- accumulate / scale / norm)
			sub.host_pipeline = 0;              // the shell's host entry point uploads the band once
			sub.device = -1;
			if (panels_split(h, P, &sub)) goto fail;
			cudaFree(h->d_col);
			cudaFree(h->d_val);
			h->d_col = nullptr;
			h->d_val = nullptr;
			h->format = B200SPMV_CSR_ADAPTIVE;
			for (int q = 0; q < h->npanels; q++) {
				h->panel[q]->acc_mode = (q > 0 ? 1 : 0) | (q + 1 < h->npanels ? 2 : 0);
				if (h->panel[q]->format == B200SPMV_SELL_C_SIGMA)
					h->format = B200SPMV_SELL_C_SIGMA;      // "some panel reads x[0] for its padding"
			}
			norm_slots = 1;
		} else if (h->format == B200SPMV_AUTO) {
			if (auto_select(h, o)) goto fail;           // convert_ms of an AUTO handle includes the trials
			norm_slots = (h->format == B200SPMV_CSR_ADAPTIVE) ? h->ntiles + h->nsplit + CSR_ZERO_CTAS
			                                                   : (h->sell_nitems + 3) / 4 + h->sell_ncuts;
		} else if (h->format == B200SPMV_CSR_ADAPTIVE) {
			if (csr_plan_tiles(h)) goto fail;
			norm_slots = h->ntiles + h->nsplit + CSR_ZERO_CTAS;
		} else {
			if (sell_convert(h)) goto fail;
			// the CSR copies of col / val are not needed any more
			cudaFree(h->d_col);
			cudaFree(h->d_val);
			h->d_col = nullptr;
			h->d_val = nullptr;
			norm_slots = (h->sell_nitems + 3) / 4 + h->sell_ncuts;
		}
		if (plan_chunks(h)) goto fail;
		h->norm_slots = norm_slots > 0 ? norm_slots : 1;
		if (cudaMalloc(&h->d_norm_partial, h->norm_slots * sizeof(double)) != cudaSuccess) break;
		cudaEventRecord(e1, h->stream);
		if (cudaStreamSynchronize(h->stream) != cudaSuccess) break;
		cudaEventElapsedTime(&h->convert_ms, e0, e1);
		rc = 0;
	} while (0);
	if (rc) {
		snprintf(g_b200_err, sizeof(g_b200_err), "b200spmv_create: CUDA failure: %s", cudaGetErrorString(cudaGetLastError()));
fail:
		if (e0) cudaEventDestroy(e0);
		if (e1) cudaEventDestroy(e1);
		free_handle(h);
		return 1;
	}
	cudaEventDestroy(e0);
	cudaEventDestroy(e1);
	*out = h;
	return 0;
}

extern "C" int b200spmv_create(b200spmv_t ** out, const int32_t * row_ptr, const int32_t * col_ind, const double * values,
                               int64_t m, int64_t n, int64_t nnz, const b200spmv_opts * opts)
{
	return create_common(out, row_ptr, col_ind, values, m, n, nnz, opts, cudaMemcpyHostToDevice);
}

extern "C" int b200spmv_create_device(b200spmv_t ** out, const int32_t * d_row_ptr, const int32_t * d_col_ind, const double * d_values,
                                      int64_t m, int64_t n, int64_t nnz, const b200spmv_opts * opts)
{
	return create_common(out, d_row_ptr, d_col_ind, d_values, m, n, nnz, opts, cudaMemcpyDeviceToDevice);
}

extern "C" void b200spmv_destroy(b200spmv_t * h)
{
	if (h) {
		cudaSetDevice(h->device);
		free_handle(h);
	}
}

extern "C" int32_t b200spmv_format(const b200spmv_t * h) { return h ? h->format : -1; }

extern "C" int b200spmv_col_range(const b200spmv_t * h, int64_t * lo, int64_t * hi)
{
	if (!h || !lo || !hi)
		B200_FAIL("NULL argument");
	*lo = h->col_lo;
	*hi = h->col_hi;
	return 0;
}

extern "C" int32_t b200spmv_launches_per_spmv(const b200spmv_t * h)
{
	if (!h)
		return 0;
	if (h->npanels > 1) {
		int32_t n = 0;
		for (int q = 0; q < h->npanels; q++)
			n += b200spmv_launches_per_spmv(h->panel[q]);
		return n;
	}
	if (h->format == B200SPMV_CSR_ADAPTIVE)
		return 1 + ((h->nsplit > 0 && !h->fold_splits) ? 1 : 0) + ((h->compressed && h->n_empty > 0) ? 1 : 0);
	return 1;                                   // cut slices are finished inside the kernel (last-arriving warp)
}

// One product: a plain handle launches its kernel(s); a panel shell launches its sub-handles in ascending column order, the
// last of which receives the caller's epilogue (scale, norm, peer destinations).
int launch_product(b200spmv * h, const double * dx, double * dy, cudaStream_t st, const b200spmv_launch * l)
{
	if (h->npanels > 1) {
		for (int q = 0; q < h->npanels; q++) {
			b200spmv * s = h->panel[q];
			const b200spmv_launch * lq = (q + 1 == h->npanels) ? l : nullptr;
			const int rc = (s->format == B200SPMV_CSR_ADAPTIVE) ? launch_csr_spmv(s, dx, dy, st, lq) : launch_sell_spmv(s, dx, dy, st, lq);
			if (rc)
				return rc;
		}
		return 0;
	}
	return (h->format == B200SPMV_CSR_ADAPTIVE) ? launch_csr_spmv(h, dx, dy, st, l) : launch_sell_spmv(h, dx, dy, st, l);
}

int launch_spmv_any(b200spmv * h, const double * dx, double * dy, cudaStream_t st)
{
	return launch_product(h, dx, dy, st, nullptr);
}

int create_from_device(b200spmv ** out, const int32_t * d_row_ptr, const int32_t * d_col, const double * d_val, int64_t m, int64_t n,
                       int64_t nnz, const b200spmv_opts * opts)
{
	return b200spmv_create_device(out, d_row_ptr, d_col, d_val, m, n, nnz, opts);
}

void destroy_handle(b200spmv * h)
{
	b200spmv_destroy(h);
}

extern "C" int b200spmv_spmv_device_ex(b200spmv_t * h, const double * dx, double * dy, void * stream, const b200spmv_launch * launch)
{
	if (!h || !dx || !dy)
		B200_FAIL("NULL argument");
	cudaStream_t st = (cudaStream_t) stream;       // NULL is the CUDA legacy default stream (what torch's default stream is)
	return launch_product(h, dx, dy, st, launch);
}

extern "C" int b200spmv_spmv_device(b200spmv_t * h, const double * dx, double * dy, void * stream)
{
	return b200spmv_spmv_device_ex(h, dx, dy, stream, nullptr);
}

// Page-locks a caller buffer once (pageable memory moves at a third of the PCIe rate and cannot overlap with kernels).
static void pin_once(b200spmv * h, const void * p, size_t bytes)
{
	if (!h->pin_host || bytes == 0)
		return;
	for (int i = 0; i < h->n_pinned; i++)
		if (h->pinned[i] == p)
			return;
	if (h->n_pinned >= 4)
		return;
	cudaPointerAttributes a;
	if (cudaPointerGetAttributes(&a, p) == cudaSuccess && a.type != cudaMemoryTypeUnregistered)
		return;                                    // already pinned / managed by the caller
	cudaGetLastError();
	if (cudaHostRegister(const_cast<void *>(p), bytes, cudaHostRegisterDefault) == cudaSuccess)
		h->pinned[h->n_pinned++] = const_cast<void *>(p);
	else
		cudaGetLastError();                        // not fatal: the copies fall back to the pageable path
}

extern "C" int b200spmv_spmv_host_ex(b200spmv_t * h, const double * x, double * y, int32_t flags)
{
	B200Range nvtx_range_("b200spmv_spmv_host");
	if (!h || !x || !y)
		B200_FAIL("NULL argument");
	CU(cudaSetDevice(h->device));
	if (!(flags & B200SPMV_SKIP_X_UPLOAD))
		pin_once(h, x, (size_t) h->n * sizeof(double));
	if (!(flags & B200SPMV_SKIP_Y_DOWNLOAD))
		pin_once(h, y, (size_t) h->m * sizeof(double));
	if (!h->d_x) {
		CU(cudaMalloc(&h->d_x, (h->n > 0 ? h->n : 1) * sizeof(double)));
		CU(cudaMalloc(&h->d_y, (h->m > 0 ? h->m : 1) * sizeof(double)));
		CU(cudaMemsetAsync(h->d_x, 0, (h->n > 0 ? h->n : 1) * sizeof(double), h->stream));
		if (h->up_stream)
			CU(cudaStreamSynchronize(h->stream));   // the range-by-range uploads run on another stream
	}
	if (!h->x_uploaded)
		flags &= ~B200SPMV_SKIP_X_UPLOAD;           // nothing to reuse yet
	h->x_uploaded = 1;
	// host_pipeline: x is uploaded range by range inside the chunk loop below
	const bool pipe = h->pipeline && h->up_stream && h->nchunks > 1 && !(flags & (B200SPMV_SKIP_X_UPLOAD | B200SPMV_SKIP_Y_DOWNLOAD));
	// only the columns the matrix references are read by the kernels: upload that band of x, not all n entries
	if (!pipe && !(flags & B200SPMV_SKIP_X_UPLOAD) && h->col_hi > h->col_lo)
		CU(cudaMemcpyAsync(h->d_x + h->col_lo, x + h->col_lo, (h->col_hi - h->col_lo) * sizeof(double), cudaMemcpyHostToDevice, h->stream));
	// SELL-C-sigma padding entries are (column 0, 0.0) and are multiplied like real ones (sellcs_format.c:150-158): x[0] is read
	// (host_pipeline: always first -- the range that brings column 0 may come after a range whose padding already reads it)
	if (!(flags & B200SPMV_SKIP_X_UPLOAD) && h->format == B200SPMV_SELL_C_SIGMA && (h->col_lo > 0 || pipe) && h->n > 0)
		CU(cudaMemcpyAsync(h->d_x, x, sizeof(double), cudaMemcpyHostToDevice, h->stream));
	if (h->nchunks > 1 && !(flags & B200SPMV_SKIP_Y_DOWNLOAD)) {
		// the rows of a unit range are final when its kernel ends: their D2H copy runs under the next range's kernel
		int rc = 0;
		int64_t up_lo = 0, up_hi = 0;                // columns of x on the device so far (one interval; empty at first)
		for (int c = 0; c < h->nchunks && !rc; c++) {
			if (pipe) {
				// what range c reads and is not there yet; the interval stays contiguous, inside [col_lo, col_hi)
				const int64_t lo = h->chunk_col_lo[c], hi = h->chunk_col_hi[c];
				if (hi > lo) {
					if (up_hi == up_lo) {
						CU(cudaMemcpyAsync(h->d_x + lo, x + lo, (hi - lo) * sizeof(double), cudaMemcpyHostToDevice, h->up_stream));
						up_lo = lo;
						up_hi = hi;
					} else {
						if (lo < up_lo) {
							CU(cudaMemcpyAsync(h->d_x + lo, x + lo, (up_lo - lo) * sizeof(double), cudaMemcpyHostToDevice, h->up_stream));
							up_lo = lo;
						}
						if (hi > up_hi) {
							CU(cudaMemcpyAsync(h->d_x + up_hi, x + up_hi, (hi - up_hi) * sizeof(double), cudaMemcpyHostToDevice, h->up_stream));
							up_hi = hi;
						}
					}
				}
				CU(cudaEventRecord(h->up_ev[c], h->up_stream));
				CU(cudaStreamWaitEvent(h->stream, h->up_ev[c], 0));
			}
			h->range_first = h->chunk_unit[c];
			h->range_count = h->chunk_unit[c + 1] - h->chunk_unit[c];
			rc = launch_spmv_any(h, h->d_x, h->d_y, h->stream);
			h->range_count = -1;
			if (rc)
				break;
			const int64_t r0 = h->chunk_row[c], r1 = h->chunk_row[c + 1];
			CU(cudaEventRecord(h->chunk_ev[c], h->stream));
			CU(cudaStreamWaitEvent(h->copy_stream, h->chunk_ev[c], 0));
			if (r1 > r0)
				CU(cudaMemcpyAsync(y + r0, h->d_y + r0, (r1 - r0) * sizeof(double), cudaMemcpyDeviceToHost, h->copy_stream));
		}
		h->range_count = -1;
		if (rc)
			return rc;
		// split rows finished by a later range than the one that owns them (plan_chunks): the copy stream is past the last range
		for (int f = 0; f < h->n_chunk_fix; f++)
			CU(cudaMemcpyAsync(y + h->chunk_fix_row[f], h->d_y + h->chunk_fix_row[f], sizeof(double), cudaMemcpyDeviceToHost, h->copy_stream));
		CU(cudaStreamSynchronize(h->stream));
		CU(cudaStreamSynchronize(h->copy_stream));
		return 0;
	}
	if (b200spmv_spmv_device_ex(h, h->d_x, h->d_y, h->stream, nullptr))
		return 1;
	if (!(flags & B200SPMV_SKIP_Y_DOWNLOAD) && h->m > 0)
		CU(cudaMemcpyAsync(y, h->d_y, h->m * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
	CU(cudaStreamSynchronize(h->stream));
	return 0;
}

extern "C" int b200spmv_spmv_host(b200spmv_t * h, const double * x, double * y)
{
	return b200spmv_spmv_host_ex(h, x, y, 0);
}

// ---------------------------------------------------------------------------------------------
// statistics / layout export
// ---------------------------------------------------------------------------------------------
extern "C" int b200spmv_stats(const b200spmv_t * h, double * out, int32_t n_out)
{
	if (!h || !out)
		B200_FAIL("NULL argument");
	double v[13];
	if (h->npanels > 1) {
		// panel shell: footprints and launches add up, the rest describes panel 0
		double w[13];
		if (b200spmv_stats(h->panel[0], v, 13))
			return 1;
		for (int q = 1; q < h->npanels; q++) {
			if (b200spmv_stats(h->panel[q], w, 13))
				return 1;
			v[0] += w[0];
		}
		v[6] = (double) h->convert_ms;
		v[12] = (double) h->npanels;
		for (int i = 0; i < n_out && i < 13; i++)
			out[i] = v[i];
		return 0;
	}
	v[12] = 1.0;
	for (int i = 0; i < 3; i++)
		v[7 + i] = (double) h->trial_ms[i];
	v[10] = (double) h->compressed;
	v[11] = (double) h->n_empty;
	if (h->format == B200SPMV_CSR_ADAPTIVE) {
		v[0] = (double) h->nnz * 12.0 + (double) (h->m + 1) * 4.0 + (double) (h->ntiles + 1) * 8.0 + (double) h->ntiles * 16.0 + (double) h->nsplit * 8.0
		       + (h->compressed ? (double) (h->m_plan + 1) * 8.0 + (double) h->n_empty * 4.0 : 0.0);
		v[1] = (double) h->ntiles;
		v[2] = (double) h->nsplit;
		v[3] = (double) h->nnz;
		v[4] = (double) h->tile_nnz;
		v[5] = (double) h->ext;
	} else {
		v[0] = (double) h->padded * 12.0 + (double) h->nslices * 4.0 + (double) (h->nslices + 1) * 8.0 + (double) h->m * 4.0
		       + (double) h->sell_nitems * 16.0 + (double) h->sell_ncuts * 16.0 + (double) h->sell_nslots * 256.0;
		v[1] = (double) h->nslices;
		v[2] = (double) h->sell_ncuts;
		v[3] = (double) h->padded;
		v[4] = (double) h->C;
		v[5] = (double) h->sigma;
	}
	v[6] = (double) h->convert_ms;
	for (int i = 0; i < n_out && i < 13; i++)
		out[i] = v[i];
	return 0;
}

// Compressed row space: plan row k is original row rows_nz[k] (rows_nz[m_plan] = m).
static int export_map_rows(const b200spmv * h, int32_t * d, int64_t count)
{
	if (!h->compressed)
		return 0;
	int32_t * map = (int32_t *) malloc((size_t) (h->m_plan + 1) * sizeof(int32_t));
	if (!map)
		B200_FAIL("out of host memory");
	if (cudaMemcpy(map, h->d_rows_nz, (size_t) (h->m_plan + 1) * sizeof(int32_t), cudaMemcpyDeviceToHost) != cudaSuccess) {
		free(map);
		B200_FAIL("export: copy of the row map failed");
	}
	for (int64_t i = 0; i < count; i++)
		d[i] = map[d[i]];
	free(map);
	return 0;
}

extern "C" int b200spmv_export(const b200spmv_t * h, int32_t what, void * dst, int64_t capacity_bytes, int64_t * count_out)
{
	if (!h || !count_out)
		B200_FAIL("NULL argument");
	if (h->npanels > 1)
		B200_FAIL("export: the handle is a shell around %d column panels (create with col_panels = 1 to export a layout)", h->npanels);
	CU(cudaSetDevice(h->device));
	const void * src = nullptr;
	int64_t count = 0, esize = 4;
	const bool sell = (h->format == B200SPMV_SELL_C_SIGMA);
	switch (what) {
	case B200SPMV_X_ROW_ORDER:   if (!sell) B200_FAIL("not a SELL handle"); src = h->d_row_order; count = h->m; break;
	case B200SPMV_X_SLICE_WIDTH: if (!sell) B200_FAIL("not a SELL handle"); src = h->d_slice_width; count = h->nslices; break;
	case B200SPMV_X_SLICE_PTR:   if (!sell) B200_FAIL("not a SELL handle"); src = h->d_slice_ptr; count = h->nslices + 1; esize = 8; break;
	case B200SPMV_X_SELL_COL:    if (!sell) B200_FAIL("not a SELL handle"); src = h->d_scol; count = h->padded; break;
	case B200SPMV_X_SELL_VAL:    if (!sell) B200_FAIL("not a SELL handle"); src = h->d_sval; count = h->padded; esize = 8; break;
	case B200SPMV_X_TILE_ROW:    if (sell) B200_FAIL("not a CSR handle"); src = h->d_tile_row; count = h->ntiles + 1; break;
	case B200SPMV_X_SPLIT_ROW:   if (sell) B200_FAIL("not a CSR handle"); src = h->d_split_row; count = h->nsplit; break;
	case B200SPMV_X_SPLIT_TILE:  if (sell) B200_FAIL("not a CSR handle"); src = h->d_split_tile; count = h->nsplit; break;
	default: B200_FAIL("unknown export selector %d", what);
	}
	*count_out = count;
	if (!dst)
		return 0;                               // sizing call
	if (capacity_bytes < count * esize)
		B200_FAIL("export buffer too small: need %lld bytes", (long long) (count * esize));
	if (count == 0)
		return 0;
	if (what == B200SPMV_X_TILE_ROW) {
		// tile_info is {row | flag, row_ptr[row]} pairs: copy the first member, drop the flag bit
		CU(cudaMemcpy2D(dst, 4, src, 8, 4, (size_t) count, cudaMemcpyDeviceToHost));
		int32_t * d = (int32_t *) dst;
		for (int64_t i = 0; i < count; i++)
			d[i] &= 0x7fffffff;
		return export_map_rows(h, d, count);
	}
	CU(cudaMemcpy(dst, src, (size_t) (count * esize), cudaMemcpyDeviceToHost));
	if (what == B200SPMV_X_SPLIT_ROW)
		return export_map_rows(h, (int32_t *) dst, count);
	return 0;
}

// ---------------------------------------------------------------------------------------------
// device memory helpers (so that a C host program needs no CUDA runtime of its own)
// ---------------------------------------------------------------------------------------------
extern "C" int b200spmv_device_count(int32_t * count)
{
	int n = 0;
	cudaError_t e = cudaGetDeviceCount(&n);
	*count = (e == cudaSuccess) ? n : 0;
	if (e != cudaSuccess)
		B200_FAIL("cudaGetDeviceCount: %s", cudaGetErrorString(e));
	return 0;
}
extern "C" int b200spmv_set_device(int32_t device) { CU(cudaSetDevice(device)); return 0; }
extern "C" int b200spmv_malloc(void ** dptr, int64_t bytes) { CU(cudaMalloc(dptr, (size_t) (bytes > 0 ? bytes : 1))); return 0; }
extern "C" int b200spmv_free(void * dptr) { CU(cudaFree(dptr)); return 0; }
extern "C" int b200spmv_memcpy_h2d(void * dst, const void * src, int64_t bytes) { CU(cudaMemcpy(dst, src, (size_t) bytes, cudaMemcpyHostToDevice)); return 0; }
extern "C" int b200spmv_memcpy_d2h(void * dst, const void * src, int64_t bytes) { CU(cudaMemcpy(dst, src, (size_t) bytes, cudaMemcpyDeviceToHost)); return 0; }
extern "C" int b200spmv_malloc_host(void ** hptr, int64_t bytes) { CU(cudaMallocHost(hptr, (size_t) (bytes > 0 ? bytes : 1))); return 0; }
extern "C" int b200spmv_free_host(void * hptr) { CU(cudaFreeHost(hptr)); return 0; }
extern "C" int b200spmv_synchronize(void) { CU(cudaDeviceSynchronize()); return 0; }

extern "C" int b200spmv_ipc_export(void * dptr, unsigned char handle_out[64])
{
	static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
	cudaIpcMemHandle_t hd;
	CU(cudaIpcGetMemHandle(&hd, dptr));
	memcpy(handle_out, &hd, 64);
	return 0;
}
extern "C" int b200spmv_ipc_import(const unsigned char handle[64], void ** dptr_out)
{
	cudaIpcMemHandle_t hd;
	memcpy(&hd, handle, 64);
	CU(cudaIpcOpenMemHandle(dptr_out, hd, cudaIpcMemLazyEnablePeerAccess));
	return 0;
}
extern "C" int b200spmv_ipc_close(void * dptr) { CU(cudaIpcCloseMemHandle(dptr)); return 0; }

// ---------------------------------------------------------------------------------------------
// device timing of back-to-back products (CUDA events on the launching stream)
// ---------------------------------------------------------------------------------------------
__global__ void flush_kernel(double * buf, int64_t n, double v)
{
	for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x)
		buf[i] = v;
}

extern "C" int b200spmv_time_device(b200spmv_t * h, const double * dx, double * dy, int32_t iters, int64_t flush_bytes, float * ms_out)
{
	if (!h || !dx || !dy || !ms_out || iters < 1)
		B200_FAIL("bad arguments");
	CU(cudaSetDevice(h->device));
	double * d_flush = nullptr;
	if (flush_bytes > 0)
		CU(cudaMalloc(&d_flush, (size_t) flush_bytes));
	cudaEvent_t * ev = (cudaEvent_t *) malloc(2 * (size_t) iters * sizeof(cudaEvent_t));
	for (int i = 0; i < 2 * iters; i++)
		cudaEventCreate(&ev[i]);
	int rc = 0;
	for (int i = 0; i < iters && !rc; i++) {
		if (d_flush)
			flush_kernel<<<148 * 8, 256, 0, h->stream>>>(d_flush, flush_bytes / 8, (double) i);
		cudaEventRecord(ev[2 * i], h->stream);
		rc = b200spmv_spmv_device_ex(h, dx, dy, h->stream, nullptr);
		cudaEventRecord(ev[2 * i + 1], h->stream);
	}
	cudaError_t e = cudaStreamSynchronize(h->stream);
	if (!rc && e == cudaSuccess)
		for (int i = 0; i < iters; i++)
			cudaEventElapsedTime(&ms_out[i], ev[2 * i], ev[2 * i + 1]);
	for (int i = 0; i < 2 * iters; i++)
		cudaEventDestroy(ev[i]);
	free(ev);
	if (d_flush)
		cudaFree(d_flush);
	if (rc)
		return rc;
	if (e != cudaSuccess)
		B200_FAIL("b200spmv_time_device: %s", cudaGetErrorString(e));
	return 0;
}

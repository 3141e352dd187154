// panels.cu -- column panels for matrices whose referenced band of x does not fit the L2.
//
// Measured (profiles/r01o_ncu_full_c5shard_csr_key_metrics.txt): a config-5 row shard gathers from an 80 MB band of x; the
// L2 (two 63 MB halves) cannot hold it, the kernel reads 5.5 GB from DRAM for 3.4 GB of compulsory bytes and runs at
// 1.04 ms, against 0.81 ms for the same shard shape with a 10 MB band.  Splitting the columns into P equal ranges
// and running one product per range (y += A_p x, ascending p) keeps every range's x resident; the price is one
// read-modify-write of y per extra panel (16 bytes per row) and one more launch.
//
// The split happens once, on the GPU, at create time: per row and panel a count, a scan per panel, a scatter that keeps
// the order of a row's entries; each panel becomes an ordinary sub-handle (CSR / SELL-C-sigma / AUTO, picked per panel).
#include "common.cuh"

constexpr int MAX_PANELS = 8;

struct PanelBounds { int32_t c[MAX_PANELS + 1]; };

__device__ __forceinline__ int panel_of(const PanelBounds & b, int P, int col)
{
	int p = 0;
	#pragma unroll
	for (int q = 1; q < MAX_PANELS; q++)
		p += (q < P && col >= b.c[q]) ? 1 : 0;
	return p;
}

// Rows longer than this are split by a whole warp (panel_*_long_kernel): a thread-per-row loop over one of the generator's
// giant skew rows (16 million entries in the mawi twin's variants) took 4 s at create time (profiles/r02b_config3_*).
constexpr int PANEL_LONG_ROW = 64;

// cnt[p * (m + 1) + i] = entries of row i in panel p (thread per row; long rows are listed for the warp-per-row pass)
__global__ void panel_count_kernel(const int32_t * __restrict__ row_ptr, const int32_t * __restrict__ col, int64_t m, int P, PanelBounds b,
                                   int64_t * __restrict__ cnt, int32_t * __restrict__ long_rows, unsigned long long * __restrict__ n_long)
{
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i > m)
		return;
	int c[MAX_PANELS] = { 0 };
	if (i < m) {
		const int64_t a = row_ptr[i], e = row_ptr[i + 1];
		if (e - a > PANEL_LONG_ROW) {
			long_rows[atomicAdd(n_long, 1ULL)] = (int32_t) i;
			return;                                   // its counts are written by panel_count_long_kernel
		}
		for (int64_t j = a; j < e; j++) {
			const int p = panel_of(b, P, col[j]);
			#pragma unroll
			for (int q = 0; q < MAX_PANELS; q++)
				c[q] += (q == p) ? 1 : 0;
		}
	}
	#pragma unroll
	for (int q = 0; q < MAX_PANELS; q++)
		if (q < P)
			cnt[(int64_t) q * (m + 1) + i] = c[q];
}

// one warp per long row
__global__ void panel_count_long_kernel(const int32_t * __restrict__ row_ptr, const int32_t * __restrict__ col, int64_t m, int P, PanelBounds b,
                                        int64_t * __restrict__ cnt, const int32_t * __restrict__ long_rows, int64_t n_long)
{
	const int64_t w = (int64_t) blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int lane = threadIdx.x & 31;
	if (w >= n_long)
		return;
	const int64_t i = long_rows[w];
	int c[MAX_PANELS] = { 0 };
	for (int64_t j = (int64_t) row_ptr[i] + lane; j < row_ptr[i + 1]; j += 32) {
		const int p = panel_of(b, P, col[j]);
		#pragma unroll
		for (int q = 0; q < MAX_PANELS; q++)
			c[q] += (q == p) ? 1 : 0;
	}
	#pragma unroll
	for (int q = 0; q < MAX_PANELS; q++) {
		int v = c[q];
		for (int o = 16; o > 0; o >>= 1)
			v += __shfl_xor_sync(0xffffffffu, v, o);
		if (lane == 0 && q < P)
			cnt[(int64_t) q * (m + 1) + i] = v;
	}
}

// after the per-panel exclusive scans: rp[p][i] = off[p][i]; entries scattered in row order
__global__ void panel_fill_kernel(const int32_t * __restrict__ row_ptr, const int32_t * __restrict__ col, const double * __restrict__ val,
                                  int64_t m, int P, PanelBounds b, const int64_t * __restrict__ off, const int64_t * __restrict__ base,
                                  int32_t * __restrict__ prp, int32_t * __restrict__ pcol, double * __restrict__ pval)
{
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i > m)
		return;
	int64_t o[MAX_PANELS];
	#pragma unroll
	for (int q = 0; q < MAX_PANELS; q++)
		if (q < P) {
			o[q] = off[(int64_t) q * (m + 1) + i];
			prp[(int64_t) q * (m + 1) + i] = (int32_t) o[q];
			o[q] += base[q];
		}
	if (i == m || row_ptr[i + 1] - row_ptr[i] > PANEL_LONG_ROW)
		return;                                           // long rows: panel_fill_long_kernel
	for (int64_t j = row_ptr[i]; j < row_ptr[i + 1]; j++) {
		const int cj = col[j];
		const int p = panel_of(b, P, cj);
		int64_t dst = 0;
		#pragma unroll
		for (int q = 0; q < MAX_PANELS; q++)
			if (q == p)
				dst = o[q]++;
		pcol[dst] = cj;
		pval[dst] = val[j];
	}
}

// one warp per long row: 32 entries at a time, each entry's slot = entries of its panel so far in the row (ballot prefix),
// so a row's entries keep their order inside every panel
__global__ void panel_fill_long_kernel(const int32_t * __restrict__ row_ptr, const int32_t * __restrict__ col, const double * __restrict__ val,
                                       int64_t m, int P, PanelBounds b, const int64_t * __restrict__ off, const int64_t * __restrict__ base,
                                       int32_t * __restrict__ pcol, double * __restrict__ pval, const int32_t * __restrict__ long_rows, int64_t n_long)
{
	const int64_t w = (int64_t) blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int lane = threadIdx.x & 31;
	if (w >= n_long)
		return;
	const int64_t i = long_rows[w];
	int64_t o[MAX_PANELS];
	#pragma unroll
	for (int q = 0; q < MAX_PANELS; q++)
		o[q] = (q < P) ? off[(int64_t) q * (m + 1) + i] + base[q] : 0;
	const int64_t a = row_ptr[i], e = row_ptr[i + 1];
	for (int64_t j0 = a; j0 < e; j0 += 32) {
		const int64_t j = j0 + lane;
		const bool live = j < e;
		const int cj = live ? col[j] : 0;
		const int p = live ? panel_of(b, P, cj) : -1;
		int64_t dst = 0;
		#pragma unroll
		for (int q = 0; q < MAX_PANELS; q++) {
			const unsigned mask = __ballot_sync(0xffffffffu, p == q);
			if (p == q)
				dst = o[q] + __popc(mask & ((1u << lane) - 1u));
			o[q] += __popc(mask);
		}
		if (live) {
			pcol[dst] = cj;
			pval[dst] = val[j];
		}
	}
}

// 0 = auto.  Measured (profiles/r01p_sweep_col_panels.txt, profiles/r01r_panels_borderline.txt): two panels win by 14-21 % on an
// 80 MB band at 60 nnz / row and by 18 % (CSR) on a 137 MB band at 5 nnz / row, but LOSE 7-24 % on 67-84 MB bands at 5-10 nnz /
// row (the second launch re-reads and re-writes y: 16 bytes per row against 60-120 bytes of matrix per row); three or four
// panels are always slower than two.  Hence: long rows split above 72 MB, short rows only when the band is far beyond L2.
int panels_auto_count(const b200spmv * h, int requested)
{
	if (requested == 1 || h->nnz == 0 || h->m == 0)
		return 1;
	if (requested > 1)
		return requested > MAX_PANELS ? MAX_PANELS : requested;
	const double band = 8.0 * (double) (h->col_hi - h->col_lo), MB = 1048576.0, avg = (double) h->nnz / (double) h->m;
	if (band <= 72.0 * MB || (avg < 24.0 && band <= 112.0 * MB))
		return 1;
	int P = (int) ((band + 88.0 * MB - 1.0) / (88.0 * MB));
	return P > MAX_PANELS ? MAX_PANELS : (P < 2 ? 2 : P);
}

// Splits h's CSR arrays (device) into P column panels and creates one sub-handle per panel.
int panels_split(b200spmv * h, int P, const b200spmv_opts * sub_opts)
{
	B200Range nvtx_range_("b200spmv panels_split");
	const int64_t m = h->m;
	PanelBounds b;
	const int64_t lo = h->col_lo, span = h->col_hi - h->col_lo;
	for (int q = 0; q <= MAX_PANELS; q++)
		b.c[q] = (int32_t) (q >= P ? h->col_hi : lo + span * q / P);
	int64_t * d_cnt = nullptr, * d_base = nullptr;
	int32_t * d_long = nullptr;
	unsigned long long * d_nlong = nullptr, n_long = 0;
	int32_t * d_prp = nullptr, * d_pcol = nullptr;
	double * d_pval = nullptr;
	int rc = 1;
	int64_t tot[MAX_PANELS], base[MAX_PANELS];
	const unsigned blocks = (unsigned) ((m + 1 + 255) / 256);
	do {
		if (cudaMalloc(&d_cnt, (size_t) P * (m + 1) * sizeof(int64_t)) != cudaSuccess) break;
		if (cudaMalloc(&d_base, MAX_PANELS * sizeof(int64_t)) != cudaSuccess) break;
		if (cudaMalloc(&d_prp, (size_t) P * (m + 1) * sizeof(int32_t)) != cudaSuccess) break;
		if (cudaMalloc(&d_pcol, (size_t) (h->nnz > 0 ? h->nnz : 1) * sizeof(int32_t)) != cudaSuccess) break;
		if (cudaMalloc(&d_pval, (size_t) (h->nnz > 0 ? h->nnz : 1) * sizeof(double)) != cudaSuccess) break;
		if (cudaMalloc(&d_long, (size_t) (m > 0 ? m : 1) * sizeof(int32_t)) != cudaSuccess) break;
		if (cudaMalloc(&d_nlong, sizeof(unsigned long long)) != cudaSuccess) break;
		if (cudaMemsetAsync(d_nlong, 0, sizeof(unsigned long long), h->stream) != cudaSuccess) break;
		panel_count_kernel<<<blocks, 256, 0, h->stream>>>(h->d_row_ptr, h->d_col, m, P, b, d_cnt, d_long, d_nlong);
		if (cudaPeekAtLastError() != cudaSuccess) break;
		if (cudaMemcpyAsync(&n_long, d_nlong, sizeof(n_long), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess) break;
		if (cudaStreamSynchronize(h->stream) != cudaSuccess) break;
		if (n_long > 0) {
			panel_count_long_kernel<<<(unsigned) ((n_long + 7) / 8), 256, 0, h->stream>>>(h->d_row_ptr, h->d_col, m, P, b, d_cnt, d_long, (int64_t) n_long);
			if (cudaPeekAtLastError() != cudaSuccess) break;
		}
		bool ok = true;
		int64_t run = 0;
		for (int q = 0; q < P && ok; q++) {
			ok = device_exclusive_scan_i64(d_cnt + (int64_t) q * (m + 1), m + 1, h->stream) == 0 &&
			     cudaMemcpyAsync(&tot[q], d_cnt + (int64_t) q * (m + 1) + m, sizeof(int64_t), cudaMemcpyDeviceToHost, h->stream) == cudaSuccess &&
			     cudaStreamSynchronize(h->stream) == cudaSuccess;
			base[q] = run;
			run += tot[q];
		}
		if (!ok || run != h->nnz) break;
		if (cudaMemcpyAsync(d_base, base, P * sizeof(int64_t), cudaMemcpyHostToDevice, h->stream) != cudaSuccess) break;
		panel_fill_kernel<<<blocks, 256, 0, h->stream>>>(h->d_row_ptr, h->d_col, h->d_val, m, P, b, d_cnt, d_base, d_prp, d_pcol, d_pval);
		if (cudaPeekAtLastError() != cudaSuccess) break;
		if (n_long > 0) {
			panel_fill_long_kernel<<<(unsigned) ((n_long + 7) / 8), 256, 0, h->stream>>>(h->d_row_ptr, h->d_col, h->d_val, m, P, b, d_cnt, d_base, d_pcol,
			                                                                         d_pval, d_long, (int64_t) n_long);
			if (cudaPeekAtLastError() != cudaSuccess) break;
		}
		if (cudaStreamSynchronize(h->stream) != cudaSuccess) break;
		rc = 0;
	} while (0);
	if (rc)
		snprintf(g_b200_err, sizeof(g_b200_err), "column panels: CUDA failure: %s", cudaGetErrorString(cudaGetLastError()));
	for (int q = 0; q < P && !rc; q++) {
		rc = create_from_device(&h->panel[q], d_prp + (int64_t) q * (m + 1), d_pcol + base[q], d_pval + base[q], m, h->n, tot[q], sub_opts);
		if (!rc) {
			h->npanels = q + 1;
			// a panel never runs the host-vector entry point (the shell stages x / y): drop what an AUTO trial left behind
			b200spmv * s = h->panel[q];
			if (s->d_x)
				cudaFree(s->d_x);
			if (s->d_y)
				cudaFree(s->d_y);
			s->d_x = nullptr;
			s->d_y = nullptr;
		}
	}
	cudaFree(d_cnt);
	cudaFree(d_long);
	cudaFree(d_nlong);
	cudaFree(d_base);
	cudaFree(d_prp);
	cudaFree(d_pcol);
	cudaFree(d_pval);
	if (rc) {
		for (int q = 0; q < h->npanels; q++) {
			destroy_handle(h->panel[q]);
			h->panel[q] = nullptr;
		}
		h->npanels = 0;
	}
	return rc;
}

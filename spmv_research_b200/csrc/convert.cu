// convert.cu -- on-GPU format construction: CSR tile plan, SELL-C-sigma conversion, scans.
//
// Integer outputs are bit-exact with the reference:
//   SELL-C-sigma : K/sell-C-s/RISC-V/radix_sort.c:36-122 (stable descending sort of row lengths
//                  inside sigma windows), sellcs_format.c:137-200 (slice widths / pointers,
//                  column-major padded slices, zero padding).
#include "common.cuh"
#ifndef CSR_EXP_PERM
#define CSR_EXP_PERM 0
#endif
#ifndef CSR_EXP_ROT
#define CSR_EXP_ROT 0
#endif
#if CSR_EXP_PERM || CSR_EXP_ROT
#include "csr_perm.h"
#endif

// ---------------------------------------------------------------------------------------------
// exclusive scan of int64, in place (three-phase, recursive over block sums)
// ---------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_BLOCK = SCAN_THREADS * SCAN_ITEMS;

__global__ void scan_block_kernel(int64_t * data, int64_t n, int64_t * block_sums)
{
	__shared__ int64_t s_warp[SCAN_THREADS / 32];
	const int64_t base = (int64_t) blockIdx.x * SCAN_BLOCK + (int64_t) threadIdx.x * SCAN_ITEMS;
	int64_t v[SCAN_ITEMS], tsum = 0;
	#pragma unroll
	for (int i = 0; i < SCAN_ITEMS; i++) {
		v[i] = (base + i < n) ? data[base + i] : 0;
		tsum += v[i];
	}
	// inclusive scan of thread sums across the warp, then across warps
	int64_t inc = tsum;
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		int64_t t = __shfl_up_sync(0xffffffffu, inc, o);
		if (lane >= o)
			inc += t;
	}
	if (lane == 31)
		s_warp[wid] = inc;
	__syncthreads();
	int64_t woff = 0;
	for (int w = 0; w < wid; w++)
		woff += s_warp[w];
	int64_t run = woff + inc - tsum;
	#pragma unroll
	for (int i = 0; i < SCAN_ITEMS; i++) {
		if (base + i < n)
			data[base + i] = run;
		run += v[i];
	}
	if (threadIdx.x == SCAN_THREADS - 1 && block_sums)
		block_sums[blockIdx.x] = run;
}

__global__ void scan_add_kernel(int64_t * data, int64_t n, const int64_t * block_offsets)
{
	const int64_t off = block_offsets[blockIdx.x];
	const int64_t base = (int64_t) blockIdx.x * SCAN_BLOCK;
	for (int i = threadIdx.x; i < SCAN_BLOCK; i += SCAN_THREADS)
		if (base + i < n)
			data[base + i] += off;
}

int device_exclusive_scan_i64(int64_t * d_data, int64_t n, cudaStream_t st)
{
	if (n <= 0)
		return 0;
	const int64_t blocks = (n + SCAN_BLOCK - 1) / SCAN_BLOCK;
	if (blocks == 1) {
		scan_block_kernel<<<1, SCAN_THREADS, 0, st>>>(d_data, n, nullptr);
		CU_LAUNCH();
		return 0;
	}
	int64_t * d_sums;
	CU(cudaMalloc(&d_sums, blocks * sizeof(int64_t)));
	scan_block_kernel<<<(unsigned) blocks, SCAN_THREADS, 0, st>>>(d_data, n, d_sums);
	CU_LAUNCH();
	int rc = device_exclusive_scan_i64(d_sums, blocks, st);
	if (!rc) {
		scan_add_kernel<<<(unsigned) blocks, SCAN_THREADS, 0, st>>>(d_data, n, d_sums);
		if (cudaPeekAtLastError() != cudaSuccess)
			rc = 1;
	}
	cudaStreamSynchronize(st);
	cudaFree(d_sums);
	if (rc)
		B200_FAIL("scan failed: %s", cudaGetErrorString(cudaGetLastError()));
	return 0;
}

// ---------------------------------------------------------------------------------------------
// CSR tile plan
// ---------------------------------------------------------------------------------------------
// tile_info[t] = { first row whose start offset is >= t*T  (bit 31: a split row enters the tile),
//                  row_ptr[that row] };  tile_info[ntiles] = { m, nnz }.
__device__ __forceinline__ int64_t lower_bound_rows(const int32_t * row_ptr, int64_t m, int64_t target)
{
	int64_t lo = 0, hi = m;          // first i in [0, m] with row_ptr[i] >= target (row_ptr[m] = nnz)
	while (lo < hi) {
		const int64_t mid = (lo + hi) >> 1;
		if (row_ptr[mid] < target)
			lo = mid + 1;
		else
			hi = mid;
	}
	return lo;
}

__global__ void tile_info_kernel(const int32_t * __restrict__ row_ptr, int64_t m, int64_t nnz, int64_t ntiles, int T, int ext,
                                 int2 * __restrict__ tile_info, int64_t * __restrict__ split_flag)
{
	const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (t > ntiles)
		return;
	if (t == ntiles) {
		tile_info[t] = make_int2((int) m, (int) nnz);
		return;
	}
	const int64_t base = t * (int64_t) T;
	const int64_t R0 = (t == 0) ? 0 : lower_bound_rows(row_ptr, m, base);
	const int64_t start0 = row_ptr[R0];
	int flag = 0;
	if (R0 > 0 && start0 > base) {
		// row R0-1 started before this window and reaches into it: split iff it leaves its
		// owner's window by more than ext
		const int64_t start_h = row_ptr[R0 - 1];
		const int64_t k = start_h / T;
		if (start0 - (k + 1) * (int64_t) T > ext)
			flag = 1;
	}
	tile_info[t] = make_int2((int) R0 | (flag ? (int) 0x80000000 : 0), (int) start0);
	// does this tile OWN a split row?  (its last owned row leaves the window by more than ext)
	const int64_t R1 = (t + 1 == ntiles) ? m : lower_bound_rows(row_ptr, m, base + T);
	const int64_t end_own = row_ptr[R1];
	split_flag[t] = (R1 > R0 && end_own - (base + T) > ext) ? 1 : 0;
}

__global__ void split_compact_kernel(const int2 * __restrict__ tile_info, const int64_t * __restrict__ split_pos, int64_t ntiles,
                                     int64_t nsplit, int32_t * __restrict__ split_row, int32_t * __restrict__ split_tile)
{
	const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= ntiles)
		return;
	const int64_t pos = split_pos[t];
	const int64_t nxt = (t + 1 < ntiles) ? split_pos[t + 1] : nsplit;
	if (nxt > pos) {
		split_row[pos] = (tile_info[t + 1].x & 0x7fffffff) - 1;
		split_tile[pos] = (int32_t) t;
	}
}

// What csr_fold_split_rows needs: per tile the split row it owns / is entered by, per split row its owner tile and the number
// of tiles that contribute a partial (owner + every tile up to the one holding the row's last entry).
__global__ void split_fold_kernel(const int2 * __restrict__ tile_info, const int64_t * __restrict__ split_pos, int64_t ntiles, int64_t nsplit,
                                  const int32_t * __restrict__ row_ptr, int T, int2 * __restrict__ tile_split, int4 * __restrict__ split_info)
{
	const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= ntiles)
		return;
	const int64_t pos = split_pos[t];
	const int64_t nxt = (t + 1 < ntiles) ? split_pos[t + 1] : nsplit;
	const int own = (nxt > pos) ? (int) pos : -1;
	const int head = (tile_info[t].x < 0) ? (int) pos - 1 : -1;      // entered by the split row of the latest owner before t
	tile_split[t] = make_int2(own, head);
	if (own >= 0) {
		const int row = (tile_info[t + 1].x & 0x7fffffff) - 1;
		const int64_t last = ((int64_t) row_ptr[row + 1] - 1) / T;
		split_info[own] = make_int4(row, (int) t, (int) (last - t + 1), 0);
	}
}

// ---- row compression ---------------------------------------------------------------------------
__global__ void nonempty_flag_kernel(const int32_t * __restrict__ row_ptr, int64_t m, int64_t * __restrict__ flag)
{
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i <= m)
		flag[i] = (i < m && row_ptr[i + 1] > row_ptr[i]) ? 1 : 0;
}

__global__ void compress_rows_kernel(const int32_t * __restrict__ row_ptr, const int64_t * __restrict__ pos, int64_t m, int64_t m_nz,
                                     int32_t * __restrict__ rp_nz, int32_t * __restrict__ rows_nz, int32_t * __restrict__ empty_rows)
{
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i > m)
		return;
	if (i == m) {
		rp_nz[m_nz] = row_ptr[m];
		rows_nz[m_nz] = (int32_t) m;
		return;
	}
	const int64_t k = pos[i];
	if (row_ptr[i + 1] > row_ptr[i]) {
		rp_nz[k] = row_ptr[i];
		rows_nz[k] = (int32_t) i;
	} else {
		empty_rows[i - k] = (int32_t) i;
	}
}

__global__ void max_owned_rows_kernel(const int2 * __restrict__ tile_info, int64_t ntiles, int * __restrict__ out)
{
	const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= ntiles)
		return;
	const int own = (tile_info[t + 1].x & 0x7fffffff) - (tile_info[t].x & 0x7fffffff);
	atomicMax(out, own);
}

static void free_tile_plan(b200spmv * h)
{
	void ** ptrs[] = { (void **) &h->d_tile_row, (void **) &h->d_split_row, (void **) &h->d_split_tile, (void **) &h->d_tile_head,
	                   (void **) &h->d_tile_tail, (void **) &h->d_perm, (void **) &h->d_perm_off, (void **) &h->d_rot,
	                   (void **) &h->d_tile_split, (void **) &h->d_split_info, (void **) &h->d_split_cnt };
	for (void ** q : ptrs) {
		if (*q)
			cudaFree(*q);
		*q = nullptr;
	}
}

static int build_compressed_rows(b200spmv * h)
{
	const int64_t m = h->m;
	int64_t * d_pos;
	CU(cudaMalloc(&d_pos, (m + 1) * sizeof(int64_t)));
	nonempty_flag_kernel<<<(unsigned) ((m + 1 + 255) / 256), 256, 0, h->stream>>>(h->d_row_ptr, m, d_pos);
	CU_LAUNCH();
	if (device_exclusive_scan_i64(d_pos, m + 1, h->stream)) {
		cudaFree(d_pos);
		return 1;
	}
	int64_t m_nz = 0;
	CU(cudaMemcpyAsync(&m_nz, d_pos + m, sizeof(int64_t), cudaMemcpyDeviceToHost, h->stream));
	CU(cudaStreamSynchronize(h->stream));
	h->m_plan = m_nz;
	h->n_empty = m - m_nz;
	CU(cudaMalloc(&h->d_rp_nz, (m_nz + 1) * sizeof(int32_t)));
	CU(cudaMalloc(&h->d_rows_nz, (m_nz + 1) * sizeof(int32_t)));
	CU(cudaMalloc(&h->d_empty_rows, (h->n_empty > 0 ? h->n_empty : 1) * sizeof(int32_t)));
	compress_rows_kernel<<<(unsigned) ((m + 1 + 255) / 256), 256, 0, h->stream>>>(h->d_row_ptr, d_pos, m, m_nz, h->d_rp_nz, h->d_rows_nz,
	                                                                              h->d_empty_rows);
	CU_LAUNCH();
	CU(cudaStreamSynchronize(h->stream));
	CU(cudaFree(d_pos));
	h->d_rp_plan = h->d_rp_nz;
	h->compressed = 1;
	return 0;
}

static int plan_tiles_on(b200spmv * h);

// Plans the tiles; switches to the compressed row space when empty rows would otherwise pile up on single tiles
// (a tile owns every row that starts in its window, so a run of 200 000 empty rows is 200 000 rows for ONE CTA:
// measured 0.70 ms -> see profiles/, the mawi twin of config 3).
int csr_plan_tiles(b200spmv * h)
{
	B200Range nvtx_range_("b200spmv csr_plan_tiles");
	h->d_rp_plan = h->d_row_ptr;
	h->m_plan = h->m;
	h->compressed = 0;
	h->n_empty = 0;
	if (h->compress_opt == 2)
		return build_compressed_rows(h) || plan_tiles_on(h);
	if (plan_tiles_on(h))
		return 1;
	if (h->compress_opt == 1 || h->m == 0)
		return 0;
	int * d_max;
	int maxown = 0;
	CU(cudaMalloc(&d_max, sizeof(int)));
	CU(cudaMemsetAsync(d_max, 0, sizeof(int), h->stream));
	max_owned_rows_kernel<<<(unsigned) ((h->ntiles + 255) / 256), 256, 0, h->stream>>>(reinterpret_cast<const int2 *>(h->d_tile_row), h->ntiles, d_max);
	CU_LAUNCH();
	CU(cudaMemcpyAsync(&maxown, d_max, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
	CU(cudaStreamSynchronize(h->stream));
	CU(cudaFree(d_max));
	// a window of T non-zeros holds at most T non-empty rows: more owned rows than 2 T means >= T empty ones on one CTA
	if (maxown <= 2 * h->tile_nnz)
		return 0;
	free_tile_plan(h);
	return build_compressed_rows(h) || plan_tiles_on(h);
}

static int plan_tiles_on(b200spmv * h)
{
	const int T = h->tile_nnz;
	h->ntiles = (h->nnz + T - 1) / T;
	if (h->ntiles < 1)
		h->ntiles = 1;                       // an all-empty matrix still needs its rows zeroed
	if (h->ntiles >= 0x7fffffffLL)
		B200_FAIL("too many tiles");
	CU(cudaMalloc(&h->d_tile_row, (h->ntiles + 1) * sizeof(int2)));
	CU(cudaMalloc(&h->d_tile_head, h->ntiles * sizeof(double)));
	CU(cudaMalloc(&h->d_tile_tail, h->ntiles * sizeof(double)));
	CU(cudaMemsetAsync(h->d_tile_head, 0, h->ntiles * sizeof(double), h->stream));
	CU(cudaMemsetAsync(h->d_tile_tail, 0, h->ntiles * sizeof(double), h->stream));
	int64_t * d_flag;
	CU(cudaMalloc(&d_flag, h->ntiles * sizeof(int64_t)));
	const unsigned blocks = (unsigned) ((h->ntiles + 1 + 255) / 256);
	tile_info_kernel<<<blocks, 256, 0, h->stream>>>(h->d_rp_plan, h->m_plan, h->nnz, h->ntiles, T, h->ext,
	                                                reinterpret_cast<int2 *>(h->d_tile_row), d_flag);
	CU_LAUNCH();
	// number of split rows = last flag + last exclusive-scan value
	int64_t last_flag = 0, last_pos = 0;
	CU(cudaMemcpyAsync(&last_flag, d_flag + h->ntiles - 1, sizeof(int64_t), cudaMemcpyDeviceToHost, h->stream));
	CU(cudaStreamSynchronize(h->stream));
	if (device_exclusive_scan_i64(d_flag, h->ntiles, h->stream)) {
		cudaFree(d_flag);
		return 1;
	}
	CU(cudaMemcpyAsync(&last_pos, d_flag + h->ntiles - 1, sizeof(int64_t), cudaMemcpyDeviceToHost, h->stream));
	CU(cudaStreamSynchronize(h->stream));
	h->nsplit = last_flag + last_pos;
	h->fold_splits = 0;
	if (h->nsplit > 0) {
		CU(cudaMalloc(&h->d_split_row, h->nsplit * sizeof(int32_t)));
		CU(cudaMalloc(&h->d_split_tile, h->nsplit * sizeof(int32_t)));
		split_compact_kernel<<<(unsigned) ((h->ntiles + 255) / 256), 256, 0, h->stream>>>(
			reinterpret_cast<const int2 *>(h->d_tile_row), d_flag, h->ntiles, h->nsplit, h->d_split_row, h->d_split_tile);
		CU_LAUNCH();
		// Few split rows (the generator's skew rows): finished inside the tile kernel by the last-arriving CTA.  Many (long-row
		// matrices: at 500 per row almost every tile is entered and left by one) keep the separate fix-up launch -- measured:
		// the fences, the counter and the dependent look-ups on EVERY tile cost the 500-per-row sample matrix 19 % (0.165 ->
		// 0.197 ms, profiles/r02f_bench_n1.json), a 4.5 us launch costs it 3 %.
		h->fold_splits = (h->nsplit * 16 <= h->ntiles) ? 1 : 0;
		CU(cudaMalloc(&h->d_tile_split, h->ntiles * sizeof(int2)));
		CU(cudaMalloc(&h->d_split_info, h->nsplit * sizeof(int4)));
		CU(cudaMalloc(&h->d_split_cnt, h->nsplit * sizeof(unsigned int)));
		CU(cudaMemsetAsync(h->d_split_cnt, 0, h->nsplit * sizeof(unsigned int), h->stream));
		split_fold_kernel<<<(unsigned) ((h->ntiles + 255) / 256), 256, 0, h->stream>>>(
			reinterpret_cast<const int2 *>(h->d_tile_row), d_flag, h->ntiles, h->nsplit, h->d_rp_plan, T,
			reinterpret_cast<int2 *>(h->d_tile_split), reinterpret_cast<int4 *>(h->d_split_info));
		CU_LAUNCH();
	}
	CU(cudaStreamSynchronize(h->stream));
	CU(cudaFree(d_flag));
#if CSR_EXP_PERM
	{
		// experiment: row -> thread table of the thread-per-row bin, built on the host (csr_perm.h)
		std::vector<int32_t> rp((size_t) h->m_plan + 1), ti((size_t) 2 * (h->ntiles + 1));
		CU(cudaMemcpy(rp.data(), h->d_rp_plan, rp.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
		CU(cudaMemcpy(ti.data(), h->d_tile_row, ti.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
		std::vector<int64_t> off;
		std::vector<uint16_t> perm;
		csr_perm_build(rp.data(), ti.data(), h->ntiles, T, h->ext, off, perm, nullptr, CSR_EXP_ROT != 0);
		CU(cudaMalloc(&h->d_perm_off, off.size() * sizeof(int64_t)));
		CU(cudaMalloc(&h->d_perm, (perm.size() + 1) * sizeof(uint16_t)));
		CU(cudaMemcpy(h->d_perm_off, off.data(), off.size() * sizeof(int64_t), cudaMemcpyHostToDevice));
		if (!perm.empty())
			CU(cudaMemcpy(h->d_perm, perm.data(), perm.size() * sizeof(uint16_t), cudaMemcpyHostToDevice));
	}
#endif
#if CSR_EXP_ROT
	{
		// experiment: start rotation of every row sum of the thread-per-row bin, built on the host (csr_perm.h)
		std::vector<int32_t> rp((size_t) h->m_plan + 1), ti((size_t) 2 * (h->ntiles + 1));
		CU(cudaMemcpy(rp.data(), h->d_rp_plan, rp.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
		CU(cudaMemcpy(ti.data(), h->d_tile_row, ti.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
		std::vector<uint8_t> rot;
		csr_rot_build(rp.data(), ti.data(), h->ntiles, h->m_plan, T, h->ext, rot);
		CU(cudaMalloc(&h->d_rot, rot.size()));
		CU(cudaMemcpy(h->d_rot, rot.data(), rot.size(), cudaMemcpyHostToDevice));
	}
#endif
	return 0;
}

// ---------------------------------------------------------------------------------------------
// SELL-C-sigma conversion
// ---------------------------------------------------------------------------------------------
// The reference's LSD radix passes (radix_sort.c:36-101) are a stable sort by descending row
// length inside each sigma window.  Sorting the unique 64-bit key ((2^32-1 - len) << 32 | row)
// ascending yields exactly that order with any comparison network, so a bitonic network is used:
// windows are padded to a power of two with all-ones keys, which sort to the end.
constexpr int BITONIC_BLOCK = 4096;                  // keys per CTA in the shared-memory phases
constexpr int BITONIC_THREADS = 512;

__global__ void sell_keys_kernel(const int32_t * __restrict__ row_ptr, int64_t m, int64_t sigma, int64_t sigma_p, int64_t nslots,
                                 uint64_t * __restrict__ keys)
{
	const int64_t slot = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (slot >= nslots)
		return;
	const int64_t w = slot / sigma_p, r = slot % sigma_p;
	const int64_t row = w * sigma + r;
	uint64_t key = ~0ULL;
	if (r < sigma && row < m) {
		const uint32_t len = (uint32_t) (row_ptr[row + 1] - row_ptr[row]);
		key = ((uint64_t) (0xFFFFFFFFu - len) << 32) | (uint64_t) (uint32_t) row;
	}
	keys[slot] = key;
}

__device__ __forceinline__ void cmpswap(uint64_t & a, uint64_t & b, bool ascending)
{
	if ((a > b) == ascending) {
		const uint64_t t = a;
		a = b;
		b = t;
	}
}

// Shared-memory phases over blocks of BITONIC_BLOCK keys.  k_first == 0: all stages k = 2 ..
// min(BITONIC_BLOCK, sigma_p).  Otherwise: the tail j = BITONIC_BLOCK/2 .. 1 of stage k_first.
__global__ void __launch_bounds__(BITONIC_THREADS)
bitonic_smem_kernel(uint64_t * __restrict__ keys, int64_t sigma_p, int64_t k_first)
{
	__shared__ uint64_t s[BITONIC_BLOCK];
	const int64_t base = (int64_t) blockIdx.x * BITONIC_BLOCK;
	for (int i = threadIdx.x; i < BITONIC_BLOCK; i += BITONIC_THREADS)
		s[i] = keys[base + i];
	__syncthreads();
	const int64_t k_lo = k_first ? k_first : 2;
	const int64_t k_hi = k_first ? k_first : (sigma_p < BITONIC_BLOCK ? sigma_p : BITONIC_BLOCK);
	for (int64_t k = k_lo; k <= k_hi; k <<= 1) {
		int j = (int) ((k >> 1) < BITONIC_BLOCK / 2 ? (k >> 1) : BITONIC_BLOCK / 2);
		for (; j > 0; j >>= 1) {
			for (int t = threadIdx.x; t < BITONIC_BLOCK / 2; t += BITONIC_THREADS) {
				const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));        // index with bit j clear
				const int64_t li = (base + i) & (sigma_p - 1);
				cmpswap(s[i], s[i | j], (li & k) == 0);
			}
			__syncthreads();
		}
	}
	for (int i = threadIdx.x; i < BITONIC_BLOCK; i += BITONIC_THREADS)
		keys[base + i] = s[i];
}

__global__ void bitonic_global_kernel(uint64_t * __restrict__ keys, int64_t npairs, int64_t sigma_p, int64_t j, int64_t k)
{
	const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= npairs)
		return;
	const int64_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
	const int64_t li = i & (sigma_p - 1);
	uint64_t a = keys[i], b = keys[i | j];
	if ((a > b) == ((li & k) == 0)) {
		keys[i] = b;
		keys[i | j] = a;
	}
}

__global__ void sell_order_kernel(const uint64_t * __restrict__ keys, int64_t m, int64_t sigma, int64_t sigma_p, int C,
                                  int32_t * __restrict__ row_order, int32_t * __restrict__ slice_width, int64_t * __restrict__ slice_size)
{
	const int64_t pos = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (pos >= m)
		return;
	const int64_t w = pos / sigma, r = pos % sigma;
	const uint64_t key = keys[w * sigma_p + r];
	row_order[pos] = (int32_t) (uint32_t) (key & 0xFFFFFFFFu);
	if (pos % C == 0) {
		const uint32_t len = 0xFFFFFFFFu - (uint32_t) (key >> 32);
		slice_width[pos / C] = (int32_t) len;
		slice_size[pos / C] = (int64_t) len * C;
	}
}

// Work items of the SpMV kernel: every slice is cut along k into chunks of SELL_KCH columns and
// across lanes into groups of 32; a slice cut into n > 1 chunks also gets one "cut" record per lane
// group that tells the finishing warp where its n partials lie.
__global__ void sell_item_count_kernel(const int32_t * __restrict__ slice_width, int64_t nslices, int G,
                                       int64_t * __restrict__ item_off, int64_t * __restrict__ slot_off, int64_t * __restrict__ cut_off)
{
	const int64_t s = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (s > nslices)
		return;
	int64_t n = 0;
	if (s < nslices) {
		n = (slice_width[s] + SELL_KCH - 1) / SELL_KCH;
		if (n < 1)
			n = 1;                       // zero-width slices still zero their rows
	}
	item_off[s] = n * G;
	slot_off[s] = (n > 1) ? n * G : 0;
	cut_off[s] = (n > 1) ? G : 0;
}

// Item record (16 bytes, everything the SpMV warp needs, no dependent look-ups):
//   x = low 32 bits of the first entry's offset in scol / sval (slice_ptr[s] + k0 * C + 32 * g)
//   y = bits 0-15: high bits of that offset, bits 16-31: number of slice columns in this item (<= SELL_KCH)
//   z = partial slot, or -1 when the item is the whole slice (the warp stores y itself)
//   w = sorted-row index of lane 0 (s * C + 32 * g)
__global__ void sell_item_fill_kernel(const int32_t * __restrict__ slice_width, const int64_t * __restrict__ slice_ptr, int C,
                                      int64_t nslices, int G, const int64_t * __restrict__ item_off,
                                      const int64_t * __restrict__ slot_off, const int64_t * __restrict__ cut_off,
                                      int4 * __restrict__ items, int4 * __restrict__ cuts, int32_t * __restrict__ slot_cut)
{
	const int64_t s = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= nslices)
		return;
	int n = (slice_width[s] + SELL_KCH - 1) / SELL_KCH;
	if (n < 1)
		n = 1;
	int64_t io = item_off[s];
	const int64_t so = slot_off[s], co = cut_off[s];
	for (int g = 0; g < G; g++) {
		for (int j = 0; j < n; j++) {
			const int64_t off = slice_ptr[s] + (int64_t) j * SELL_KCH * C + 32 * g;
			int kc = slice_width[s] - j * SELL_KCH;
			kc = kc < 0 ? 0 : (kc > SELL_KCH ? SELL_KCH : kc);
			items[io++] = make_int4((int) (uint32_t) (off & 0xffffffffLL), (int) ((off >> 32) & 0xffff) | (kc << 16),
			                        (n > 1) ? (int) (so + (int64_t) g * n + j) : -1, (int) (s * C + 32 * g));
			if (n > 1)
				slot_cut[so + (int64_t) g * n + j] = (int32_t) (co + g);
		}
		if (n > 1)
			cuts[co + g] = make_int4((int) s, (int) (so + (int64_t) g * n), n, g);
	}
}

static int sell_plan_items(b200spmv * h)
{
	cudaStream_t st = h->stream;
	const int64_t ns = h->nslices;
	const int G = h->C >= 32 ? h->C / 32 : 1;
	h->sell_nitems = h->sell_ncuts = h->sell_nslots = 0;
	if (ns == 0)
		return 0;
	int64_t * d_off;
	CU(cudaMalloc(&d_off, 3 * (ns + 1) * sizeof(int64_t)));
	int64_t * d_item = d_off, * d_slot = d_off + (ns + 1), * d_cut = d_off + 2 * (ns + 1);
	sell_item_count_kernel<<<(unsigned) ((ns + 1 + 255) / 256), 256, 0, st>>>(h->d_slice_width, ns, G, d_item, d_slot, d_cut);
	CU_LAUNCH();
	if (device_exclusive_scan_i64(d_item, ns + 1, st) || device_exclusive_scan_i64(d_slot, ns + 1, st) || device_exclusive_scan_i64(d_cut, ns + 1, st)) {
		cudaFree(d_off);
		return 1;
	}
	int64_t tot[3];
	CU(cudaMemcpyAsync(&tot[0], d_item + ns, 8, cudaMemcpyDeviceToHost, st));
	CU(cudaMemcpyAsync(&tot[1], d_slot + ns, 8, cudaMemcpyDeviceToHost, st));
	CU(cudaMemcpyAsync(&tot[2], d_cut + ns, 8, cudaMemcpyDeviceToHost, st));
	CU(cudaStreamSynchronize(st));
	h->sell_nitems = tot[0];
	h->sell_nslots = tot[1];
	h->sell_ncuts = tot[2];
	{
		// chunk boundaries of the host-vector path (capi.cu plan_chunks): 4 ranges on sort-window boundaries; a slice's items are
		// contiguous, so a range holds whole slices and every cut is finished inside its own range's launch
		const int64_t spw = h->sigma / h->C, nwin = (h->m + h->sigma - 1) / h->sigma;
		for (int c = 0; c <= 4; c++) {
			int64_t s = (nwin * c / 4) * spw;
			s = s > ns ? ns : s;
			if (c == 4)
				s = ns;
			CU(cudaMemcpyAsync(&h->sell_chunk_item[c], d_item + s, 8, cudaMemcpyDeviceToHost, st));
		}
		CU(cudaStreamSynchronize(st));
	}
	if (tot[1] >= 0x7fffffffLL / 32 || tot[0] >= 0x7fffffffLL) {
		cudaFree(d_off);
		B200_FAIL("SELL-C-sigma plan too large (%lld items)", (long long) tot[0]);
	}
	CU(cudaMalloc(&h->d_sell_items, (size_t) tot[0] * sizeof(int4)));
	CU(cudaMalloc(&h->d_sell_cuts, (size_t) (tot[2] > 0 ? tot[2] : 1) * sizeof(int4)));
	CU(cudaMalloc(&h->d_sell_partial, (size_t) (tot[1] > 0 ? tot[1] : 1) * 32 * sizeof(double)));
	CU(cudaMalloc(&h->d_sell_slot_cut, (size_t) (tot[1] > 0 ? tot[1] : 1) * sizeof(int32_t)));
	CU(cudaMalloc(&h->d_sell_cut_cnt, (size_t) (tot[2] > 0 ? tot[2] : 1) * sizeof(unsigned int)));
	CU(cudaMemsetAsync(h->d_sell_cut_cnt, 0, (size_t) (tot[2] > 0 ? tot[2] : 1) * sizeof(unsigned int), st));
	if (h->padded >= (1LL << 47)) {
		cudaFree(d_off);
		B200_FAIL("SELL-C-sigma padded size does not fit the item record");
	}
	sell_item_fill_kernel<<<(unsigned) ((ns + 255) / 256), 256, 0, st>>>(h->d_slice_width, h->d_slice_ptr, h->C, ns, G, d_item, d_slot, d_cut,
	                                                                    reinterpret_cast<int4 *>(h->d_sell_items), reinterpret_cast<int4 *>(h->d_sell_cuts), h->d_sell_slot_cut);
	CU_LAUNCH();
	CU(cudaStreamSynchronize(st));
	CU(cudaFree(d_off));
	return 0;
}

// sorted row r, k-th entry -> slice_ptr[r / C] + r % C + k * C   (sellcs_format.c:173-186)
//
// One warp per WORK ITEM of the product plan (slice, group of 32 lanes, chunk of SELL_KCH slice columns; sell_plan_items):
// the CSR -> column-major move is a transpose, done through a 32 x 32 shared-memory tile so that BOTH sides are coalesced --
// each row's entries are read as one contiguous run (lane = column), each slice column is written as one contiguous run
// (lane = row).  Items, not slices, are the unit so that a slice holding one of the generator's giant skew rows (500 100
// entries) is spread over thousands of warps.  Narrow items (<= 16 columns, the short-row matrices) skip the tile: thread
// per row, strided reads served by L1, coalesced writes.
// Round 1's kernel (thread per row; rows >= 64 copied by the warp with stride-C writes, i.e. 32 different lines per store)
// took 2.4 - 8.6 ms on the 44 M-nnz matrices of the bench sample with rows >= 100 long or a giant row, 14 - 50 x the HBM time
// of the move (profiles/r02f_launches_bench_summary.txt).  Padding entries stay as the memset left them: (0, 0.0).
constexpr int SCAT_WARPS = 2;

__global__ void __launch_bounds__(32 * SCAT_WARPS)
sell_scatter_items_kernel(const int4 * __restrict__ items, int64_t nitems, const int32_t * __restrict__ row_ptr, const int32_t * __restrict__ col,
                          const double * __restrict__ val, const int32_t * __restrict__ row_order, const int64_t * __restrict__ slice_ptr,
                          int64_t m, int C, int32_t * __restrict__ scol, double * __restrict__ sval)
{
	__shared__ double s_v[SCAT_WARPS][32][33];
	__shared__ int s_c[SCAT_WARPS][32][33];
	const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
	const int64_t it = (int64_t) blockIdx.x * SCAT_WARPS + wp;
	if (it >= nitems)
		return;                                      // warp-uniform; only warp-level synchronisation below
	const int4 item = items[it];
	const int kc = (int) ((uint32_t) item.y >> 16);
	const int64_t off = ((int64_t) (item.y & 0xffff) << 32) + (int64_t) (uint32_t) item.x;
	const int64_t w = item.w;
	const int64_t s = w / C;
	const int k0 = (int) ((off - slice_ptr[s] - (w - s * C)) / C);          // first slice column of this item
	const int64_t r = w + lane;
	int64_t src = 0;
	int cnt = 0;                                         // this lane's (= row's) entries inside the item's columns
	if (lane < C && r < m) {
		const int orig = row_order[r];
		src = row_ptr[orig];
		cnt = row_ptr[orig + 1] - (int) src - k0;
		cnt = cnt < 0 ? 0 : (cnt > kc ? kc : cnt);
		src += k0;
	}
	const int64_t dst = off + lane;
	int wmax = cnt;
	#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		const int t = __shfl_xor_sync(0xffffffffu, wmax, o);
		wmax = t > wmax ? t : wmax;
	}
	if (wmax == 0)
		return;
	if (wmax <= 16) {
		for (int k = 0; k < cnt; k++) {
			scol[dst + (int64_t) k * C] = col[src + k];
			sval[dst + (int64_t) k * C] = val[src + k];
		}
		return;
	}
	for (int kb = 0; kb < wmax; kb += 32) {
		unsigned mask = __ballot_sync(0xffffffffu, cnt > kb);
		while (mask) {
			const int l = __ffs(mask) - 1;
			mask &= mask - 1;
			const int64_t s2 = __shfl_sync(0xffffffffu, src, l);
			const int n2 = __shfl_sync(0xffffffffu, cnt, l);
			const int k = kb + lane;
			if (k < n2) {
				s_c[wp][l][lane] = col[s2 + k];
				s_v[wp][l][lane] = val[s2 + k];
			}
		}
		__syncwarp();
		int lim = wmax - kb;
		lim = lim > 32 ? 32 : lim;
		for (int kk = 0; kk < lim; kk++)
			if (kb + kk < cnt) {
				scol[dst + (int64_t) (kb + kk) * C] = s_c[wp][lane][kk];
				sval[dst + (int64_t) (kb + kk) * C] = s_v[wp][lane][kk];
			}
		__syncwarp();
	}
}

int sell_convert(b200spmv * h)
{
	B200Range nvtx_range_("b200spmv sell_convert");
	const int64_t m = h->m;
	const int C = h->C;
	const int64_t sigma = h->sigma;
	cudaStream_t st = h->stream;
	h->nslices = (m + C - 1) / C;
	CU(cudaMalloc(&h->d_row_order, (m > 0 ? m : 1) * sizeof(int32_t)));
	CU(cudaMalloc(&h->d_slice_width, (h->nslices > 0 ? h->nslices : 1) * sizeof(int32_t)));
	CU(cudaMalloc(&h->d_slice_ptr, (h->nslices + 1) * sizeof(int64_t)));
	CU(cudaMemsetAsync(h->d_slice_ptr, 0, (h->nslices + 1) * sizeof(int64_t), st));
	if (m > 0) {
		int64_t win = sigma < m ? sigma : m;
		int64_t sigma_p = 1;
		while (sigma_p < win)
			sigma_p <<= 1;
		const int64_t nwin = (m + sigma - 1) / sigma;
		int64_t nslots = nwin * sigma_p;
		// the shared-memory phases work on whole blocks: round the key array up with sentinels
		const int64_t nslots_alloc = (nslots + BITONIC_BLOCK - 1) / BITONIC_BLOCK * BITONIC_BLOCK;
		uint64_t * d_keys;
		CU(cudaMalloc(&d_keys, nslots_alloc * sizeof(uint64_t)));
		CU(cudaMemsetAsync(d_keys, 0xFF, nslots_alloc * sizeof(uint64_t), st));
		sell_keys_kernel<<<(unsigned) ((nslots + 255) / 256), 256, 0, st>>>(h->d_row_ptr, m, sigma, sigma_p, nslots, d_keys);
		CU_LAUNCH();
		const unsigned nblk = (unsigned) (nslots_alloc / BITONIC_BLOCK);
		if (sigma_p > 1) {
			bitonic_smem_kernel<<<nblk, BITONIC_THREADS, 0, st>>>(d_keys, sigma_p, 0);
			CU_LAUNCH();
			for (int64_t k = 2 * (int64_t) BITONIC_BLOCK; k <= sigma_p; k <<= 1) {
				const int64_t npairs = nslots_alloc / 2;
				for (int64_t j = k >> 1; j >= BITONIC_BLOCK; j >>= 1) {
					bitonic_global_kernel<<<(unsigned) ((npairs + 255) / 256), 256, 0, st>>>(d_keys, npairs, sigma_p, j, k);
					CU_LAUNCH();
				}
				bitonic_smem_kernel<<<nblk, BITONIC_THREADS, 0, st>>>(d_keys, sigma_p, k);
				CU_LAUNCH();
			}
		}
		sell_order_kernel<<<(unsigned) ((m + 255) / 256), 256, 0, st>>>(d_keys, m, sigma, sigma_p, C, h->d_row_order, h->d_slice_width,
		                                                                h->d_slice_ptr);
		CU_LAUNCH();
		CU(cudaStreamSynchronize(st));
		CU(cudaFree(d_keys));
		if (device_exclusive_scan_i64(h->d_slice_ptr, h->nslices + 1, st))
			return 1;
	}
	CU(cudaMemcpyAsync(&h->padded, h->d_slice_ptr + h->nslices, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
	CU(cudaStreamSynchronize(st));
	if (h->sell_pad_limit > 0 && h->padded > h->sell_pad_limit)
		return 2;                                   // AUTO trial: too much padding, candidate skipped (not an error)
	const int64_t alloc = h->padded + 8;
	CU(cudaMalloc(&h->d_scol, alloc * sizeof(int32_t)));
	CU(cudaMalloc(&h->d_sval, alloc * sizeof(double)));
	CU(cudaMemsetAsync(h->d_scol, 0, alloc * sizeof(int32_t), st));
	CU(cudaMemsetAsync(h->d_sval, 0, alloc * sizeof(double), st));
	// the product's work items first: they are also the units of the scatter
	if (sell_plan_items(h))
		return 1;
	if (h->sell_nitems > 0) {
		sell_scatter_items_kernel<<<(unsigned) ((h->sell_nitems + SCAT_WARPS - 1) / SCAT_WARPS), 32 * SCAT_WARPS, 0, st>>>(
			reinterpret_cast<const int4 *>(h->d_sell_items), h->sell_nitems, h->d_row_ptr, h->d_col, h->d_val, h->d_row_order, h->d_slice_ptr, m, C,
			h->d_scol, h->d_sval);
		CU_LAUNCH();
	}
	CU(cudaStreamSynchronize(st));
	return 0;
}

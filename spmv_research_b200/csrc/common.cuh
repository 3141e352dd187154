// common.cuh -- shared declarations of libb200spmv: error plumbing, sm_100a load helpers, handle.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <nvtx3/nvToolsExt.h>      // header-only; a no-op unless a profiler (nsys / ncu --nvtx) is attached
#include "../../include/b200spmv.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libb200spmv is written for sm_100a (B200) only"
#endif

extern thread_local char g_b200_err[512];

#define B200_FAIL(...)                                                    \
	do {                                                              \
		snprintf(g_b200_err, sizeof(g_b200_err), __VA_ARGS__);    \
		return 1;                                                 \
	} while (0)

#define CU(call)                                                                                            \
	do {                                                                                                \
		cudaError_t e_ = (call);                                                                    \
		if (e_ != cudaSuccess)                                                                      \
			B200_FAIL("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));     \
	} while (0)

#define CU_LAUNCH() CU(cudaPeekAtLastError())

// NVTX range over a C-ABI entry point / a create-time phase (the reference's SDV / TIME_IT region markers,
// bench_spmv.cpp:321-324, cuda_csr.cu:138-204, become profiler ranges).  Not placed on the device-resident product path.
struct B200Range {
	explicit B200Range(const char * name) { nvtxRangePushA(name); }
	~B200Range() { nvtxRangePop(); }
};

// ---------------------------------------------------------------------------------------------
// Tunables of the nnz-tiled CSR kernel
// ---------------------------------------------------------------------------------------------
#ifndef CSR_EXP_THREADS
#define CSR_EXP_THREADS 256               // experiment switch (tools/sweep_exp.py): 128 = half-size CTAs, twice as many resident
#endif
constexpr int CSR_THREADS = CSR_EXP_THREADS;      // threads per tile CTA
constexpr int CSR_ZERO_CTAS = 148 * 8;    // most CTAs csr_zero_rows_kernel is launched with (compressed rows); one norm slot each
constexpr int SELL_THREADS = 128;         // threads per SELL CTA: four warps, one work item each
constexpr int SELL_KCH = 128;             // slice columns per work item

// ---------------------------------------------------------------------------------------------
// Blackwell load helpers.
//  * matrix streams (values / col_ind): read exactly once -> non-coherent loads that do not
//    allocate in L1 (SASS: LDG.E.NA...CONSTANT) and carry an L2 evict-first policy, so they do
//    not push x out of the 126 MB L2.
//  * x gathers: L1-allocating loads with an L2 evict_last cache hint (LDG.E.EL.64.CONSTANT).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t make_policy_evict_first()
{
	uint64_t pol;
	asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
	return pol;
}

__device__ __forceinline__ double ld_stream_f64(const double * p, uint64_t pol)
{
	double r;
	asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(r) : "l"(p), "l"(pol));
	return r;
}

__device__ __forceinline__ int ld_stream_i32(const int32_t * p, uint64_t pol)
{
	int r;
	asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.s32 %0, [%1], %2;" : "=r"(r) : "l"(p), "l"(pol));
	return r;
}


__device__ __forceinline__ uint64_t make_policy_evict_last()
{
	uint64_t pol;
	asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
	return pol;
}

// (A fractional evict_last policy -- createpolicy.fractional with budget / band -- was measured for x bands larger than
// L2: <= 2 % at any budget, profiles/r01l_sweep_x_l2_budget.txt, and the run-time policy cost the long-row matrices 2-3 %;
// removed.  Column panels, panels.cu, are what fixes those bands.)
#ifndef CSR_EXP_XNA
#define CSR_EXP_XNA 0                     // experiment switch: 1 = x gathers do not allocate in L1 (L2 evict_last kept)
#endif
template <bool HINT>
__device__ __forceinline__ double ld_x(const double * p, uint64_t pol)
{
	double r;
	if (HINT && CSR_EXP_XNA)
		asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(r) : "l"(p), "l"(pol));
	else if (HINT)
		asm volatile("ld.global.nc.L1::evict_last.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(r) : "l"(p), "l"(pol));
	else
		asm volatile("ld.global.nc.f64 %0, [%1];" : "=d"(r) : "l"(p));
	return r;
}

// y destinations of one product: the caller's y plus optional peer buffers, optional scale / norm.
struct YOut {
	double * y;
	double * peer[8];
	int n_peer;
	int64_t peer_lo, peer_hi; // rows stored to the peers: [peer_lo, peer_hi)
	int accum;                // column panels: add the partial already in y[row] (earlier panels) to this panel's row sum
	int partial;              // column panels: not the last panel -> store the raw sum to y only (no scale / peers / norm)
	const double * scale;     // device scalar or nullptr
	double * norm_partial;    // one slot per CTA or nullptr
};

__device__ __forceinline__ void y_store(const YOut & o, int64_t row, double v)
{
	o.y[row] = v;
	const bool send = (row >= o.peer_lo) && (row < o.peer_hi);
	#pragma unroll
	for (int d = 0; d < 8; d++)          // unrolled + predicated: keeps the parameter struct out of local memory
		if (d < o.n_peer && send)
			o.peer[d][row] = v;
}

// Finishes one row: returns the value that entered y (what the norm accumulates), 0 for a partial panel.
__device__ __forceinline__ double y_finish(const YOut & o, int64_t row, double sum, double scale)
{
	if (o.accum)
		sum += o.y[row];
	if (o.partial) {
		o.y[row] = sum;
		return 0.0;
	}
	sum *= scale;
	y_store(o, row, sum);
	return sum;
}

__device__ __forceinline__ double warp_sum(double v)
{
	#pragma unroll
	for (int o = 16; o > 0; o >>= 1)
		v += __shfl_xor_sync(0xffffffffu, v, o);
	return v;
}

// ---------------------------------------------------------------------------------------------
// The handle
// ---------------------------------------------------------------------------------------------
struct b200spmv {
	int format, device;           // format: the format that runs (AUTO resolves to CSR or SELL at create time)
	int requested_format;
	int x_uploaded;               // the host-vector entry point has staged an x at least once
	float trial_ms[4];            // AUTO: measured product time of {CSR, SELL sigma, SELL sigma_alt, -}; 0 = not tried
	int64_t sell_pad_limit;       // AUTO: a SELL candidate padded beyond this many entries is skipped (0 = no limit)
	int64_t m, n, nnz;
	int64_t col_lo, col_hi;       // referenced columns lie in [col_lo, col_hi); x outside is never read
	int x_hint;
	cudaStream_t stream;
	float convert_ms;

	// CSR (device).  col / val are padded with zeros up to a multiple of the tile size + ext.
	int32_t * d_row_ptr;
	int32_t * d_col;
	double * d_val;

	// Row compression (matrices with many / long runs of empty rows): the tile plan and the kernels then work on the
	// non-empty rows only -- rp_plan[k] = row_ptr of the k-th non-empty row, rows_nz[k] = its index -- and the empty rows
	// are zeroed by a separate, balanced launch.  Without compression rp_plan = d_row_ptr, m_plan = m, rows_nz = nullptr.
	int compress_opt;         // 0 = auto, 1 = never, 2 = always
	int compressed;
	int64_t m_plan, n_empty;
	int32_t * d_rp_plan;      // alias of d_row_ptr or d_rp_nz
	int32_t * d_rp_nz;        // [m_plan + 1]
	int32_t * d_rows_nz;      // [m_plan + 1], sentinel m
	int32_t * d_empty_rows;   // [n_empty]

	// CSR tile plan
	int npt;                  // nnz per thread
	int tile_nnz;             // CSR_THREADS * npt
	int ext;                  // tail-extension capacity
	int64_t ntiles;
	int32_t * d_tile_row;     // [ntiles + 1]
	int64_t nsplit;
	int32_t * d_split_row;    // [nsplit]
	int32_t * d_split_tile;   // [nsplit]
	int32_t * d_tile_split;   // int2[ntiles]: {split row owned, split row entering} or -1 (csr_fold_split_rows); null when nsplit = 0
	int32_t * d_split_info;   // int4[nsplit]: {row, owner tile, contributing tiles, 0}
	unsigned int * d_split_cnt;   // [nsplit] arrival counters, zero between products
	int fold_splits;              // 1: split rows are finished inside the tile kernel (few of them); 0: by csr_fixup_kernel
	double * d_tile_head;     // [ntiles] partial of a split row entering the tile
	double * d_tile_tail;     // [ntiles] partial of a split row leaving the tile
	uint16_t * d_perm;        // CSR_EXP_PERM builds only (csr_perm.h): thread slot -> tile-local row; null otherwise
	int64_t * d_perm_off;     // [ntiles + 1] slot offsets
	uint8_t * d_rot;          // CSR_EXP_ROT builds only (csr_perm.h): start rotation of each row's sum; null otherwise

	// SELL-C-sigma (device)
	int C;
	int sell_variant;         // kernel shape (k steps in flight x resident warps / SM): 0 = by average slice width,
	                          // 1 = 8 x 32, 2 = 8 x 48, 3 = 4 x 64, 4 = 16 x 16
	void * d_trial_flush;     // AUTO trial scratch (freed before create returns)
	int64_t sigma, nslices, padded;
	int32_t * d_row_order;
	int32_t * d_slice_width;
	int64_t * d_slice_ptr;
	int32_t * d_scol;
	double * d_sval;
	int32_t * d_sell_items;   // int4[sell_nitems]  {slice, first column, partial slot or -1, lane group}
	int32_t * d_sell_cuts;    // int4[sell_ncuts]   {slice, first partial slot, items, lane group}
	double * d_sell_partial;  // [sell_nslots * 32]
	int32_t * d_sell_slot_cut;        // [sell_nslots] cut record of each partial slot
	unsigned int * d_sell_cut_cnt;    // [sell_ncuts] arrival counters, zero between products
	int64_t sell_nitems, sell_ncuts, sell_nslots;

	// diag(A) for the Jacobi-preconditioned CG (keep_diag), number of rows without a non-zero diagonal entry
	double * d_diag;
	int diag_bad;

	// Host-vector entry point: the product is launched in up to 4 unit ranges (tiles / SELL items) whose rows are final
	// when the range's kernel ends, so the D2H copy of a range overlaps the kernel of the next one.
	int nchunks;                  // 0 / 1 = one launch
	int64_t chunk_unit[5], chunk_row[5];
	int64_t sell_chunk_item[5];   // SELL: first item of the slice at each chunk boundary (set by sell_plan_items; slices with cuts have several items)
	int n_chunk_fix;              // CSR: split rows whose tiles straddle a chunk boundary (<= 3): final only after a LATER range,
	int64_t chunk_fix_row[4];     //      so their y entry is copied once more after the last range
	int64_t range_first, range_count;      // range_count < 0: whole matrix (set only around the chunked launches)
	cudaStream_t copy_stream;
	cudaEvent_t chunk_ev[4];
	// host_pipeline: x travels range by range on up_stream (columns [chunk_col_lo, chunk_col_hi) are what range c reads)
	int pipeline;
	int64_t blk_rows;             // rows per entry of the column-range table built at create time (0 = no table)
	int blk_lo[256], blk_hi[256]; // smallest / largest column of each block of blk_rows rows (lo > hi: no entries)
	int64_t chunk_col_lo[4], chunk_col_hi[4];
	cudaStream_t up_stream;
	cudaEvent_t up_ev[4];

	// Column panels (x band larger than L2): this handle is then a shell around npanels sub-handles, one per column
	// range, launched in turn; sub-handle p > 0 adds to the partial in y, the last one finishes the rows (scale, peers, norm).
	int npanels;
	struct b200spmv * panel[8];
	int acc_mode;                 // bit 0: accumulate into y, bit 1: partial (not the last panel)

	int pin_host;                 // page-lock the caller's x / y on first sight
	void * pinned[4];             // buffers this handle registered (unregistered at destroy)
	int n_pinned;

	// staging for the host-vector entry point, norm partials
	double * d_x;
	double * d_y;
	double * d_norm_partial;
	int64_t norm_slots;
};

// convert.cu
int csr_plan_tiles(b200spmv * h);
int sell_convert(b200spmv * h);
int device_exclusive_scan_i64(int64_t * d_data, int64_t n, cudaStream_t st);
// kernels
int launch_csr_spmv(b200spmv * h, const double * dx, double * dy, cudaStream_t st, const b200spmv_launch * l);
int launch_sell_spmv(b200spmv * h, const double * dx, double * dy, cudaStream_t st, const b200spmv_launch * l);
int launch_norm_reduce(b200spmv * h, int64_t slots, double * d_out, cudaStream_t st);
int launch_csr_tiles_range(b200spmv * h, const double * dx, double * dy, cudaStream_t st, const b200spmv_launch * l, int64_t first, int64_t count);
int launch_product(b200spmv * h, const double * dx, double * dy, cudaStream_t st, const b200spmv_launch * l);      // capi.cu
int launch_spmv_any(b200spmv * h, const double * dx, double * dy, cudaStream_t st);     // capi.cu: dispatch on h->format
int pcg_extract_diag(b200spmv * h);
// panels.cu
int panels_auto_count(const b200spmv * h, int requested);
int panels_split(b200spmv * h, int P, const b200spmv_opts * sub_opts);
// capi.cu
int create_from_device(b200spmv ** out, const int32_t * d_row_ptr, const int32_t * d_col, const double * d_val, int64_t m, int64_t n,
                       int64_t nnz, const b200spmv_opts * opts);
void destroy_handle(b200spmv * h);                                                     // pcg.cu

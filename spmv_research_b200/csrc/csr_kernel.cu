// csr_kernel.cu -- the adaptive, nnz-tiled fp64 CSR SpMV kernel for sm_100a.
//
// Computes what the reference's subkernel_csr_scalar computes (K/csr.cpp:334-350:
// y[i] = sum_j a[j] * x[ja[j]], every row written, empty rows -> 0), laid out for B200:
//
//  * the nnz stream is cut into fixed windows ("tiles") of T = 256 * NPT non-zeros, one CTA per
//    tile, so every values / col_ind load is aligned and fully coalesced regardless of the
//    row-length distribution -- the nnz-balanced decomposition of merge-path
//    (K/merge.cpp:256-318) with the row search done once at convert time (tile_info);
//  * lane l of a warp takes non-zero 32*k + l of the warp's share ("lane-consecutive"): measured
//    on B200 (tools/hbm_probe.cu, profiles/) the x gather is limited to one 128-byte L1 wavefront
//    per clock per SM, and a warp-wide gather costs one wavefront per DISTINCT line, so putting
//    the neighbouring non-zeros of a row (clustered columns, similar adjacent rows) into ONE
//    gather instruction is worth up to 1.6x over giving a thread 8 consecutive non-zeros;
//  * all 2*NPT stream loads of a thread are issued before the first x gather (L1 no-allocate,
//    L2 evict-first policy), products go to shared memory with conflict-free 64-bit stores;
//  * a row belongs to the tile in which it STARTS.  The owner finishes the row even when it
//    leaves the window, by loading up to EXT further non-zeros ("tail extension"), so ordinary
//    matrices need one launch and no carry fix-up.  Only rows sticking out by more than EXT
//    ("split" rows: the generator's giant skew rows) are completed from per-tile head/tail partials in a fixed
//    order by whichever of their tiles finishes last (csr_fold_split_rows) -- deterministic, no atomics on y
//    (cf. cuda_csr_const_nnz_per_thread.cu:578-659,864 which uses atomicAdd + memset);
//  * per tile the row reduction is binned by the number of rows it owns: thread per row (short
//    rows, sequential ascending-j sum = the oracle's order), 2..32 lanes per row with shuffle
//    reduction (long rows), whole CTA for split-row partials.
#include "common.cuh"

// Experiment switches for the row reduction (tools/sweep_exp.py builds extra libraries with them; measured on a B200 in round 2,
// profiles/r02c_sweep_exp_row_reduction.txt):
//   CSR_EXP_LDS128  = 1: thread-per-row bins read the products two at a time (LDS.128 on even positions) -- same additions
//                        in the same ascending-j order, half the shared-memory load instructions;
//   CSR_EXP_MIN_LPR = 2 / 4 / 8: no thread-per-row bin; every tile reduces with at least that many lanes per row (lanes of a
//                        row read adjacent products, so a half-warp touches few distinct rows: fewer bank conflicts, at the
//                        price of log2(lanes) shuffles per row and of the sequential summation order).
#ifndef CSR_EXP_LDS128
#define CSR_EXP_LDS128 1      // shipped since round 2: the only variant that won (-1 % on the sample, -4.5 % on config 1, y bit-identical;
#endif                        // profiles/r02c_sweep_exp_row_reduction.txt).  MIN_LPR / PERM / ROT all measured slower or flat.
#ifndef CSR_EXP_MIN_LPR
#define CSR_EXP_MIN_LPR 0
#endif
//   CSR_EXP_PERM    = 1: the bins with 1 / 2 / 4 / 8 lanes per row take their rows through a per-tile table (csr_perm.h) in
//                        which the rows of a half-warp start at positions whose bank windows cannot collide (thread per
//                        row: lane b only gets rows starting at b mod 16): the row sums are free of bank conflicts; y is
//                        bit-identical (the same lanes add the same products in the same order, elsewhere in the CTA).
#ifndef CSR_EXP_PERM
#define CSR_EXP_PERM 0
#endif
#if CSR_EXP_PERM && CSR_EXP_MIN_LPR
#error "CSR_EXP_PERM replays the shipped binning rule (csr_perm_lpr): not with CSR_EXP_MIN_LPR"
#endif
//   CSR_EXP_ROT     = 1: thread-per-row bins keep row = thread but start each row at a per-row rotation (csr_perm.h,
//                        csr_rot_build) chosen so that the 16 lanes of a half-warp collide least; wraps around the row.
//                        No holes, y stays coalesced; the sum is a cyclic rotation of the ascending-j one (not bit-equal).
#ifndef CSR_EXP_ROT
#define CSR_EXP_ROT 0
#endif
// Both: rotation in the thread-per-row bin, table in the lanes-per-row bins (the table builder then skips the former).
#if CSR_EXP_PERM && CSR_EXP_ROT
#define CSR_PERM_PARAMS , const uint16_t * __restrict__ perm, const int64_t * __restrict__ perm_off, const uint8_t * __restrict__ rot
#define CSR_PERM_ARGS , h->d_perm, h->d_perm_off, h->d_rot
#elif CSR_EXP_PERM
#define CSR_PERM_PARAMS , const uint16_t * __restrict__ perm, const int64_t * __restrict__ perm_off
#define CSR_PERM_ARGS , h->d_perm, h->d_perm_off
#elif CSR_EXP_ROT
#define CSR_PERM_PARAMS , const uint8_t * __restrict__ rot
#define CSR_PERM_ARGS , h->d_rot
#else
#define CSR_PERM_PARAMS
#define CSR_PERM_ARGS
#endif

__device__ __forceinline__ double block_sum_range(const double * s_prod, int a, int b, double * s_red)
{
	double acc = 0.0;
	for (int j = a + (int) threadIdx.x; j < b; j += CSR_THREADS)
		acc += s_prod[j];
	acc = warp_sum(acc);
	__syncthreads();
	if ((threadIdx.x & 31) == 0)
		s_red[threadIdx.x >> 5] = acc;
	__syncthreads();
	double tot = 0.0;
	#pragma unroll
	for (int w = 0; w < CSR_THREADS / 32; w++)
		tot += s_red[w];
	return tot;
}

// Split rows are finished inside the tile kernel (round 2; round 1 used a second launch, csr_fixup_kernel: 4.5 us + a launch gap
// on a ~115 us product).  Every tile a split row touches -- its owner (tail partial) and the tiles it enters (head partials) --
// bumps the row's counter after publishing its partial; the CTA that arrives last sums tail[owner], head[owner + 1], ...,
// head[last] in a FIXED order (thread j takes entries j, j + 256, ...; butterfly; warps in order), so y does not depend on
// which CTA happened to be last: deterministic, still no atomics on y.  The counter is reset by the finisher.
struct SplitFold {
	const int2 * tile_split;      // per tile: {split row owned (tail) or -1, split row entering (head) or -1}
	const int4 * info;            // per split row: {row (plan space), owner tile, number of contributing tiles, 0}
	unsigned int * cnt;           // per split row: contributions so far
	int64_t ntiles;
};

template <int MODE>
__device__ __noinline__ void csr_fold_split_rows(int64_t tile, bool tail_split, bool head_split, SplitFold sf, const double * tile_head,
                                                 const double * tile_tail, const int32_t * rowmap, YOut out, double scale, double * s_red)
{
	__shared__ int s_last;
	const int tid = threadIdx.x;
	const int2 ts = sf.tile_split[tile];
	#pragma unroll 1
	for (int which = 0; which < 2; which++) {
		const int s = (which == 0) ? (head_split ? ts.y : -1) : (tail_split ? ts.x : -1);
		if (s < 0)
			continue;                               // uniform over the CTA
		const int4 info = sf.info[s];
		if (tid == 0) {
			__threadfence();                        // this tile's partial (written by this thread) before the count
			const unsigned int old = atomicAdd(&sf.cnt[s], 1u);
			s_last = (old + 1u == (unsigned int) info.z) ? 1 : 0;
		}
		__syncthreads();
		if (s_last) {
			__threadfence();
			double acc = 0.0;
			for (int j = tid; j < info.z; j += CSR_THREADS)
				acc += (j == 0) ? __ldcg(tile_tail + info.y) : __ldcg(tile_head + info.y + j);
			acc = warp_sum(acc);
			if ((tid & 31) == 0)
				s_red[tid >> 5] = acc;
			__syncthreads();
			if (tid == 0) {
				double tot = 0.0;
				#pragma unroll
				for (int w = 0; w < CSR_THREADS / 32; w++)
					tot += s_red[w];
				if constexpr (MODE == 0) {
					out.y[info.x] = tot;
				} else {
					tot = y_finish(out, (MODE == 2 && rowmap) ? rowmap[info.x] : info.x, tot, scale);
					if (out.norm_partial)
						out.norm_partial[sf.ntiles + s] = tot * tot;
				}
				sf.cnt[s] = 0u;                         // ready for the next product of this handle
			}
		}
		__syncthreads();
	}
}

// PLAIN: the product as the harness calls it (y = A x, nothing else): no row map, scale, norm, peer destinations or
// panel accumulation.  The kernel sits at its 40-register cap (6 CTAs per SM); each of those features measurably slowed
// the plain product when it was a run-time branch (+5 % in total), so they are compiled out of this instantiation.
// MODE 0 = PLAIN; 1 = EPILOGUE (scale, norm partials, peer destinations: the power-iteration product);
// 2 = FULL (adds the compressed-row map, the column-panel accumulate / partial modes and unit-range launches).
template <int NPT, bool HINT, int MODE>
__global__ void __launch_bounds__(CSR_THREADS, ((NPT <= 4) ? 8 : ((NPT <= 8) ? 6 : ((NPT <= 16) ? 3 : 2))) * (256 / CSR_THREADS))
csr_tile_kernel(const int32_t * __restrict__ row_ptr, const int32_t * __restrict__ col, const double * __restrict__ val,
                const int2 * __restrict__ tile_info, const double * __restrict__ x,
                double * __restrict__ tile_head, double * __restrict__ tile_tail,
                int64_t nnz, int ext_cap, const int32_t * __restrict__ rowmap, int64_t tile_base, YOut out, SplitFold sf CSR_PERM_PARAMS)
{
	constexpr bool PLAIN = (MODE == 0);
	constexpr int T = CSR_THREADS * NPT;
	extern __shared__ __align__(16) double s_prod[];      // T + ext_cap products
	__shared__ double s_red[CSR_THREADS / 32];
	__shared__ int s_rp[CSR_THREADS + 1];                 // window-relative row_ptr of the first 256 owned rows

	const int tid = threadIdx.x;
	const int64_t tile = (MODE == 2) ? tile_base + blockIdx.x : (int64_t) blockIdx.x;      // unit ranges run the FULL instantiation
	const int64_t base = tile * (int64_t) T;
	const uint64_t pol = make_policy_evict_last();
	const uint64_t pol_stream = make_policy_evict_first();

	// 1. all streaming loads of the window, issued back to back (lane-consecutive, coalesced)
	int c[NPT];
	double v[NPT];
	#pragma unroll
	for (int k = 0; k < NPT; k++) {
		const int64_t idx = base + k * CSR_THREADS + tid;
		c[k] = ld_stream_i32(col + idx, pol_stream);
		v[k] = ld_stream_f64(val + idx, pol_stream);
	}

	// 2. tile bookkeeping (independent of the stream loads)
	const int2 ti0 = tile_info[tile];
	const int2 ti1 = tile_info[tile + 1];
	const int R0 = ti0.x & 0x7fffffff;
	const bool head_split = (ti0.x < 0);
	const int R1 = ti1.x & 0x7fffffff;
	const int64_t start0 = ti0.y;                       // row_ptr[R0]
	const int64_t end_own = ti1.y;                      // row_ptr[R1]
	const int nrows = R1 - R0;
	const int64_t lim = (base + T < nnz) ? base + T : nnz;
	const int64_t over = end_own - (base + T);          // > 0: the last owned row leaves the window
	const bool tail_split = (nrows > 0) && (over > ext_cap);
	double scale = 1.0;
	if constexpr (!PLAIN)
		scale = out.scale ? *out.scale : 1.0;
	int rp_pre = 0, rp_pre2 = 0;
	const int npre = nrows < CSR_THREADS ? nrows : CSR_THREADS;
	const bool do_ext = (nrows > 0 && over > 0 && !tail_split);
	int ce = 0;
	double ve = 0.0;
	{
		// the first CSR_THREADS extension entries travel with the window's own loads (one per thread)
		if (do_ext && tid < over) {
			ce = ld_stream_i32(col + base + T + tid, pol_stream);
			ve = ld_stream_f64(val + base + T + tid, pol_stream);
		}
		if (tid <= npre)
			rp_pre = (int) (row_ptr[R0 + tid] - base);
		if (tid == 0 && npre == CSR_THREADS)
			rp_pre2 = (int) (row_ptr[R0 + CSR_THREADS] - base);
	}
#if CSR_EXP_PERM || CSR_EXP_ROT
	// The thread's first table entry / its rows' rotations travel with the stream loads (they would otherwise be dependent
	// global loads right after the barrier).  One register: a tile uses either the rotations (thread-per-row bin: low byte
	// row tid, next byte row tid + CSR_THREADS) or a table entry (slot tid there, slot tid / LPR in the lanes-per-row bins).
	int meta_pre = 0xFFFF;
	const bool tpr_bin = (tail_split ? nrows - 1 : nrows) > CSR_THREADS / 2;
#if CSR_EXP_PERM
	const int64_t pslot0 = perm_off[tile];
	const int pslots = (int) (perm_off[tile + 1] - pslot0);
	if (!(CSR_EXP_ROT && tpr_bin)) {
		const int rf = tail_split ? nrows - 1 : nrows;
		int idx = tid;
		if (!tpr_bin) {
			int lp = 2;
			while (lp < 32 && rf * (lp * 2) <= CSR_THREADS)
				lp *= 2;
			idx = tid / lp;
		}
		if (idx < pslots)
			meta_pre = perm[pslot0 + idx];
	}
#endif
#if CSR_EXP_ROT
	if (tpr_bin) {
		meta_pre = 0;
		if (tid < nrows)
			meta_pre = rot[R0 + tid];
		if (tid + CSR_THREADS < nrows)
			meta_pre |= (int) rot[R0 + tid + CSR_THREADS] << 8;
	}
#endif
#define perm_pre meta_pre
#define rot_pre meta_pre
#endif

	// 3. gathers and products -> shared memory in nnz order (conflict-free STS.64)
	#pragma unroll
	for (int k = 0; k < NPT; k++)
		v[k] *= ld_x<HINT>(x + c[k], pol);
	if (do_ext && tid < over)
		ve *= ld_x<HINT>(x + ce, pol);
	#pragma unroll
	for (int k = 0; k < NPT; k++)
		s_prod[k * CSR_THREADS + tid] = v[k];

	{
		if (do_ext && tid < over)
			s_prod[T + tid] = ve;
		if (tid <= npre)
			s_rp[tid] = rp_pre;
		if (tid == 0 && npre == CSR_THREADS)
			s_rp[CSR_THREADS] = rp_pre2;
	}
	// 4. tail extension: finish the last owned row past the window end
	if (do_ext) {
		for (int j = tid + CSR_THREADS; j < (int) over; j += CSR_THREADS) {
			const int64_t idx = base + T + j;
			s_prod[T + j] = ld_stream_f64(val + idx, pol_stream) * ld_x<HINT>(x + ld_stream_i32(col + idx, pol_stream), pol);
		}
	}
	__syncthreads();

	// 5. row reduction, binned by rows owned
	double nrm = 0.0;
	const int rows_full = tail_split ? nrows - 1 : nrows;      // rows this tile completes
	if (CSR_EXP_MIN_LPR == 0 && rows_full > CSR_THREADS / 2) {
		// thread per row: sequential ascending-j sum (the oracle's summation order)
#if CSR_EXP_PERM && !CSR_EXP_ROT
		const int64_t slot0 = pslot0;
		const int nslots = pslots;
		const int nwork = nslots > 0 ? nslots : rows_full;         // no table for this tile: row = thread, as shipped
		for (int sl = tid; sl < nwork; sl += CSR_THREADS) {
			int r = sl;
			if (nslots > 0) {
				r = sl == tid ? perm_pre : (int) perm[slot0 + sl];
				if (r == 0xFFFF)
					continue;                           // a hole of the table: no row for this lane in this half-warp
			}
#else
		for (int r = tid; r < rows_full; r += CSR_THREADS) {
#endif
			const int i = R0 + r;
			int a, b;
			if (r < CSR_THREADS) {
				a = s_rp[r];
				b = s_rp[r + 1];
			} else {
				a = (int) (row_ptr[i] - base);
				b = (int) (row_ptr[i + 1] - base);
			}
			double sum = 0.0;
#if CSR_EXP_ROT
			{
				int idx = a + (r == tid ? (rot_pre & 0xff) : (r == tid + CSR_THREADS ? (rot_pre >> 8) : (int) rot[i]));   // rotation < row length
				for (int j = a; j < b; j++) {
					sum += s_prod[idx];
					idx = (idx + 1 == b) ? a : idx + 1;
				}
			}
#elif CSR_EXP_LDS128
			if ((a & 1) && a < b)
				sum += s_prod[a++];
			for (; a + 1 < b; a += 2) {
				const double2 v2 = *reinterpret_cast<const double2 *>(s_prod + a);      // even position: 16-byte aligned
				sum += v2.x;
				sum += v2.y;
			}
			if (a < b)
				sum += s_prod[a];
#else
			for (; a < b; a++)
				sum += s_prod[a];
#endif
			if constexpr (PLAIN) {
				out.y[i] = sum;
			} else if constexpr (MODE == 1) {
				sum *= scale;
				y_store(out, i, sum);
				nrm += sum * sum;
			} else {
				sum = y_finish(out, rowmap ? rowmap[i] : i, sum, scale);
				nrm += sum * sum;
			}
		}
	} else if (rows_full > 0) {
		// LPR lanes per row, LPR = largest power of two with rows_full * LPR <= CSR_THREADS (max 32)
		int lpr = 2;
		while (lpr < 32 && rows_full * (lpr * 2) <= CSR_THREADS)
			lpr *= 2;
#if CSR_EXP_MIN_LPR
		if (lpr < CSR_EXP_MIN_LPR)
			lpr = CSR_EXP_MIN_LPR;
#endif
		const int groups = CSR_THREADS / lpr;
		const int g = tid / lpr, lane = tid % lpr;
#if CSR_EXP_PERM
		// group slots through the tile's table (csr_perm.h): the groups of a half-warp then read disjoint bank windows
		const int64_t gslot0 = pslot0;
		const int ngslots = pslots;
		const int nwork = ngslots > 0 ? ngslots : rows_full;
#else
		const int nwork = rows_full;
#endif
		for (int rb = 0; rb < nwork; rb += groups) {
#if CSR_EXP_PERM
			int r = rb + g;
			if (ngslots > 0)
				r = rb == 0 ? perm_pre : (r < ngslots ? (int) perm[gslot0 + r] : 0xFFFF);      // a hole (0xFFFF) is >= rows_full: skipped below
#else
			const int r = rb + g;
#endif
			double sum = 0.0;
			if (r < rows_full) {
#if CSR_EXP_MIN_LPR
				// more than CSR_THREADS rows can reach this bin: the staged row_ptr slice covers the first CSR_THREADS
				const int a = r < CSR_THREADS ? s_rp[r] : (int) (row_ptr[R0 + r] - base);
				const int b = r < CSR_THREADS ? s_rp[r + 1] : (int) (row_ptr[R0 + r + 1] - base);
#else
				const int a = s_rp[r], b = s_rp[r + 1];
#endif
				for (int j = a + lane; j < b; j += lpr)
					sum += s_prod[j];
			}
			for (int o = lpr >> 1; o > 0; o >>= 1)
				sum += __shfl_xor_sync(0xffffffffu, sum, o);
			if (r < rows_full && lane == 0) {
				if constexpr (PLAIN) {
					out.y[R0 + r] = sum;
				} else if constexpr (MODE == 1) {
					sum *= scale;
					y_store(out, R0 + r, sum);
					nrm += sum * sum;
				} else {
					sum = y_finish(out, rowmap ? rowmap[R0 + r] : R0 + r, sum, scale);
					nrm += sum * sum;
				}
			}
		}
	}

	// 6. split rows: whole-CTA partial sums over the part of the window they cover
	if (tail_split) {
		const int a = (int) (row_ptr[R1 - 1] - base);
		const double part = block_sum_range(s_prod, a, T, s_red);
		if (tid == 0)
			tile_tail[tile] = part;
	}
	if (head_split) {
		const int64_t hend = (start0 < lim) ? start0 : lim;
		const double part = block_sum_range(s_prod, 0, (int) (hend - base), s_red);
		if (tid == 0)
			tile_head[tile] = part;
	}
	if ((tail_split || head_split) && sf.tile_split)
		csr_fold_split_rows<MODE>(tile, tail_split, head_split, sf, tile_head, tile_tail, rowmap, out, scale, s_red);

	// 7. optional sum of squares of the rows written here (power iteration norm, no atomics)
	if constexpr (!PLAIN)
	if (out.norm_partial) {
		nrm = warp_sum(nrm);
		__syncthreads();
		if ((tid & 31) == 0)
			s_red[tid >> 5] = nrm;
		__syncthreads();
		if (tid == 0) {
			double tot = 0.0;
			#pragma unroll
			for (int w = 0; w < CSR_THREADS / 32; w++)
				tot += s_red[w];
			out.norm_partial[tile] = tot;
		}
	}
}

// Handles with MANY split rows (fold_splits = 0): one warp per split row after the tile kernel,
// y[row] = tail[owner] + head[owner+1] + ... + head[last], fixed order.
__global__ void csr_fixup_kernel(const int32_t * __restrict__ row_ptr, const int32_t * __restrict__ split_row,
                                 const int32_t * __restrict__ split_tile, const double * __restrict__ tile_head,
                                 const double * __restrict__ tile_tail, int64_t nsplit, int tile_nnz, int64_t ntiles,
                                 const int32_t * __restrict__ rowmap, YOut out)
{
	const int64_t s = blockIdx.x;
	if (s >= nsplit)
		return;
	const int lane = threadIdx.x;
	const int row = split_row[s];
	const int64_t k = split_tile[s];
	const int64_t e = row_ptr[row + 1];
	const int64_t last = (e - 1) / tile_nnz;
	double acc = (lane == 0) ? tile_tail[k] : 0.0;
	for (int64_t t = k + 1 + lane; t <= last; t += 32)
		acc += tile_head[t];
	acc = warp_sum(acc);
	if (lane == 0) {
		const double scale = out.scale ? *out.scale : 1.0;
		acc = y_finish(out, rowmap ? rowmap[row] : row, acc, scale);
		if (out.norm_partial)
			out.norm_partial[ntiles + s] = acc * acc;
	}
}

// Compressed row space: the empty rows, listed at plan time, are finished by a balanced grid-stride pass -- written 0 in a plain
// handle; in a column panel "empty" means empty in THIS panel, so the row goes through y_finish like any other (the partial
// of earlier panels is kept / scaled / sent to the peers / added to the norm).
__global__ void csr_zero_rows_kernel(const int32_t * __restrict__ empty_rows, int64_t n_empty, YOut out, int64_t norm_base)
{
	__shared__ double s_red[8];
	const double scale = out.scale ? *out.scale : 1.0;
	double nrm = 0.0;
	for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n_empty; i += (int64_t) gridDim.x * blockDim.x) {
		const double v = y_finish(out, empty_rows[i], 0.0, scale);
		nrm += v * v;
	}
	if (out.norm_partial) {
		nrm = warp_sum(nrm);
		if ((threadIdx.x & 31) == 0)
			s_red[threadIdx.x >> 5] = nrm;
		__syncthreads();
		if (threadIdx.x == 0) {
			double tot = 0.0;
			for (int w = 0; w < (int) (blockDim.x >> 5); w++)
				tot += s_red[w];
			out.norm_partial[norm_base + blockIdx.x] = tot;
		}
	}
}

// Deterministic final reduction of the per-CTA norm partials (single CTA, fixed tree).
__global__ void norm_reduce_kernel(const double * __restrict__ partial, int64_t n, double * __restrict__ out)
{
	__shared__ double s_red[32];
	double acc = 0.0;
	for (int64_t i = threadIdx.x; i < n; i += blockDim.x)
		acc += partial[i];
	acc = warp_sum(acc);
	if ((threadIdx.x & 31) == 0)
		s_red[threadIdx.x >> 5] = acc;
	__syncthreads();
	if (threadIdx.x == 0) {
		double tot = 0.0;
		for (int w = 0; w < (int) (blockDim.x >> 5); w++)
			tot += s_red[w];
		*out = tot;
	}
}

int launch_norm_reduce(b200spmv * h, int64_t slots, double * d_out, cudaStream_t st)
{
	norm_reduce_kernel<<<1, 1024, 0, st>>>(h->d_norm_partial, slots, d_out);
	CU_LAUNCH();
	return 0;
}

template <int NPT, bool HINT, int MODE>
static int launch_tiles_impl(b200spmv * h, const double * dx, cudaStream_t st, const YOut & out)
{
	const size_t smem = (size_t) (h->tile_nnz + h->ext) * sizeof(double);
	// the opt-in for > 48 KB of dynamic shared memory is per device (a process may drive several: opts.device)
	static unsigned char attr_set[64] = { 0 };
	const int dv = (h->device >= 0 && h->device < 64) ? h->device : 0;
	if (!attr_set[dv]) {
		CU(cudaFuncSetAttribute(csr_tile_kernel<NPT, HINT, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
		attr_set[dv] = 1;
	}
	const int64_t first = h->range_count >= 0 ? h->range_first : 0;
	const int64_t count = h->range_count >= 0 ? h->range_count : h->ntiles;
	if (count <= 0)
		return 0;
	SplitFold sf;
	sf.tile_split = h->fold_splits ? reinterpret_cast<const int2 *>(h->d_tile_split) : nullptr;
	sf.info = reinterpret_cast<const int4 *>(h->d_split_info);
	sf.cnt = h->d_split_cnt;
	sf.ntiles = h->ntiles;
	csr_tile_kernel<NPT, HINT, MODE><<<(unsigned) count, CSR_THREADS, smem, st>>>(
		h->d_rp_plan, h->d_col, h->d_val, reinterpret_cast<const int2 *>(h->d_tile_row), dx,
		h->d_tile_head, h->d_tile_tail, h->nnz, h->ext, h->compressed ? h->d_rows_nz : nullptr, first, out, sf CSR_PERM_ARGS);
	CU_LAUNCH();
	return 0;
}

template <int NPT, bool HINT>
static int launch_tiles(b200spmv * h, const double * dx, cudaStream_t st, const YOut & out)
{
	const bool full = h->compressed || out.accum || out.partial || h->range_count >= 0;
	const bool plain = !full && out.n_peer == 0 && !out.scale && !out.norm_partial;
	if (plain)
		return launch_tiles_impl<NPT, HINT, 0>(h, dx, st, out);
	return full ? launch_tiles_impl<NPT, HINT, 2>(h, dx, st, out) : launch_tiles_impl<NPT, HINT, 1>(h, dx, st, out);
}

// y destinations / epilogue of one launch from the caller's b200spmv_launch and the handle's panel role
static int csr_make_out(const b200spmv * h, double * dy, const b200spmv_launch * l, YOut & out)
{
	memset(&out, 0, sizeof(out));
	out.y = dy;
	if (l) {
		if (l->n_peer < 0 || l->n_peer > 8)
			B200_FAIL("n_peer must be in [0, 8]");
		out.n_peer = l->n_peer;
		for (int d = 0; d < l->n_peer; d++)
			out.peer[d] = l->y_peer[d];
		out.peer_lo = 0;
		out.peer_hi = h->m;
		if (l->peer_row_lo != 0 || l->peer_row_hi != 0) {
			out.peer_lo = l->peer_row_lo;
			out.peer_hi = l->peer_row_hi;
		}
		out.scale = l->scale;
		out.norm_partial = l->norm2 ? h->d_norm_partial : nullptr;
	}
	out.accum = h->acc_mode & 1;
	out.partial = (h->acc_mode >> 1) & 1;
	return 0;
}

static int csr_launch_tiles(b200spmv * h, const double * dx, cudaStream_t st, const YOut & out)
{
	const bool hint = (h->x_hint == 0);
	switch (h->npt) {
	case 4:  return hint ? launch_tiles<4, true>(h, dx, st, out) : launch_tiles<4, false>(h, dx, st, out);
	case 8:  return hint ? launch_tiles<8, true>(h, dx, st, out) : launch_tiles<8, false>(h, dx, st, out);
	case 16: return hint ? launch_tiles<16, true>(h, dx, st, out) : launch_tiles<16, false>(h, dx, st, out);
	case 32: return hint ? launch_tiles<32, true>(h, dx, st, out) : launch_tiles<32, false>(h, dx, st, out);
	default: B200_FAIL("unsupported csr_nnz_per_thread %d", h->npt);
	}
}

int launch_csr_spmv(b200spmv * h, const double * dx, double * dy, cudaStream_t st, const b200spmv_launch * l)
{
	YOut out;
	if (csr_make_out(h, dy, l, out))
		return 1;
	int64_t zero_ctas = 0;
	// rows that are empty in THIS handle: a plain handle writes them 0; a column panel that accumulates already has the earlier
	// panels' partial in y, so a middle panel has nothing to do for them, nor has the last panel of a product without epilogue
	const bool rows_keep_y = out.accum && (out.partial || (!out.scale && out.n_peer == 0 && !out.norm_partial));
	if (h->compressed && h->n_empty > 0 && !rows_keep_y) {
		const int64_t want = (h->n_empty + 255) / 256;
		zero_ctas = want < CSR_ZERO_CTAS ? want : CSR_ZERO_CTAS;
		csr_zero_rows_kernel<<<(unsigned) zero_ctas, 256, 0, st>>>(h->d_empty_rows, h->n_empty, out, h->ntiles + h->nsplit);
		CU_LAUNCH();
	}
	if (csr_launch_tiles(h, dx, st, out))
		return 1;
	if (h->nsplit > 0 && !h->fold_splits) {
		csr_fixup_kernel<<<(unsigned) h->nsplit, 32, 0, st>>>(h->d_rp_plan, h->d_split_row, h->d_split_tile, h->d_tile_head,
		                                                     h->d_tile_tail, h->nsplit, h->tile_nnz, h->ntiles,
		                                                     h->compressed ? h->d_rows_nz : nullptr, out);
		CU_LAUNCH();
	}
	if (l && l->norm2)
		return launch_norm_reduce(h, h->ntiles + h->nsplit + zero_ctas, l->norm2, st);
	return 0;
}

// Tiles [first, first + count) only -- the rows STARTING in them are final when the launch ends.  For handles whose rows are
// all finished by the tile kernel (no split rows, no row compression); no norm reduction: the caller (power.cu) composes
// several ranges -- of several column panels -- and reduces the norm partials once, after the last.
int launch_csr_tiles_range(b200spmv * h, const double * dx, double * dy, cudaStream_t st, const b200spmv_launch * l, int64_t first, int64_t count)
{
	if (h->format != B200SPMV_CSR_ADAPTIVE || h->npanels > 1 || h->nsplit > 0 || h->compressed)
		B200_FAIL("tile-range launches need a plain CSR handle without split rows / row compression");
	if (first < 0 || count < 0 || first + count > h->ntiles)
		B200_FAIL("tile range [%lld, +%lld) outside [0, %lld)", (long long) first, (long long) count, (long long) h->ntiles);
	YOut out;
	if (csr_make_out(h, dy, l, out))
		return 1;
	h->range_first = first;
	h->range_count = count;
	const int rc = csr_launch_tiles(h, dx, st, out);
	h->range_count = -1;
	return rc;
}

/*
 * spmv_oracle.c -- CPU restatement of the reference algorithms on the SpMV hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this library, and there only as the checker or as the timed CPU baseline.
 *
 * Parity status: PINNED.  Every function below is checked in tests/test_oracle_pin.py
 * against the reference's own code compiled from /root/reference into oracle/_ref/
 * (see oracle/Makefile) and against the committed fixtures in tests/golden/ that were
 * produced by those reference builds (tests/golden/make_golden.py).
 *
 * Reference paths are relative to /root/reference.  K/ = benchmark_code/BENCH/src/spmv_kernels/.
 */
#define _GNU_SOURCE
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ---------------------------------------------------------------------------------------
 * A2. serial CSR kernel -- K/csr.cpp:334-350 (subkernel_csr_scalar)
 * one running nnz index across rows, one accumulator, ascending j, every row stored.
 * ------------------------------------------------------------------------------------- */
void oracle_csr_spmv_scalar(const int32_t *row_ptr, const int32_t *col_ind, const double *val,
                            const double *x, double *y, int64_t i_s, int64_t i_e)
{
	int64_t j = row_ptr[i_s];
	for (int64_t i = i_s; i < i_e; i++) {
		int64_t j_e = row_ptr[i + 1];
		double sum = 0;
		for (; j < j_e; j++)
			sum += val[j] * x[col_ind[j]];
		y[i] = sum;
	}
}

/* The same kernel in the reference's `_f` build (make.sh:215-216: ValueType = float; the plugin's constructor narrows the
 * double matrix values with a C cast, K/csr.cpp:58-167): float products, one float accumulator, ascending j. */
void oracle_csr_spmv_scalar_f32(const int32_t *row_ptr, const int32_t *col_ind, const float *val,
                                const float *x, float *y, int64_t i_s, int64_t i_e)
{
	int64_t j = row_ptr[i_s];
	for (int64_t i = i_s; i < i_e; i++) {
		int64_t j_e = row_ptr[i + 1];
		float sum = 0;
		for (; j < j_e; j++)
			sum += val[j] * x[col_ind[j]];
		y[i] = sum;
	}
}

/* fp32 parity bound, per row: bound[i] = sum_j |a_ij * x_j| evaluated in double from the float operands. */
void oracle_row_abs_bound_f32(const int32_t *row_ptr, const int32_t *col_ind, const float *val,
                              const float *x, double *bound, int64_t m)
{
	#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < m; i++) {
		double s = 0;
		for (int64_t j = row_ptr[i]; j < row_ptr[i + 1]; j++)
			s += fabs((double)val[j] * (double)x[col_ind[j]]);
		bound[i] = s;
	}
}

/* The north-star per-row parity bound: bound[i] = sum_j |a_ij * x_j|  (BASELINE.md section 2). */
void oracle_row_abs_bound(const int32_t *row_ptr, const int32_t *col_ind, const double *val,
                          const double *x, double *bound, int64_t m)
{
	#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < m; i++) {
		double s = 0;
		for (int64_t j = row_ptr[i]; j < row_ptr[i + 1]; j++)
			s += fabs(val[j] * x[col_ind[j]]);
		bound[i] = s;
	}
}

/* ---------------------------------------------------------------------------------------
 * A1. nnz-balanced partitioner -- lib/parallel_util.h:156-184
 *     (loop_partitioner_balance_prefix_sums) + lib/macros/macrolib.h:506-581
 *     (macros_binary_search with the default comparator and distance).
 * Sums[N] is never read.  Ties in distance go to the upper index; equal values resolve
 * to the leftmost index holding the value.
 * ------------------------------------------------------------------------------------- */
static int64_t closest_index(const int32_t *A, int64_t lo, int64_t hi, int64_t target)
{
	int64_t s = lo, e = hi, mid;
	if (target < A[s])
		return s;
	if (target > A[e])
		return e;
	for (;;) {
		mid = (s + e) / 2;
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
	int64_t ds = llabs(target - (int64_t)A[s]);
	int64_t de = llabs(target - (int64_t)A[e]);
	return (ds < de) ? s : e;
}

void oracle_partition_prefix_sums(int64_t num_workers, int64_t worker_pos, const int32_t *sums,
                                  int64_t N, int64_t total, int64_t *start, int64_t *end)
{
	int64_t target = sums[0] + (total * worker_pos) / num_workers;
	int64_t target_next = sums[0] + (total * (worker_pos + 1)) / num_workers;
	*start = (worker_pos == 0) ? 0 : closest_index(sums, 0, N - 1, target);
	*end = (worker_pos == num_workers - 1) ? N : closest_index(sums, 0, N - 1, target_next);
}

/* CPU baseline "port": OpenMP CSR, rows split per thread by nnz exactly as the reference does
 * (K/csr.cpp:140 partition in the ctor, K/csr.cpp:381-404 compute_csr).  bounds has 2*T slots
 * and is filled on first use when bounds[1] < 0. */
void oracle_csr_spmv_omp(const int32_t *row_ptr, const int32_t *col_ind, const double *val,
                         const double *x, double *y, int64_t m, int64_t nnz,
                         int64_t *bounds, int num_threads)
{
	if (bounds[1] < 0) {
		for (int t = 0; t < num_threads; t++)
			oracle_partition_prefix_sums(num_threads, t, row_ptr, m, nnz, &bounds[2 * t], &bounds[2 * t + 1]);
	}
	#pragma omp parallel num_threads(num_threads)
	{
		#ifdef _OPENMP
		int t = omp_get_thread_num();
		#else
		int t = 0;
		#endif
		oracle_csr_spmv_scalar(row_ptr, col_ind, val, x, y, bounds[2 * t], bounds[2 * t + 1]);
	}
}

int oracle_max_threads(void)
{
	#ifdef _OPENMP
	return omp_get_max_threads();
	#else
	return 1;
	#endif
}

/* ---------------------------------------------------------------------------------------
 * A6. merge-path diagonal search -- K/merge.cpp:230-251 (MergePathSearch); list A is
 * row_end = row_ptr + 1 (length m), list B is the counting sequence 0..nnz-1.
 * ------------------------------------------------------------------------------------- */
void oracle_merge_path_search(int64_t diagonal, const int32_t *row_end, int64_t m, int64_t nnz,
                              int64_t *out_row, int64_t *out_nz)
{
	int64_t x_min = diagonal - nnz > 0 ? diagonal - nnz : 0;
	int64_t x_max = diagonal < m ? diagonal : m;
	while (x_min < x_max) {
		int64_t pivot = (x_min + x_max) >> 1;
		if (row_end[pivot] <= diagonal - pivot - 1)
			x_min = pivot + 1;
		else
			x_max = pivot;
	}
	*out_row = x_min < m ? x_min : m;
	*out_nz = diagonal - x_min;
}

/* Merge-based CsrMV with T workers -- K/merge.cpp:256-318 (OmpMergeCsrmv): items per worker
 * = ceil((m+nnz)/T), whole rows then the partial last row, serial carry-out fix-up.
 * coords (2*(T+1) int64) receives the (row, nz) start coordinate of every worker + the end. */
void oracle_merge_csrmv(int64_t T, const int32_t *row_ptr, const int32_t *col_ind, const double *val,
                        const double *x, double *y, int64_t m, int64_t nnz, int64_t *coords)
{
	const int32_t *row_end = row_ptr + 1;
	int64_t items = m + nnz;
	int64_t per = (items + T - 1) / T;
	int64_t *carry_row = (int64_t *)malloc(T * sizeof(int64_t));
	double *carry_val = (double *)malloc(T * sizeof(double));
	for (int64_t t = 0; t < T; t++) {
		int64_t d0 = per * t < items ? per * t : items;
		int64_t d1 = d0 + per < items ? d0 + per : items;
		int64_t r, z, r1, z1;
		oracle_merge_path_search(d0, row_end, m, nnz, &r, &z);
		oracle_merge_path_search(d1, row_end, m, nnz, &r1, &z1);
		if (coords) {
			coords[2 * t] = r;
			coords[2 * t + 1] = z;
			if (t == T - 1) {
				coords[2 * T] = r1;
				coords[2 * T + 1] = z1;
			}
		}
		for (; r < r1; ++r) {
			double tot = 0.0;
			for (; z < row_end[r]; ++z)
				tot += val[z] * x[col_ind[z]];
			y[r] = tot;
		}
		double tot = 0.0;
		for (; z < z1; ++z)
			tot += val[z] * x[col_ind[z]];
		carry_row[t] = r1;
		carry_val[t] = tot;
	}
	for (int64_t t = 0; t < T - 1; ++t)
		if (carry_row[t] < m)
			y[carry_row[t]] += carry_val[t];
	free(carry_row);
	free(carry_val);
}

/* ---------------------------------------------------------------------------------------
 * A3 + A4. SELL-C-sigma conversion -- K/sell-C-s/RISC-V/radix_sort.c:36-122,
 * sellcs_utils.c:36-48, sellcs_format.c:137-200 (csr_to_sellcs).
 * The reference's LSD radix passes amount to a STABLE sort by DESCENDING row length inside
 * each sigma window (ties keep ascending original row index); restated here as a stable
 * counting sort over the window's length range.  Outputs use 64-bit slice pointers
 * (the reference's -DINDEX64 build; its int32 build agrees after widening while it fits).
 * Caller allocates col_out / val_out with slice_ptr[nslices] entries after a sizing call
 * (col_out == NULL -> only row_order / slice_width / slice_ptr are produced).
 * ------------------------------------------------------------------------------------- */
int64_t oracle_sellcs_convert(const int32_t *row_ptr, const int32_t *col_ind, const double *val,
                              int64_t m, int64_t C, int64_t sigma,
                              int32_t *row_order, int32_t *slice_width, int64_t *slice_ptr,
                              int32_t *col_out, double *val_out)
{
	int64_t nslices = (m + C - 1) / C;
	int64_t *len_sorted = (int64_t *)malloc((m > 0 ? m : 1) * sizeof(int64_t));
	for (int64_t k = 0; k < m; k += sigma) {
		int64_t k_end = k + sigma > m ? m : k + sigma;
		int64_t maxlen = 0;
		for (int64_t i = k; i < k_end; i++) {
			int64_t l = row_ptr[i + 1] - row_ptr[i];
			if (l > maxlen)
				maxlen = l;
		}
		int64_t *cnt = (int64_t *)calloc(maxlen + 2, sizeof(int64_t));
		for (int64_t i = k; i < k_end; i++)
			cnt[row_ptr[i + 1] - row_ptr[i]]++;
		/* descending: position of first element of length l = number of rows longer than l */
		int64_t run = 0;
		for (int64_t l = maxlen; l >= 0; l--) {
			int64_t c = cnt[l];
			cnt[l] = run;
			run += c;
		}
		for (int64_t i = k; i < k_end; i++) {
			int64_t l = row_ptr[i + 1] - row_ptr[i];
			int64_t pos = k + cnt[l]++;
			row_order[pos] = (int32_t)i;
			len_sorted[pos] = l;
		}
		free(cnt);
	}
	int64_t total = 0;
	for (int64_t s = 0; s < nslices; s++) {
		int64_t w = len_sorted[s * C];
		slice_ptr[s] = total;
		slice_width[s] = (int32_t)w;
		total += w * C;
	}
	slice_ptr[nslices] = total;
	free(len_sorted);
	if (col_out == NULL)
		return total;
	memset(col_out, 0, total * sizeof(int32_t));
	memset(val_out, 0, total * sizeof(double));
	for (int64_t r = 0; r < m; r++) {
		int64_t w = slice_ptr[r / C] + (r % C);
		int64_t orig = row_order[r];
		for (int64_t j = row_ptr[orig]; j < row_ptr[orig + 1]; j++) {
			val_out[w] = val[j];
			col_out[w] = col_ind[j];
			w += C;
		}
	}
	return total;
}

/* a7. SELL-C-sigma SpMV semantics -- K/sell-C-s/RISC-V/sellcs_mv_kernels_epi.c:178-257:
 * per slice, lane-wise fused multiply-accumulate down the slice width, then scatter through
 * row_order; the ragged last slice uses min(C, m - s*C) lanes.  (The reference kernel is
 * RISC-V vector intrinsics; vfmacc is a fused multiply-add, hence fma() here.) */
void oracle_sellcs_spmv(int64_t m, int64_t C, const int32_t *row_order, const int32_t *slice_width,
                        const int64_t *slice_ptr, const int32_t *col, const double *val,
                        const double *x, double *y)
{
	int64_t nslices = (m + C - 1) / C;
	#pragma omp parallel for schedule(static)
	for (int64_t s = 0; s < nslices; s++) {
		int64_t row0 = s * C;
		int64_t lanes = row0 + C > m ? m - row0 : C;
		for (int64_t l = 0; l < lanes; l++) {
			double acc = 0.0;
			int64_t p = slice_ptr[s] + l;
			for (int64_t k = 0; k < slice_width[s]; k++, p += C)
				acc = fma(x[col[p]], val[p], acc);
			y[row_order[row0 + l]] = acc;
		}
	}
}

/* ---------------------------------------------------------------------------------------
 * a11. gold vector + pass criterion -- benchmark_code/BENCH/src/bench_spmv.cpp:108-235
 * (check_accuracy): per-row Kahan sum in _Float128; max relative difference over rows with
 * y_gold > eps (signed test, as in the reference); plus mae / max_ae / mse.
 * out[0]=maxRelDiff out[1]=mae out[2]=max_ae out[3]=mse ; gold (may be NULL) gets (double)y_gold.
 * ------------------------------------------------------------------------------------- */
void oracle_check_accuracy(const int32_t *row_ptr, const int32_t *col_ind, const double *val,
                           const double *x, const double *y, int64_t m, double eps,
                           double *gold, double *out)
{
	_Float128 *yg = (_Float128 *)malloc((m > 0 ? m : 1) * sizeof(_Float128));
	#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < m; i++) {
		_Float128 sum = 0, comp = 0, v, tmp;
		for (int64_t j = row_ptr[i]; j < row_ptr[i + 1]; j++) {
			v = (_Float128)val[j] * (_Float128)x[col_ind[j]] - comp;
			tmp = sum + v;
			comp = (tmp - sum) - v;
			sum = tmp;
		}
		yg[i] = sum;
	}
	_Float128 maxrel = 0, sae = 0, sse = 0, maxae = 0;
	for (int64_t i = 0; i < m; i++) {
		_Float128 d = yg[i] - (_Float128)y[i];
		if (d < 0)
			d = -d;
		sae += d;
		sse += d * d;
		if (d > maxae)
			maxae = d;
		if (yg[i] > (_Float128)eps) {
			_Float128 r = d / yg[i];
			if (r > maxrel)
				maxrel = r;
		}
		if (gold)
			gold[i] = (double)yg[i];
	}
	out[0] = (double)maxrel;
	out[1] = m ? (double)(sae / m) : 0;
	out[2] = (double)maxae;
	out[3] = m ? (double)(sse / m) : 0;
	free(yg);
}

/* ---------------------------------------------------------------------------------------
 * a10. CSR-Adaptive greedy row blocking (behavioural reference for nnz-tiled binning) --
 * K/GPU_clean/csr_adaptive_cuda.cu:46-79: blocks of consecutive rows holding at most
 * block_nnz non-zeros, a lone over-long row gets a block of its own, list padded with m to a
 * multiple of multiblock and terminated with m.  row_blocks needs m + multiblock + 1 slots.
 * ------------------------------------------------------------------------------------- */
int64_t oracle_adaptive_rowblocks(const int32_t *row_ptr, int64_t m, int64_t block_nnz,
                                  int64_t multiblock, int32_t *row_blocks)
{
	int64_t cnt = 1, sum = 0, last_i = 0;
	row_blocks[0] = 0;
	for (int64_t i = 1; i < m; i++) {
		sum += row_ptr[i] - row_ptr[i - 1];
		if (sum == block_nnz) {
			last_i = i;
			row_blocks[cnt++] = (int32_t)i;
			sum = 0;
		} else if (sum > block_nnz) {
			if (i - last_i > 1) {
				row_blocks[cnt++] = (int32_t)(i - 1);
				i--;
			} else if (i - last_i == 1) {
				row_blocks[cnt++] = (int32_t)i;
			}
			last_i = i;
			sum = 0;
		}
	}
	while (cnt % multiblock != 0)
		row_blocks[cnt++] = (int32_t)m;
	row_blocks[cnt++] = (int32_t)m;
	return cnt;
}

/* FNV-1a-64 over raw bytes: the fingerprint used for the generator pins in SURVEY.md App. C. */
uint64_t oracle_fnv1a64(const void *data, int64_t nbytes)
{
	const unsigned char *p = (const unsigned char *)data;
	uint64_t h = 0xcbf29ce484222325ULL;
	for (int64_t i = 0; i < nbytes; i++) {
		h ^= p[i];
		h *= 0x100000001b3ULL;
	}
	return h;
}

/*
 * Jacobi-preconditioned CG, serial restatement of preconditioned_cg()
 * (benchmark_code/BENCH/src/bench_cg.cpp:93-320): x0 = 0, K = diag(A) (first entry with j == i, :121-131),
 * eps = 1e-15 * ||b||, eps_counter = 1e-7 * ||b|| (:159-175), every restart_k = 100 iterations the explicit residual
 * b - A x is evaluated, x_best updated, and CG restarted from it when err_explicit / err > 1e3 while err_best >
 * eps_counter (:186-236), stop when ||r|| < eps (:238) or k == max_iterations; afterwards one more explicit
 * residual and x_best update, x_best is the answer (:290-320).  out[0..4] = iterations, err, err_explicit, err_best, eps.
 * Returns 0, or 1 when a diagonal entry is missing / zero ("bad K, zero in diagonal").
 */
static double o_dot(const double *a, const double *b, int64_t n)
{
	double s = 0;
	for (int64_t i = 0; i < n; i++)
		s += a[i] * b[i];
	return s;
}

static void o_spmv(const int32_t *rp, const int32_t *ci, const double *v, const double *x, double *y, int64_t m)
{
	for (int64_t i = 0; i < m; i++) {
		double s = 0;
		for (int64_t j = rp[i]; j < rp[i + 1]; j++)
			s += v[j] * x[ci[j]];
		y[i] = s;
	}
}

int oracle_pcg(const int32_t *rp, const int32_t *ci, const double *v, int64_t m, const double *b, double *x_out,
               int64_t max_iterations, double *out)
{
	double *mem = (double *)calloc((size_t)(8 * m + 8), sizeof(double));
	if (!mem)
		return 2;
	double *rk = mem, *rke = mem + m, *pk = mem + 2 * m, *zk = mem + 3 * m, *Ap = mem + 4 * m, *K = mem + 5 * m, *x = mem + 6 * m,
	       *xb = mem + 7 * m;
	for (int64_t i = 0; i < m; i++) {
		K[i] = 0;
		for (int64_t j = rp[i]; j < rp[i + 1]; j++)
			if (ci[j] == i) {
				K[i] = v[j];
				break;
			}
		if (K[i] == 0) {
			free(mem);
			return 1;
		}
	}
	o_spmv(rp, ci, v, x, Ap, m);
	for (int64_t i = 0; i < m; i++) {
		rk[i] = b[i] + -1 * Ap[i];
		zk[i] = rk[i] / K[i];
		pk[i] = zk[i];
	}
	double err = sqrt(o_dot(rk, rk, m)), bn = sqrt(o_dot(b, b, m));
	const double eps = 1.0e-15 * bn, eps_counter = 1.0e-7 * bn;
	double err_explicit = err, err_best = err;
	int64_t k = 0;
	const int64_t restart_k = 100;
	while (k < max_iterations) {
		const int check = (k > 0) && !(k % restart_k);
		if (check) {
			o_spmv(rp, ci, v, x, Ap, m);
			for (int64_t i = 0; i < m; i++)
				rke[i] = b[i] + -1 * Ap[i];
			err_explicit = sqrt(o_dot(rke, rke, m));
			if (err_explicit < err_best) {
				memcpy(xb, x, (size_t)m * sizeof(double));
				err_best = err_explicit;
			}
		}
		err = sqrt(o_dot(rk, rk, m));
		if (check && (err_best > eps_counter) && (err_explicit / err > 1e3)) {
			memcpy(rk, rke, (size_t)m * sizeof(double));
			for (int64_t i = 0; i < m; i++)
				zk[i] = rk[i] / K[i];
			memcpy(pk, zk, (size_t)m * sizeof(double));
		}
		if (err < eps)
			break;
		o_spmv(rp, ci, v, pk, Ap, m);
		const double old_zr = o_dot(zk, rk, m);
		const double ak = o_dot(zk, rk, m) / o_dot(pk, Ap, m);
		for (int64_t i = 0; i < m; i++)
			x[i] = x[i] + ak * pk[i];
		for (int64_t i = 0; i < m; i++)
			rk[i] = rk[i] + -ak * Ap[i];
		for (int64_t i = 0; i < m; i++)
			zk[i] = rk[i] / K[i];
		const double bk = o_dot(zk, rk, m) / old_zr;
		for (int64_t i = 0; i < m; i++)
			pk[i] = zk[i] + bk * pk[i];
		k++;
	}
	o_spmv(rp, ci, v, x, Ap, m);
	for (int64_t i = 0; i < m; i++)
		rke[i] = b[i] + -1 * Ap[i];
	err_explicit = sqrt(o_dot(rke, rke, m));
	if (err_explicit < err_best) {
		memcpy(xb, x, (size_t)m * sizeof(double));
		err_best = err_explicit;
	}
	memcpy(x_out, xb, (size_t)m * sizeof(double));
	out[0] = (double)k;
	out[1] = err;
	out[2] = err_explicit;
	out[3] = err_best;
	out[4] = eps;
	free(mem);
	return 0;
}

/* Jacobi-preconditioned BiCGSTAB, serial restatement of preconditioned_bicgstab()
 * (benchmark_code/BENCH/src/bench_bicg.cpp:150-456).  Like the reference it has NO stop test (the "if (err < eps) break" is
 * commented out at :333-334): it runs max_iterations iterations, keeps x_best = the iterate with the smallest explicit
 * residual seen at the checkpoints (every 100 iterations, :307-331) and at the end (:427-446), and returns x_best.
 * trace (nullable, max_iterations doubles) receives ||r_k|| at the top of iteration k (the "k = .. error = .." line, :343).
 * out[0] iterations, out[1] ||r|| at the top of the last iteration, out[2] last explicit residual, out[3] err_best, out[4] eps.
 * Returns 1 on "bad K, zero in diagonal" (:209-210). */
int oracle_bicgstab(const int32_t *rp, const int32_t *ci, const double *v, int64_t m, const double *b, double *x_out,
                    int64_t max_iterations, double *trace, double *out)
{
	double *mem = (double *)calloc((size_t)(12 * m + 8), sizeof(double));
	if (!mem)
		return 2;
	double *r0 = mem, *rk = mem + m, *rke = mem + 2 * m, *pk = mem + 3 * m, *z = mem + 4 * m, *h = mem + 5 * m, *s = mem + 6 * m,
	       *vv = mem + 7 * m, *buf = mem + 8 * m, *K = mem + 9 * m, *xk = mem + 10 * m, *xb = mem + 11 * m;
	for (int64_t i = 0; i < m; i++) {
		K[i] = 0;
		for (int64_t j = rp[i]; j < rp[i + 1]; j++)
			if (ci[j] == i) {
				K[i] = v[j];
				break;
			}
		if (K[i] == 0) {
			free(mem);
			return 1;
		}
	}
	o_spmv(rp, ci, v, xk, buf, m);                              /* :218-250 */
	for (int64_t i = 0; i < m; i++) {
		rk[i] = b[i] + -1 * buf[i];
		r0[i] = rk[i];
		pk[i] = rk[i];
	}
	double s_pk_p = o_dot(r0, rk, m), s_a = 0;
	double err = sqrt(o_dot(rk, rk, m));
	const double bn = sqrt(o_dot(b, b, m));
	const double eps = 1.0e-15 * bn;
	double err_explicit = err, err_best = err;
	int64_t k = 0;
	const int64_t restart_k = 100;
	while (k < max_iterations) {
		if ((k > 0) && !(k % restart_k)) {                      /* :307-331 */
			o_spmv(rp, ci, v, xk, buf, m);
			for (int64_t i = 0; i < m; i++)
				rke[i] = b[i] + -1 * buf[i];
			err_explicit = sqrt(o_dot(rke, rke, m));
			if (err_explicit < err_best) {
				memcpy(xb, xk, (size_t)m * sizeof(double));
				err_best = err_explicit;
			}
		}
		err = sqrt(o_dot(rk, rk, m));                           /* :333-341 */
		if (trace)
			trace[k] = err;
		double *y = buf;                                        /* y = pk / K, v = A y   (:349-361) */
		for (int64_t i = 0; i < m; i++)
			y[i] = pk[i] / K[i];
		o_spmv(rp, ci, v, y, vv, m);
		s_a = s_pk_p / o_dot(r0, vv, m);                        /* :368-386 */
		for (int64_t i = 0; i < m; i++)
			h[i] = xk[i] + s_a * y[i];
		for (int64_t i = 0; i < m; i++)
			s[i] = rk[i] + -s_a * vv[i];
		for (int64_t i = 0; i < m; i++)
			z[i] = s[i] / K[i];
		double *t = buf;                                        /* t = A z   (:389-392) */
		o_spmv(rp, ci, v, z, t, m);
		double p1 = 0, p2 = 0;                                  /* :399-416 */
		for (int64_t i = 0; i < m; i++) {
			const double v1 = t[i] / K[i], v2 = s[i] / K[i];
			p1 += v1 * v2;
			p2 += v1 * v1;
		}
		const double s_w = p1 / p2;
		for (int64_t i = 0; i < m; i++)                         /* :418-422 */
			rk[i] = s[i] + -s_w * t[i];
		for (int64_t i = 0; i < m; i++)
			xk[i] = h[i] + s_w * z[i];
		const double s_pk = o_dot(r0, rk, m);                   /* :426-437 */
		const double s_b = (s_pk / s_pk_p) * (s_a / s_w);
		for (int64_t i = 0; i < m; i++)
			pk[i] = rk[i] + s_b * (pk[i] - s_w * vv[i]);
		s_pk_p = s_pk;
		k++;
	}
	o_spmv(rp, ci, v, xk, buf, m);                              /* :427-446 */
	for (int64_t i = 0; i < m; i++)
		rke[i] = b[i] + -1 * buf[i];
	err_explicit = sqrt(o_dot(rke, rke, m));
	if (err_explicit < err_best) {
		memcpy(xb, xk, (size_t)m * sizeof(double));
		err_best = err_explicit;
	}
	memcpy(x_out, xb, (size_t)m * sizeof(double));
	out[0] = (double)k;
	out[1] = err;
	out[2] = err_explicit;
	out[3] = err_best;
	out[4] = eps;
	free(mem);
	return 0;
}

/* Plain SELL of the reference (benchmark_code/BENCH/src/spmv_kernels/sell.cpp:98-192): slices
 * ("row clusters") of VL consecutive rows in their original order (VL = vec_len_default_f64: 4 with AVX2, 8 with AVX-512),
 * each padded to its longest row and stored column-major (entry k of the slice's rows side by side).  What is peculiar:
 *   - row_cluster_ptr is already multiplied by VL (:137-143);
 *   - a padding entry has value 0 and the LAST REAL COLUMN the converting thread has seen (0 at the start of the thread's
 *     range, :158-185), so the layout depends on the number of threads: the slices are dealt to `nthreads` workers with
 *     loop_partitioner_balance_prefix_sums over row_cluster_ptr (:155);
 *   - the product accumulates with fused multiply-adds in ascending k and stores all VL lanes of the last slice, i.e.
 *     y[m .. S*VL) is written too (:232-273; the harness allocates m + 64).
 * Pass 1 (ja == NULL): fills rcp[S + 1] and returns nnz_ext.  Pass 2: fills ja / a. */
static int64_t sell_plain_widths(const int32_t *rp, int64_t m, int64_t VL, int32_t *rcp)
{
	const int64_t S = (m + VL - 1) / VL;
	int64_t run = 0;
	for (int64_t s = 0; s < S; s++) {
		int64_t width = 0;
		const int64_t k_e = (s + 1) * VL < m ? (s + 1) * VL : m;
		for (int64_t k = s * VL; k < k_e; k++)
			if (rp[k + 1] - rp[k] > width)
				width = rp[k + 1] - rp[k];
		rcp[s] = (int32_t)run;
		run += VL * width;
	}
	rcp[S] = (int32_t)run;
	return run;
}

/* one converting thread's slices [ii_s, ii_e): sell.cpp:158-188 / sell_sorted.cpp:262-290 */
static void sell_plain_fill(const int32_t *rp, const int32_t *ci, const double *v, int64_t m, int64_t VL, const int32_t *rcp,
                            int64_t ii_s, int64_t ii_e, int32_t *ja, double *a)
{
	int64_t col = 0;
	for (int64_t ii = ii_s; ii < ii_e; ii++) {
		const int64_t width = (rcp[ii + 1] - rcp[ii]) / VL;
		const int64_t i_s = VL * ii;
		for (int64_t lane = 0; lane < VL; lane++) {
			const int64_t i = i_s + lane;
			int64_t k = 0;
			if (i < m)
				for (int64_t j = rp[i]; j < rp[i + 1]; j++, k++) {
					col = ci[j];
					a[rcp[ii] + k * VL + lane] = v[j];
					ja[rcp[ii] + k * VL + lane] = (int32_t)col;
				}
			for (; k < width; k++) {        /* padding of shorter rows, and of the rows missing in the last slice */
				a[rcp[ii] + k * VL + lane] = 0;
				ja[rcp[ii] + k * VL + lane] = (int32_t)col;
			}
		}
	}
}

int64_t oracle_sell_plain_convert(const int32_t *rp, const int32_t *ci, const double *v, int64_t m, int64_t VL, int64_t nthreads,
                                  int32_t *rcp, int32_t *ja, double *a)
{
	const int64_t S = (m + VL - 1) / VL;
	const int64_t run = sell_plain_widths(rp, m, VL, rcp);
	if (!ja)
		return run;
	for (int64_t t = 0; t < nthreads; t++) {
		int64_t ii_s, ii_e;
		oracle_partition_prefix_sums(nthreads, t, rcp, S, run, &ii_s, &ii_e);
		sell_plain_fill(rp, ci, v, m, VL, rcp, ii_s, ii_e, ja, a);
	}
	return run;
}

/* Per-thread-sorted SELL of the reference (K/sell_sorted.cpp:125-298): the slices are dealt to the threads by nnz on the
 * ORIGINAL row_ptr (:158-175), each thread sorts the rows of its range by ASCENDING length with a stable bucket sort
 * (:186-191; permutation[i] = new position of row i, rev_permutation its inverse), the matrix is rewritten in that order and
 * then stored exactly like the plain SELL, with the same thread ranges (:262-290).  The product writes
 * y[rev_permutation[i]] (:392-395), so per row it is the plain SELL's chain of fused multiply-adds.
 * Pass 1 (ja == NULL): permutation / rev_permutation / rcp, returns nnz_ext.  Pass 2: ja / a as well. */
int64_t oracle_sell_sorted_convert(const int32_t *rp, const int32_t *ci, const double *v, int64_t m, int64_t VL, int64_t nthreads,
                                   int32_t *permutation, int32_t *rev_permutation, int32_t *rcp, int32_t *ja, double *a)
{
	const int64_t S = (m + VL - 1) / VL, nnz = rp[m];
	int32_t *rcp0 = (int32_t *)malloc((size_t)(S + 1) * sizeof(int32_t));
	int32_t *rp2 = (int32_t *)malloc((size_t)(m + 1) * sizeof(int32_t));
	int32_t *ci2 = (int32_t *)malloc((size_t)(nnz > 0 ? nnz : 1) * sizeof(int32_t));
	double *v2 = (double *)malloc((size_t)(nnz > 0 ? nnz : 1) * sizeof(double));
	int64_t *bounds = (int64_t *)malloc((size_t)(2 * nthreads) * sizeof(int64_t));
	for (int64_t s = 0; s < S; s++)
		rcp0[s] = rp[s * VL];
	rcp0[S] = rp[m];
	for (int64_t t = 0; t < nthreads; t++) {
		int64_t ii_s, ii_e;
		oracle_partition_prefix_sums(nthreads, t, rcp0, S, nnz, &ii_s, &ii_e);
		bounds[2 * t] = ii_s;
		bounds[2 * t + 1] = ii_e;
		const int64_t i_s = ii_s * VL, i_e = (t == nthreads - 1) ? m : ii_e * VL;
		int64_t dmax = 0;
		for (int64_t i = i_s; i < i_e; i++)
			if (rp[i + 1] - rp[i] > dmax)
				dmax = rp[i + 1] - rp[i];
		int64_t *off = (int64_t *)calloc((size_t)(dmax + 2), sizeof(int64_t));   /* stable counting sort, ascending length */
		for (int64_t i = i_s; i < i_e; i++)
			off[rp[i + 1] - rp[i] + 1]++;
		for (int64_t d = 0; d <= dmax; d++)
			off[d + 1] += off[d];
		for (int64_t i = i_s; i < i_e; i++) {
			permutation[i] = (int32_t)(i_s + off[rp[i + 1] - rp[i]]++);
			rev_permutation[permutation[i]] = (int32_t)i;
		}
		free(off);
	}
	rp2[0] = 0;
	for (int64_t i = 0; i < m; i++) {
		const int64_t r = rev_permutation[i];
		rp2[i + 1] = rp2[i] + (rp[r + 1] - rp[r]);
		for (int64_t j = rp[r], k = rp2[i]; j < rp[r + 1]; j++, k++) {
			ci2[k] = ci[j];
			v2[k] = v[j];
		}
	}
	const int64_t run = sell_plain_widths(rp2, m, VL, rcp);
	if (ja)
		for (int64_t t = 0; t < nthreads; t++)
			sell_plain_fill(rp2, ci2, v2, m, VL, rcp, bounds[2 * t], bounds[2 * t + 1], ja, a);
	free(rcp0);
	free(rp2);
	free(ci2);
	free(v2);
	free(bounds);
	return run;
}

/* y must hold S * VL entries (sell.cpp:266 stores whole vectors). */
void oracle_sell_plain_spmv(int64_t m, int64_t VL, const int32_t *rcp, const int32_t *ja, const double *a, const double *x, double *y)
{
	const int64_t S = (m + VL - 1) / VL;
	for (int64_t ii = 0; ii < S; ii++)
		for (int64_t lane = 0; lane < VL; lane++) {
			double sum = 0;
			for (int64_t jj = rcp[ii]; jj < rcp[ii + 1]; jj += VL)
				sum = fma(a[jj + lane], x[ja[jj + lane]], sum);
			y[VL * ii + lane] = sum;
		}
}

/* COO -> CSR as the reference's import_file() does it (bench.cpp:208-229 -> coo_to_csr, lib/storage_formats/csr/csr_gen.c:
 * 178-213): a stable bucket sort of the entries by row (row_ptr = the bucket offsets), then the columns of each row
 * sorted ascending (csr_sort_columns).  For a matrix without duplicate (row, col) pairs the result is simply the entries
 * in (row, col) order.  The reference orders duplicates with an unstable quicksort; here they keep their input order
 * (stable), and the return value says how many there are so that callers can refuse them.
 * row_ptr[m + 1], col_out[nnz], val_out[nnz].  Returns the number of duplicate pairs, or -1 on an index out of range. */
int64_t oracle_coo_to_csr(const int32_t *R, const int32_t *Cc, const double *V, int64_t m, int64_t n, int64_t nnz,
                          int32_t *row_ptr, int32_t *col_out, double *val_out)
{
	int64_t *cnt = (int64_t *)calloc((size_t)(m + 1), sizeof(int64_t));
	int64_t *perm = (int64_t *)malloc((size_t)(nnz > 0 ? nnz : 1) * sizeof(int64_t));
	for (int64_t e = 0; e < nnz; e++) {
		if (R[e] < 0 || R[e] >= m || Cc[e] < 0 || Cc[e] >= n) {
			free(cnt);
			free(perm);
			return -1;
		}
		cnt[R[e] + 1]++;
	}
	for (int64_t i = 0; i < m; i++)
		cnt[i + 1] += cnt[i];
	for (int64_t i = 0; i <= m; i++)
		row_ptr[i] = (int32_t)cnt[i];
	for (int64_t e = 0; e < nnz; e++)
		perm[cnt[R[e]]++] = e;                      /* stable: entries of a row in input order */
	int64_t dups = 0;
	for (int64_t i = 0; i < m; i++) {
		const int64_t a = row_ptr[i], b = row_ptr[i + 1];
		for (int64_t j = a + 1; j < b; j++) {       /* stable insertion sort by column (rows are short in tests) */
			const int64_t p = perm[j];
			int64_t k = j;
			while (k > a && Cc[perm[k - 1]] > Cc[p]) {
				perm[k] = perm[k - 1];
				k--;
			}
			perm[k] = p;
		}
		for (int64_t j = a; j < b; j++) {
			col_out[j] = Cc[perm[j]];
			val_out[j] = V[perm[j]];
			if (j > a && col_out[j] == col_out[j - 1])
				dups++;
		}
	}
	free(cnt);
	free(perm);
	return dups;
}

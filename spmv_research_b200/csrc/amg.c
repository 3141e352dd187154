/*
 * amg.c -- host-side synthetic CSR matrix source for the B200 SpMV engine.
 *
 * A from-scratch restatement of the reference's artificial matrix generator, producing the
 * SAME row_ptr / col_ind / values arrays bit for bit:
 *   reference: artificial-matrix-generator/artificial_matrix_generation.c:163-523 (generator),
 *              :78-150 (bandwidth correction), random.h:127-308 (uniform / Box-Muller normal /
 *              Marsaglia-Tsang gamma over glibc random_r), ordered_set.c (a set + sorted dump).
 * What is different by design:
 *   - the glibc TYPE_4 additive-feedback generator is implemented here (no random_r call), so
 *     a row segment needs nothing but its seed -- the property a device generator needs;
 *   - rows are produced per "anchor segment" (the reference reseeds with seed+i every
 *     anchors_period rows, :259-288), so any row range [i_s, i_e) can be generated on its own:
 *     each rank of a row-sharded run builds only its shard (b200amg_fill);
 *   - duplicate detection is an open-addressing hash set + sort instead of an AVL tree
 *     (the tree's shape never influences the output).
 * Must be compiled WITHOUT -ffast-math and with -ffp-contract=off: the output depends on the
 * exact double rounding sequence (and on glibc's libm exp/log/cos/sqrt/pow/log2).
 */
#define _GNU_SOURCE
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <float.h>
#include <stdio.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../../include/b200amg.h"

#define RANDOM_MAX    0x7FFFFFFFL
#define RANDOM_MAX_64 0x3FFFFFFFFFFFFFFFL

static __thread char amg_err[256];
const char *b200amg_last_error(void) { return amg_err; }

/* ---- glibc random_r, TYPE_4 (256-byte state: degree 63, separation 1) ---- */
#define R_DEG 63
#define R_SEP 1
typedef struct {
	int32_t st[R_DEG];
	int f, r;
} rng_t;

static inline int32_t rng_next(rng_t *g)
{
	uint32_t v = (uint32_t)g->st[g->f] + (uint32_t)g->st[g->r];
	g->st[g->f] = (int32_t)v;
	if (++g->f >= R_DEG) {
		g->f = 0;
		++g->r;
	} else if (++g->r >= R_DEG) {
		g->r = 0;
	}
	return (int32_t)(v >> 1);
}

static void rng_seed(rng_t *g, unsigned int seed)
{
	int32_t word = seed ? (int32_t)seed : 1;
	g->st[0] = word;
	for (int i = 1; i < R_DEG; i++) {
		long hi = word / 127773, lo = word % 127773;
		long w = 16807 * lo - 2836 * hi;
		if (w < 0)
			w += 2147483647;
		word = (int32_t)w;
		g->st[i] = word;
	}
	g->f = R_SEP;
	g->r = 0;
	for (int k = 0; k < R_DEG * 10; k++)
		(void)rng_next(g);
}

/* random.h:227-237 */
static inline double rnd_uniform(rng_t *g, double min, double max)
{
	double res = ((double)rng_next(g)) / ((double)(RANDOM_MAX + 1L));
	res = (max - min) * res + min;
	return res;
}

/* random.h:177-222 : modulo-bias-corrected integers over [min, max) */
static inline long rnd_uniform_integer(rng_t *g, long min, long max)
{
	int64_t range = max - min;
	if (range >= RANDOM_MAX) {
		int64_t lim = RANDOM_MAX_64 - RANDOM_MAX_64 % range, res;
		do {
			res = rng_next(g);
			res <<= 31;
			res |= rng_next(g);
		} while (res >= lim);
		return (long)(res % range + min);
	}
	int64_t lim = RANDOM_MAX - RANDOM_MAX % range;
	int32_t v;
	do {
		v = rng_next(g);
	} while (v >= lim);
	return (long)(v % range + min);
}

/* random.h:243-255 : Box-Muller, cosine branch only */
static inline double rnd_normal(rng_t *g, double mean, double std)
{
	double u1 = rnd_uniform(g, 2 * DBL_EPSILON, 1);
	double u2 = rnd_uniform(g, 0, 1);
	double magnitude = std * sqrt(-2.0 * log(u1));
	return magnitude * cos(2.0 * M_PI * u2) + mean;
}

/* random.h:271-296 : Marsaglia-Tsang */
static double rnd_gamma(rng_t *g, double k, double theta)
{
	if (k < 1) {
		double u = rnd_uniform(g, 2 * DBL_EPSILON, 1);
		return rnd_gamma(g, 1.0 + k, theta) * pow(u, 1.0 / k);
	}
	double x, v, u;
	double d = k - 1.0 / 3.0;
	double c = (1.0 / 3.0) / sqrt(d);
	for (;;) {
		do {
			x = rnd_normal(g, 0, 1);
			v = 1.0 + c * x;
		} while (v <= 0);
		v = v * v * v;
		u = rnd_uniform(g, 2 * DBL_EPSILON, 1);
		if (u < 1 - 0.0331 * x * x * x * x)
			break;
		if (log(u) < 0.5 * x * x + d * (1 - v + log(v)))
			break;
	}
	return theta * d * v;
}

/* ---- bandwidth correction: artificial_matrix_generation.c:105-150 ---- */
static double expected_bw(double B, double n)
{
	double E, a, b, b_e, x;
	long i;
	b_e = 0;
	for (i = 0; i < (long)n; i++)
		b_e += log2(n - i) - log2(B - i);
	b = pow(2, b_e);
	E = 0;
	for (i = n; i <= (long)B; i++) {
		x = i;
		a = (B - x + 1) * n / x * (n - 1) / (x - 1);
		E += a * b * (double)i;
		b_e += log2(x + 1) - log2(x + 1 - n);
		b = pow(2, b_e);
	}
	return E;
}

static double corrected_bw(double B, double n)
{
	double E;
	n = floor(n + 0.5);
	if (n < 2)
		n = 2;
	E = expected_bw(B, n);
	if (E == 0)
		return n;
	return B * (B / E);
}

/* ---- derived model parameters: artificial_matrix_generation.c:206-237 ---- */
typedef struct {
	double MAX, C, avg_distr, std_distr;
	long period;
} model_t;

static int model_init(const b200amg_params *p, model_t *M)
{
	if (p->nr_rows < 100 || p->nr_cols < 1) {
		snprintf(amg_err, sizeof(amg_err), "generator needs nr_rows >= 100 (reference divides by nr_rows/100) and nr_cols >= 1");
		return 1;
	}
	double avg = p->avg_nnz_per_row, std = p->std_nnz_per_row;
	double MAX = avg * (1 + p->skew);
	double c1 = MAX / avg;
	double c2 = (std > 0) ? MAX * MAX / (2 * std * std) : 0;
	double C = (c1 > c2) ? c1 : c2;
	if (C <= 2 || p->skew < 1) {
		M->MAX = 0;
		M->C = 0;
		M->avg_distr = avg;
		M->std_distr = std;
	} else {
		double avg_exp = MAX / C;
		double std_exp = sqrt(avg_exp * avg_exp * (C / 2 - 1));
		M->MAX = MAX;
		M->C = C;
		M->avg_distr = avg - MAX / C;
		M->std_distr = sqrt(std * std - std_exp * std_exp);
		if (p->distribution == 'n')
			if (M->std_distr > M->avg_distr / 3)
				M->std_distr = M->avg_distr / 3;
	}
	long anchors = p->nr_rows / 100;
	if (anchors > (1 << 12))
		anchors = 1 << 12;
	M->period = p->nr_rows / anchors;
	return 0;
}

static inline long row_degree(const b200amg_params *p, const model_t *M, rng_t *g, long i)
{
	double degree_exp = M->MAX * exp(-M->C / p->nr_rows * (double)i);
	double degree_distr;
	if (M->std_distr > 0) {
		if (p->distribution == 'n')
			degree_distr = rnd_normal(g, M->avg_distr, M->std_distr);
		else {
			double k = (M->avg_distr * M->avg_distr) / (M->std_distr * M->std_distr);
			double theta = (M->avg_distr > 0) ? (M->std_distr * M->std_distr) / M->avg_distr : 1;
			degree_distr = rnd_gamma(g, k, theta);
		}
	} else
		degree_distr = M->avg_distr;
	double degree = degree_exp + degree_distr;
	degree = (degree > 0) ? floor(degree + 0.5) : 0;
	if (degree > p->nr_cols)
		degree = p->nr_cols;
	return (long)degree;
}

/*
 * Pass 1 (artificial_matrix_generation.c:279-354): row lengths -> row_ptr (exclusive prefix).
 * Returns nnz (>= 0) or -1 on error.  bw_corrected_out receives the corrected band width that
 * pass 2 needs (:348).
 */
int64_t b200amg_row_ptr(const b200amg_params *p, int32_t *row_ptr, double *bw_corrected_out)
{
	model_t M;
	if (model_init(p, &M))
		return -1;
	long m = p->nr_rows;
	long nseg = (m + M.period - 1) / M.period;
	#pragma omp parallel for schedule(dynamic, 4)
	for (long s = 0; s < nseg; s++) {
		rng_t g;
		long i0 = s * M.period, i1 = i0 + M.period > m ? m : i0 + M.period;
		rng_seed(&g, p->seed + (unsigned int)i0);
		for (long i = i0; i < i1; i++)
			row_ptr[i + 1] = (int32_t)row_degree(p, &M, &g, i);
	}
	int64_t tot = 0;
	row_ptr[0] = 0;
	for (long i = 0; i < m; i++) {
		tot += row_ptr[i + 1];
		if (tot > 0x7FFFFFFFL) {
			snprintf(amg_err, sizeof(amg_err), "nnz exceeds int32 (the reference's row_ptr is int)");
			return -1;
		}
		row_ptr[i + 1] = (int32_t)tot;
	}
	if (bw_corrected_out)
		*bw_corrected_out = b200amg_bw_corrected(p, tot);
	return tot;
}

/* The corrected band width alone (:340-348), for a matrix whose pass 1 gave nnz non-zeros.  Pass 1 depends on
 * (rows, avg, std, distribution, skew, seed) only, so a parameter sweep over bw / neighbours / similarity computes
 * row_ptr once and this per line. */
double b200amg_bw_corrected(const b200amg_params *p, int64_t nnz)
{
	double p_right = p->avg_num_neighbours / 2;
	double cluster = 1 / (1 - p_right);
	double avg_real = ((double)nnz) / p->nr_rows;
	return corrected_bw(p->bw_scaled > 1 ? (double)p->nr_cols : p->bw_scaled * p->nr_cols, avg_real / cluster);
}

/* ---- per-thread column set: open addressing with generation stamps ---- */
typedef struct {
	int32_t *key;
	uint32_t *stamp;
	uint32_t gen;
	uint64_t mask;
} cset_t;

static int cset_init(cset_t *S, long max_degree)
{
	uint64_t cap = 64;
	while (cap < (uint64_t)max_degree * 2 + 2)
		cap <<= 1;
	S->key = (int32_t *)malloc(cap * sizeof(int32_t));
	S->stamp = (uint32_t *)calloc(cap, sizeof(uint32_t));
	S->gen = 0;
	S->mask = cap - 1;
	return (S->key && S->stamp) ? 0 : 1;
}

static inline void cset_clear(cset_t *S)
{
	if (++S->gen == 0) {
		memset(S->stamp, 0, (S->mask + 1) * sizeof(uint32_t));
		S->gen = 1;
	}
}

static inline int cset_insert(cset_t *S, long k)
{
	uint64_t h = ((uint64_t)k * 0x9E3779B97F4A7C15ULL) >> 20;
	for (;; h++) {
		uint64_t s = h & S->mask;
		if (S->stamp[s] != S->gen) {
			S->stamp[s] = S->gen;
			S->key[s] = (int32_t)k;
			return 1;
		}
		if (S->key[s] == (int32_t)k)
			return 0;
	}
}

static int cmp_i32(const void *a, const void *b)
{
	int32_t x = *(const int32_t *)a, y = *(const int32_t *)b;
	return (x > y) - (x < y);
}

static void sort_i32(int32_t *a, long n)
{
	if (n < 24) {
		for (long i = 1; i < n; i++) {
			int32_t v = a[i];
			long j = i - 1;
			for (; j >= 0 && a[j] > v; j--)
				a[j + 1] = a[j];
			a[j + 1] = v;
		}
	} else
		qsort(a, n, sizeof(int32_t), cmp_i32);
}

/* One anchor segment of pass 2 (artificial_matrix_generation.c:366-523).  Writes the rows
 * [i0, i1) of the segment into col / val, which are indexed by (row_ptr[i] - base). */
static int fill_segment(const b200amg_params *p, const model_t *M, double bw_corrected, const int32_t *row_ptr,
                        long i0, long i1, int64_t base, int32_t *col, double *val, cset_t *S)
{
	rng_t g;
	const long nr_cols = p->nr_cols;
	const double p_right = p->avg_num_neighbours / 2;
	const double crs = p->cross_row_similarity;
	long bound_l, bound_r, brl, brr, range, half_range;
	rng_seed(&g, p->seed + (unsigned int)i0);
	for (long i = i0; i < i1; i++) {
		long degree = row_ptr[i + 1] - row_ptr[i];
		if (degree == 0)
			continue;
		int64_t j_s = row_ptr[i] - base, j_e = j_s + degree, j;
		cset_clear(S);
		range = floor(bw_corrected + 0.5);
		if (range < degree)
			range = degree;
		half_range = range / 2;
		if (p->placement == 's') {
			for (j = j_s; j < j_e; j++) {
				col[j] = (int32_t)(j - j_s);
				val[j] = rnd_uniform(&g, 0, 1);
			}
			continue;
		} else if (p->placement == 'd') {
			brl = i - half_range;
			brr = i + half_range + range % 2;
		} else {
			brl = nr_cols / 2 - half_range;
			brr = nr_cols / 2 + half_range + range % 2;
		}
		bound_l = brl;
		bound_r = brr;
		if (bound_l < 0) {
			bound_l = 0;
			if (bound_r < degree)
				bound_r = degree;
		}
		if (bound_r > nr_cols) {
			bound_r = nr_cols;
			if (bound_l > nr_cols - degree)
				bound_l = nr_cols - degree;
		}
		j = j_s;
		int32_t *out = col + j_s;       /* insertion order; sorted at the end of the row */
		long cnt = 0, k = 0;
		if ((crs > 0) && (i % M->period != 0) && (rnd_uniform(&g, 0, 1) <= crs)) {
			long degree_prev = row_ptr[i] - row_ptr[i - 1];
			long degree_min = degree < degree_prev ? degree : degree_prev;
			int64_t jp = j_s - degree_prev, jpe = jp + degree_min;
			/* the previous row may be the output area we are about to overwrite only if
			 * degree_prev == 0, in which case the loop is empty */
			while (jp < jpe) {
				k = col[jp];
				if (k >= bound_r)
					k = bound_l;
				if (!cset_insert(S, k)) {
					snprintf(amg_err, sizeof(amg_err), "row %ld: similarity copy hit a duplicate column (the reference aborts here too)", i);
					return 1;
				}
				out[cnt++] = (int32_t)k;
				val[j] = rnd_uniform(&g, 0, 1);
				j++;
				jp++;
			}
		}
		long d1 = 0, d2 = 0;
		while (j < j_e) {
			long retries = 0;
			do {
				if (retries < 20) {
					k = rnd_uniform_integer(&g, brl, brr);
					if (k >= bound_r)
						k = bound_r - 1;
					else if (k < bound_l)
						k = bound_l;
					retries++;
				} else if (d1 < d2) {
					k = bound_l + d1;
					d1++;
				} else {
					k = bound_r - d2 - 1;
					d2++;
				}
			} while (!cset_insert(S, k));
			out[cnt++] = (int32_t)k;
			val[j] = rnd_uniform(&g, 0, 1);
			j++;
			if (p_right > 0) {
				while (j < j_e) {
					if (rnd_uniform(&g, 0, 1) > p_right)
						break;
					k++;
					if (k >= bound_r)
						k = bound_l;
					if (cset_insert(S, k)) {
						out[cnt++] = (int32_t)k;
						val[j] = rnd_uniform(&g, 0, 1);
						j++;
					}
				}
			}
		}
		sort_i32(out, cnt);
	}
	return 0;
}

/*
 * Pass 2 for the rows [i_s, i_e): col_ind / values receive row_ptr[i_e] - row_ptr[i_s] entries,
 * entry j of the full matrix landing at j - row_ptr[i_s].  Whole anchor segments are generated
 * (a row depends on the random stream since its segment start) into scratch where they stick
 * out of the requested range.
 */
int b200amg_fill(const b200amg_params *p, const int32_t *row_ptr, double bw_corrected,
                 int64_t i_s, int64_t i_e, int32_t *col_ind, double *values)
{
	model_t M;
	if (model_init(p, &M))
		return 1;
	long m = p->nr_rows;
	if (i_s < 0 || i_e > m || i_s > i_e) {
		snprintf(amg_err, sizeof(amg_err), "bad row range");
		return 1;
	}
	if (i_s == i_e)
		return 0;
	long max_degree = 0;
	for (long i = 0; i < m; i++) {
		long d = row_ptr[i + 1] - row_ptr[i];
		if (d > max_degree)
			max_degree = d;
	}
	long seg_first = i_s / M.period, seg_last = (i_e - 1) / M.period;
	int64_t base = row_ptr[i_s];
	int err = 0;
	#pragma omp parallel
	{
		cset_t S;
		int bad = cset_init(&S, max_degree);
		#pragma omp for schedule(dynamic, 1)
		for (long s = seg_first; s <= seg_last; s++) {
			if (bad) {
				#pragma omp atomic write
				err = 2;
				continue;
			}
			long i0 = s * M.period, i1 = i0 + M.period > m ? m : i0 + M.period;
			int rc;
			if (i0 >= i_s && i1 <= i_e) {
				rc = fill_segment(p, &M, bw_corrected, row_ptr, i0, i1, base, col_ind, values, &S);
			} else {
				/* segment sticks out of the shard: generate into scratch, copy the overlap */
				int64_t sb = row_ptr[i0], sn = row_ptr[i1] - sb;
				int32_t *tc = (int32_t *)malloc((sn > 0 ? sn : 1) * sizeof(int32_t));
				double *tv = (double *)malloc((sn > 0 ? sn : 1) * sizeof(double));
				rc = (!tc || !tv) ? 2 : fill_segment(p, &M, bw_corrected, row_ptr, i0, i1, sb, tc, tv, &S);
				if (!rc) {
					long a = i0 > i_s ? i0 : i_s, b = i1 < i_e ? i1 : i_e;
					int64_t off = row_ptr[a] - sb, cntj = row_ptr[b] - row_ptr[a];
					memcpy(col_ind + (row_ptr[a] - base), tc + off, cntj * sizeof(int32_t));
					memcpy(values + (row_ptr[a] - base), tv + off, cntj * sizeof(double));
				}
				free(tc);
				free(tv);
			}
			if (rc) {
				#pragma omp atomic write
				err = rc;
			}
		}
		free(S.key);
		free(S.stamp);
	}
	if (err == 2)
		snprintf(amg_err, sizeof(amg_err), "out of memory");
	return err;
}

/* FNV-1a-64 fingerprint of a raw array (regression pins). */
uint64_t b200amg_fnv1a64(const void *data, int64_t nbytes)
{
	const unsigned char *q = (const unsigned char *)data;
	uint64_t h = 0xcbf29ce484222325ULL;
	for (int64_t i = 0; i < nbytes; i++) {
		h ^= q[i];
		h *= 0x100000001b3ULL;
	}
	return h;
}

/*
 * amg_fill.h -- pass 2 of the synthetic matrix source (column placement + values) written once for the device.
 *
 * SURVEY.md section 8f row N1.  The reference generator (artificial-matrix-generator/artificial_matrix_generation.c)
 * works in two passes: row lengths (:279-354, Box-Muller / Marsaglia-Tsang over glibc's log / cos / pow / exp, m draws)
 * and the non-zeros themselves (:366-523).  Only pass 1 touches libm; pass 2 is integer arithmetic on the glibc TYPE_4
 * additive-feedback stream plus exact power-of-two scalings, so it can run on the GPU bit for bit: pass 1 stays on the
 * host (libb200amg.so, m libm calls), pass 2 -- all the nnz -- is this file.
 *
 * The reference reseeds with seed + i every `period` rows (:259-288): a segment of `period` rows is one sequential
 * random stream, at most 4096 of them.  One WARP owns one segment: lane 0 walks the stream (the draws are data
 * dependent: retries on duplicate columns, neighbour runs that end on a draw), all 32 lanes sort the finished row
 * (a same-direction bitonic network, which tolerates virtual +inf padding) and copy it out.
 *
 * The same source compiles as plain C++ (AMGD_HOST_EMULATION, tests/emul/amg_fill_emul.cpp): there a "warp" is one
 * lane, every comparator of a sort phase is executed by it, and the output is compared bit for bit with csrc/amg.c and
 * with the reference generator on the CPU -- the test this round can run without a GPU.  That build is test
 * infrastructure; the product library contains only the device instantiation.
 */
#pragma once
#include <stdint.h>
#include <math.h>

#if defined(__CUDACC__) && !defined(AMGD_HOST_EMULATION)
#define AMGD_FN __device__ __forceinline__
#define AMGD_LANE ((int) (threadIdx.x & 31))
#define AMGD_NLANES 32
#define AMGD_SYNC() __syncwarp()
#define AMGD_BCAST(v) __shfl_sync(0xffffffffu, (v), 0)
#else
#define AMGD_FN static inline
#define AMGD_LANE 0
#define AMGD_NLANES 1
#define AMGD_SYNC() ((void) 0)
#define AMGD_BCAST(v) (v)
#endif

#define AMGD_DEG 63                         /* glibc TYPE_4: degree 63, separation 1 */
#define AMGD_RANDOM_MAX 0x7FFFFFFFLL
#define AMGD_RANDOM_MAX_64 0x3FFFFFFFFFFFFFFFLL
#define AMGD_STAGE 512                      /* rows up to this long are built and sorted in shared memory */
#define AMGD_SMALL 12                       /* rows up to this long are insertion-sorted by lane 0 */

typedef struct {
	int32_t * st;                       /* AMGD_DEG words (shared memory on the device) */
	int f, r;
} amgd_rng;

typedef struct {
	uint64_t * slot;                    /* (generation << 32) | column; zero-initialised, generation starts at 1 */
	uint64_t mask;
	uint32_t gen;
} amgd_set;

typedef struct {
	int64_t nr_cols;
	int64_t period;                     /* rows per anchor segment */
	double bw_corrected;                /* from pass 1 (b200amg_row_ptr) */
	double p_right;                     /* avg_num_neighbours / 2 */
	double crs;                         /* cross_row_similarity */
	uint32_t seed;
	char placement;                     /* 'r' | 'd' | 's' */
} amgd_model;

AMGD_FN int32_t amgd_next(amgd_rng * g)
{
	uint32_t v = (uint32_t) g->st[g->f] + (uint32_t) g->st[g->r];
	g->st[g->f] = (int32_t) v;
	if (++g->f >= AMGD_DEG) {
		g->f = 0;
		++g->r;
	} else if (++g->r >= AMGD_DEG)
		g->r = 0;
	return (int32_t) (v >> 1);
}

AMGD_FN void amgd_seed(amgd_rng * g, uint32_t seed)
{
	int32_t word = seed ? (int32_t) seed : 1;
	g->st[0] = word;
	for (int i = 1; i < AMGD_DEG; i++) {
		int64_t hi = word / 127773, lo = word % 127773;
		int64_t w = 16807 * lo - 2836 * hi;
		if (w < 0)
			w += 2147483647;
		word = (int32_t) w;
		g->st[i] = word;
	}
	g->f = 1;
	g->r = 0;
	for (int k = 0; k < AMGD_DEG * 10; k++)
		(void) amgd_next(g);
}

/* random.h:227-237 with min = 0, max = 1: v / 2^31, then (1 - 0) * res + 0 -- every step exact, so one exact scaling. */
AMGD_FN double amgd_u01(amgd_rng * g)
{
	return (double) amgd_next(g) * (1.0 / 2147483648.0);
}

/* random.h:177-222: integers over [min, min + range) with the modulo bias rejected.  The rejection limit depends on the
 * range only, so a row computes it once (amgd_draw_init) instead of once per draw. */
typedef struct {
	int64_t min, range, lim;
} amgd_draw;

AMGD_FN void amgd_draw_init(amgd_draw * D, int64_t min, int64_t max)
{
	D->min = min;
	D->range = max - min;
	if (D->range >= AMGD_RANDOM_MAX)
		D->lim = AMGD_RANDOM_MAX_64 - AMGD_RANDOM_MAX_64 % D->range;
	else
		D->lim = AMGD_RANDOM_MAX - (int64_t) ((uint32_t) AMGD_RANDOM_MAX % (uint32_t) D->range);
}

AMGD_FN int64_t amgd_uniform_integer(amgd_rng * g, const amgd_draw * D)
{
	if (D->range >= AMGD_RANDOM_MAX) {
		int64_t res;
		do {
			res = amgd_next(g);
			res <<= 31;
			res |= amgd_next(g);
		} while (res >= D->lim);
		return res % D->range + D->min;
	}
	int32_t v;
	do {
		v = amgd_next(g);
	} while (v >= D->lim);
	return (int64_t) ((uint32_t) v % (uint32_t) D->range) + D->min;
}

AMGD_FN void amgd_set_clear(amgd_set * S)
{
	++S->gen;                           /* a segment has far fewer than 2^32 rows: no wrap */
}

AMGD_FN int amgd_set_insert(amgd_set * S, int64_t k)
{
	const uint64_t want = ((uint64_t) S->gen << 32) | (uint32_t) k;
	uint64_t h = ((uint64_t) k * 0x9E3779B97F4A7C15ULL) >> 20;
	for (;; h++) {
		uint64_t s = h & S->mask;
		uint64_t v = S->slot[s];
		if ((uint32_t) (v >> 32) != S->gen) {
			S->slot[s] = want;
			return 1;
		}
		if (v == want)
			return 0;
	}
}

/*
 * One row with degree > 0 (artificial_matrix_generation.c:380-523), lane 0 only.  Columns go to out[] in insertion
 * order (sorted afterwards by the warp), values to val[0 .. degree).  prev = the previous row's SORTED columns.
 * Returns the number of columns written (= degree), or -1 when the similarity copy meets a duplicate column (the
 * reference aborts there too).
 */
AMGD_FN int64_t amgd_row(const amgd_model * M, amgd_rng * g, amgd_set * S, int64_t i, int64_t degree, int64_t degree_prev,
                         const int32_t * prev, int32_t * out, double * val)
{
	const int64_t nr_cols = M->nr_cols;
	int64_t range = (int64_t) floor(M->bw_corrected + 0.5);
	if (range < degree)
		range = degree;
	const int64_t half_range = range / 2;
	int64_t brl, brr, bound_l, bound_r;
	if (M->placement == 's') {
		for (int64_t j = 0; j < degree; j++) {
			out[j] = (int32_t) j;
			val[j] = amgd_u01(g);
		}
		return degree;
	}
	if (M->placement == 'd') {
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
	amgd_set_clear(S);
	amgd_draw D;
	amgd_draw_init(&D, brl, brr);
	int64_t j = 0, k = 0;
	if ((M->crs > 0) && (i % M->period != 0) && (amgd_u01(g) <= M->crs)) {
		const int64_t degree_min = degree < degree_prev ? degree : degree_prev;
		for (int64_t jp = 0; jp < degree_min; jp++) {
			k = prev[jp];
			if (k >= bound_r)
				k = bound_l;
			if (!amgd_set_insert(S, k))
				return -1;
			out[j] = (int32_t) k;
			val[j] = amgd_u01(g);
			j++;
		}
	}
	int64_t d1 = 0, d2 = 0;
	while (j < degree) {
		int retries = 0;
		do {
			if (retries < 20) {
				k = amgd_uniform_integer(g, &D);
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
		} while (!amgd_set_insert(S, k));
		out[j] = (int32_t) k;
		val[j] = amgd_u01(g);
		j++;
		if (M->p_right > 0) {
			while (j < degree) {
				if (amgd_u01(g) > M->p_right)
					break;
				k++;
				if (k >= bound_r)
					k = bound_l;
				if (amgd_set_insert(S, k)) {
					out[j] = (int32_t) k;
					val[j] = amgd_u01(g);
					j++;
				}
			}
		}
	}
	return degree;
}

/*
 * Ascending sort of a[0 .. n) by all lanes of the warp: a bitonic network in which every comparator orders
 * (low index, high index) ascending -- each merge starts with a "flip" stage (i <-> block end - i), the rest are
 * half-cleaners -- so elements past n, taken as +inf, never move and n need not be a power of two.  Columns of a row
 * are distinct, so the result does not depend on the algorithm.
 */
AMGD_FN void amgd_sort(int32_t * a, int64_t n)
{
	const int lane = AMGD_LANE;
	if (n <= AMGD_SMALL) {
		if (lane == 0)
			for (int64_t i = 1; i < n; i++) {
				int32_t v = a[i];
				int64_t q = i - 1;
				for (; q >= 0 && a[q] > v; q--)
					a[q + 1] = a[q];
				a[q + 1] = v;
			}
		AMGD_SYNC();
		return;
	}
	int lp = 0;
	while (((int64_t) 1 << lp) < n)
		lp++;
	const int64_t half = ((int64_t) 1 << lp) >> 1;
	for (int lk = 1; lk <= lp; lk++) {                                /* merge blocks of k = 2^lk */
		const int64_t k = (int64_t) 1 << lk, hk = k >> 1;
		for (int64_t t = lane; t < half; t += AMGD_NLANES) {      /* flip stage */
			int64_t blk = t >> (lk - 1), off = t & (hk - 1);
			int64_t lo = blk * k + off, hi = blk * k + k - 1 - off;
			if (hi < n) {
				int32_t x = a[lo], y = a[hi];
				if (x > y) {
					a[lo] = y;
					a[hi] = x;
				}
			}
		}
		AMGD_SYNC();
		for (int ld = lk - 2; ld >= 0; ld--) {                    /* half-cleaners, distance d = 2^ld */
			const int64_t d = (int64_t) 1 << ld;
			for (int64_t t = lane; t < half; t += AMGD_NLANES) {
				int64_t blk = t >> ld, off = t & (d - 1);
				int64_t lo = blk * 2 * d + off, hi = lo + d;
				if (hi < n) {
					int32_t x = a[lo], y = a[hi];
					if (x > y) {
						a[lo] = y;
						a[hi] = x;
					}
				}
			}
			AMGD_SYNC();
		}
	}
}

/*
 * One anchor segment, rows [i0, i1), by one warp.  row_ptr is the whole matrix's; entry j of the matrix is written at
 * col[j - base] / val[j - base].  rng_words: AMGD_DEG ints, stage: AMGD_STAGE ints (shared memory on the device),
 * slots: the segment's zero-initialised set storage with mask + 1 entries, mask + 1 >= 2 * (longest row) + 2.
 * Returns 0, or 1 + the row at which the similarity copy met a duplicate.
 */
AMGD_FN int64_t amgd_segment(const amgd_model * M, const int32_t * row_ptr, int64_t i0, int64_t i1, int64_t base, int32_t * col,
                             double * val, int32_t * rng_words, int32_t * stage, uint64_t * slots, uint64_t mask)
{
	const int lane = AMGD_LANE;
	amgd_rng g;
	amgd_set S;
	g.st = rng_words;
	g.f = g.r = 0;
	S.slot = slots;
	S.mask = mask;
	S.gen = 0;
	if (lane == 0)
		amgd_seed(&g, M->seed + (uint32_t) i0);
	int64_t rp_i = row_ptr[i0];
	int64_t degree_prev = 0;
	for (int64_t i = i0; i < i1; i++) {
		const int64_t rp_n = row_ptr[i + 1];
		const int64_t degree = rp_n - rp_i;
		if (degree > 0) {
			const int64_t j_s = rp_i - base;
			const int staged = degree <= AMGD_STAGE;
			int32_t * out = staged ? stage : col + j_s;
			int64_t cnt = 0;
			if (lane == 0)
				cnt = amgd_row(M, &g, &S, i, degree, degree_prev, col + j_s - degree_prev, out, val + j_s);
			AMGD_SYNC();
			cnt = AMGD_BCAST(cnt);
			if (cnt < 0)
				return 1 + i;
			if (M->placement != 's')
				amgd_sort(out, cnt);
			if (staged) {
				for (int64_t t = lane; t < cnt; t += AMGD_NLANES)
					col[j_s + t] = stage[t];
				AMGD_SYNC();
			}
		}
		degree_prev = degree;
		rp_i = rp_n;
	}
	return 0;
}

/*
 * b200amg.h -- C-ABI of the host-side synthetic matrix source (libb200amg.so).
 *
 * Replaces, for the SpMV hot path's input side, the reference entry point
 *   struct csr_matrix * artificial_matrix_generation(long nr_rows, long nr_cols, double avg_nnz_per_row,
 *       double std_nnz_per_row, char * distribution, unsigned int seed, char * placement, double bw_scaled,
 *       double skew, double avg_num_neighbours, double cross_row_similarity)
 *   (artificial-matrix-generator/artificial_matrix_generation.h:92, called from
 *    benchmark_code/BENCH/src/bench.cpp:250), producing bit-identical row_ptr / col_ind / values.
 * Split in two calls so that a row-sharded run generates only its own rows.
 * All functions return 0 / a non-negative count on success; on failure a non-zero / negative
 * value and b200amg_last_error() describes it.  Plain pointers and sizes only.
 */
#ifndef B200AMG_H
#define B200AMG_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200amg_params {
	int64_t nr_rows;
	int64_t nr_cols;
	double avg_nnz_per_row;
	double std_nnz_per_row;
	char distribution;          /* 'n' normal, 'g' gamma  (first letter of the reference's string) */
	char placement;             /* 'r' random, 'd' diagonal, 's' simple */
	char pad_[6];
	double bw_scaled;
	double skew;
	double avg_num_neighbours;
	double cross_row_similarity;
	uint32_t seed;
	uint32_t pad2_;
} b200amg_params;

/* pass 1: fills row_ptr[0..nr_rows]; returns nnz or -1.  *bw_corrected feeds b200amg_fill. */
int64_t b200amg_row_ptr(const b200amg_params * p, int32_t * row_ptr, double * bw_corrected);

/* bw_corrected alone, given pass 1's nnz: pass 1 does not depend on bw / neighbours / similarity, so a sweep over those
 * computes row_ptr once per (rows, avg, std, distribution, skew, seed) and this per line. */
double b200amg_bw_corrected(const b200amg_params * p, int64_t nnz);

/* pass 2 for rows [i_s, i_e): entry j of the full matrix is written at j - row_ptr[i_s]. */
int b200amg_fill(const b200amg_params * p, const int32_t * row_ptr, double bw_corrected,
                 int64_t i_s, int64_t i_e, int32_t * col_ind, double * values);

uint64_t b200amg_fnv1a64(const void * data, int64_t nbytes);
const char * b200amg_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* B200AMG_H */

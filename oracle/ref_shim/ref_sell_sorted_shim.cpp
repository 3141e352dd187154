/*
 * ref_sell_sorted_shim.cpp -- builds the REFERENCE's own per-thread-sorted SELL plugin (K/sell_sorted.cpp, slices of VEC_LEN =
 * vec_len_default_f64 rows) into oracle/_ref/libref_sell_sorted.so and exposes its layout arrays.
 * The reference source is #included from where it lies under $(REF); nothing is copied.
 * TEST INFRASTRUCTURE ONLY (see oracle/spmv_oracle.c header).
 */
#include <stdlib.h>
#include <stdint.h>

#include REF_SELL_SORTED_CPP         /* -> $(REF)/benchmark_code/BENCH/src/spmv_kernels/sell_sorted.cpp */

extern "C" {

/* lib/functools/functools_gen.c:334 asks the topology library for the L1 size to pick the block length of its parallel scan
 * (an integer prefix sum: the result does not depend on it).  Answered here instead of linking the harness's topology / io
 * / mmap sources. */
long topohw_get_cache_size(long cpu_id, long level, enum cache_types type) { return 32 * 1024; }

/* lib/omp_functions.c:43-50 (the team size of the enclosing, else of a following, parallel region); that file also pulls
 * the harness's affinity / pthread sources, which nothing here needs. */
int safe_omp_get_num_threads()
{
	if (omp_get_level() > 0)
		return omp_get_num_threads();
	return omp_get_max_threads();
}

void * ref_sells_create(int32_t * row_ptr, int32_t * col_ind, double * values, long m, long n, long nnz)
{
	return (void *) csr_to_format(row_ptr, col_ind, values, m, n, nnz, 0, 0);
}

void ref_sells_spmv(void * h, double * x, double * y)
{
	((struct Matrix_Format *) h)->spmv(x, y);
}

const char * ref_sells_format_name(void * h) { return ((struct Matrix_Format *) h)->format_name; }
long ref_sells_vec_len(void) { return VEC_LEN; }
long ref_sells_num_row_clusters(void * h) { return ((SELLArray *) h)->num_row_clusters; }
long ref_sells_nnz_ext(void * h) { return ((SELLArray *) h)->nnz_ext; }
const int32_t * ref_sells_row_cluster_ptr(void * h) { return ((SELLArray *) h)->row_cluster_ptr; }
const int32_t * ref_sells_ja(void * h) { return ((SELLArray *) h)->ja; }
const double * ref_sells_a(void * h) { return ((SELLArray *) h)->a; }
const int * ref_sells_permutation(void * h) { return ((SELLArray *) h)->permutation; }
const int * ref_sells_rev_permutation(void * h) { return ((SELLArray *) h)->rev_permutation; }

}

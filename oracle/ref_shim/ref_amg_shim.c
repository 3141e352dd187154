/*
 * ref_amg_shim.c -- C entry points over the REFERENCE's artificial matrix generator
 * (artificial-matrix-generator/artificial_matrix_generation.c:163-579), compiled from $(REF)
 * into oracle/_ref/libref_amg.so.  Used to pin the product's own generator restatement
 * (spmv_research_b200/csrc/amg.c) bit for bit.  TEST INFRASTRUCTURE ONLY.
 */
#include <stdint.h>
#include <stdlib.h>
#include "artificial_matrix_generation.h"

void * ref_amg_generate(long nr_rows, long nr_cols, double avg, double std, char * distribution, unsigned int seed, char * placement, double bw, double skew, double neigh, double crs)
{
	return artificial_matrix_generation(nr_rows, nr_cols, avg, std, distribution, seed, placement, bw, skew, neigh, crs);
}
int64_t ref_amg_nnz(void * h) { return ((struct csr_matrix *) h)->nr_nzeros; }
int * ref_amg_row_ptr(void * h) { return ((struct csr_matrix *) h)->row_ptr; }
int * ref_amg_col_ind(void * h) { return ((struct csr_matrix *) h)->col_ind; }
double * ref_amg_values(void * h) { return ((struct csr_matrix *) h)->values; }
void ref_amg_free(void * h) { free_csr_matrix((struct csr_matrix *) h); }

/*
 * ref_sellcs_shim.c -- thin C entry points over the REFERENCE's SELL-C-sigma format library
 * (K/sell-C-s/RISC-V/{sellcs_format,radix_sort,sellcs_utils}.c), compiled from $(REF) into
 * oracle/_ref/libref_sellcs{32,64}.so (without / with -DINDEX64).  TEST INFRASTRUCTURE ONLY.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "sellcs-spmv.h"

int ref_sellcs_index_bytes(void) { return (int) sizeof(index_t); }

void * ref_sellcs_create(int64_t m, int64_t n, const index_t * row_ptr, const index_t * col_ind, const double * values, uint64_t C, uint64_t sigma)
{
	sellcs_matrix_t * M = (sellcs_matrix_t *) calloc(1, sizeof(*M));
	sellcs_init_params(C, sigma, M);
	sellcs_create_matrix_from_CSR_rd((index_t) m, (index_t) n, row_ptr, col_ind, values, 0, 0, M);
	return M;
}

int64_t ref_sellcs_nslices(void * h) { return ((sellcs_matrix_t *) h)->nslices; }
int64_t ref_sellcs_padded(void * h) { sellcs_matrix_t * M = h; return M->slice_pointers[M->nslices]; }
const index_t * ref_sellcs_row_order(void * h) { return ((sellcs_matrix_t *) h)->row_order; }
const index_t * ref_sellcs_slice_widths(void * h) { return ((sellcs_matrix_t *) h)->slice_widths; }
const index_t * ref_sellcs_slice_pointers(void * h) { return ((sellcs_matrix_t *) h)->slice_pointers; }
const index_t * ref_sellcs_column_indices(void * h) { return ((sellcs_matrix_t *) h)->column_indices; }
const double * ref_sellcs_values(void * h) { return ((sellcs_matrix_t *) h)->values; }

void ref_sellcs_free(void * h)
{
	sellcs_matrix_t * M = h;
	free(M->row_order); free(M->slice_widths); free(M->slice_pointers); free(M->column_indices); free(M->values); free(M);
}

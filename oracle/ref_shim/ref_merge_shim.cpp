/*
 * ref_merge_shim.cpp -- builds the REFERENCE's CPU merge-path CsrMV (K/merge.cpp:230-318)
 * into oracle/_ref/libref_merge.so.  Sources are #included from $(REF); nothing is copied.
 * TEST INFRASTRUCTURE ONLY.
 */
#include <stdlib.h>
#include <stdint.h>

#include REF_MERGE_CPP        /* -> $(REF)/benchmark_code/BENCH/src/spmv_kernels/merge.cpp */

extern "C" {

void ref_merge_path_search(int32_t diagonal, int32_t * row_end, int32_t m, int32_t nnz, int32_t * out_xy)
{
	int2 c;
	CountingInputIterator nz(0);
	MergePathSearch(diagonal, row_end, nz, m, nnz, c);
	out_xy[0] = c.x;
	out_xy[1] = c.y;
}

void ref_merge_csrmv(int num_threads, int32_t * row_ptr, int32_t * col_ind, double * values, double * x, double * y, int32_t m, int32_t nnz)
{
	OmpMergeCsrmv(num_threads, row_ptr + 1, col_ind, values, x, y, m, nnz);
}

}

/*
 * ref_csr_shim.cpp -- builds the REFERENCE's own CPU CSR plugin (K/csr.cpp) and its
 * nnz-balanced partitioner macro (lib/parallel_util.h:156-184) into oracle/_ref/libref_csr.so.
 * The reference sources are #included from where they lie under $(REF); nothing is copied.
 * Built twice: ValueType = double (libref_csr.so, the `_d` build) and ValueType = float (libref_csr_f.so, the `_f` build,
 * make.sh:215-216: the matrix still arrives as double and is narrowed by the plugin's constructor).
 * TEST INFRASTRUCTURE ONLY (see oracle/spmv_oracle.c header).
 */
#include <stdlib.h>
#include <stdint.h>

int prefetch_distance = 1;   /* K/csr.cpp:30 declares it extern; BENCH never defines it (SURVEY 8c gotcha 1) */

#include REF_CSR_CPP          /* -> $(REF)/benchmark_code/BENCH/src/spmv_kernels/csr.cpp */

extern "C" {

void * ref_csr_create(int32_t * row_ptr, int32_t * col_ind, double * values, long m, long n, long nnz)
{
	setenv("USE_PROCESSES", "0", 0);    /* K/csr.cpp:89 does atoi(getenv("USE_PROCESSES")) */
	return (void *) csr_to_format(row_ptr, col_ind, values, m, n, nnz, 0, 0);
}

void ref_csr_spmv(void * h, ValueType * x, ValueType * y)
{
	((struct Matrix_Format *) h)->spmv(x, y);
}

const char * ref_csr_format_name(void * h)
{
	return ((struct Matrix_Format *) h)->format_name;
}

/* the serial oracle: subkernel_csr_scalar over all rows (K/csr.cpp:334-350) */
void ref_csr_spmv_serial(void * h, ValueType * x, ValueType * y)
{
	CSRArrays * csr = (CSRArrays *) h;
	subkernel_csr_scalar(csr, x, y, 0, csr->m);
}

void ref_partition_prefix_sums(long num_workers, long worker_pos, int32_t * sums, long N, long total, long * s, long * e)
{
	long i_s, i_e;
	loop_partitioner_balance_prefix_sums(num_workers, worker_pos, sums, N, total, &i_s, &i_e);
	*s = i_s;
	*e = i_e;
}

}

/*
 * ref_mtx_dump.cpp -- the REFERENCE's own Matrix Market ingest (bench.cpp:128-239: mtx_read + mtx_values_convert_to_real
 * + coo_to_csr, from lib/storage_formats/matrix_market/matrix_market.c and lib/aux/csr_converter.c) as a tiny
 * executable: reads a .mtx file the way import_file() does and dumps the CSR arrays in binary.  Linked with the harness's
 * library objects compiled from where they lie under $(REF); nothing is copied.  TEST INFRASTRUCTURE ONLY.
 *
 *   ref_mtx_dump.exe in.mtx out.bin      out.bin: int64 m, n, nnz, symmetric, nnz_diag, nnz_non_diag; int32 ia[m+1]; int32 ja[nnz]; double a[nnz]
 */
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>

#include "macros/cpp_defines.h"

extern "C" {
	#include "storage_formats/matrix_market/matrix_market.h"
	#include "aux/csr_converter_reference.h"
}

int main(int argc, char ** argv)
{
	if (argc != 3)
		return 2;
	struct Matrix_Market * MTX = mtx_read(argv[1], 1, 1);
	int64_t hdr[6] = { MTX->m, MTX->n, MTX->nnz, MTX->symmetric, MTX->nnz_diag, MTX->nnz_non_diag };
	mtx_values_convert_to_real(MTX);
	long m = MTX->m, n = MTX->n, nnz = MTX->nnz;
	INT_T * ia = (INT_T *) calloc(m + 1, sizeof(INT_T));
	INT_T * ja = (INT_T *) calloc(nnz > 0 ? nnz : 1, sizeof(INT_T));
	ValueTypeReference * a = (ValueTypeReference *) calloc(nnz > 0 ? nnz : 1, sizeof(ValueTypeReference));
	coo_to_csr((INT_T *) MTX->R, (INT_T *) MTX->C, (ValueTypeReference *) MTX->V, m, n, nnz, ia, ja, a, 1, 0);
	FILE * f = fopen(argv[2], "wb");
	if (!f)
		return 3;
	fwrite(hdr, sizeof(hdr), 1, f);
	fwrite(ia, sizeof(INT_T), m + 1, f);
	fwrite(ja, sizeof(INT_T), nnz, f);
	fwrite(a, sizeof(ValueTypeReference), nnz, f);
	fclose(f);
	return 0;
}

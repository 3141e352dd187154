/*
 * b200mtx.h -- Matrix Market (.mtx) reader for the B200 SpMV engine: the step before the hot path when the
 * input is a real matrix file instead of the synthetic generator.
 *
 * Replaces, on the host, what the reference harness does in import_file() (benchmark_code/BENCH/src/bench.cpp:128-206):
 *   mtx_read(filename, expand_symmetry = 1, pattern_dummy_vals = 1)     lib/storage_formats/matrix_market/matrix_market.c:257-325
 *   mtx_values_convert_to_real()                                        matrix_market.c:420-452
 * with the same conventions: "coordinate" format only here (the reference also reads dense "array" files, which are not
 * sparse inputs); fields real / integer / pattern (values 1.0) / complex (values |z|); symmetry general / symmetric /
 * skew-symmetric (mirrored value negated) / Hermitian; indices converted to base 0; numbers parsed with strtol / strtod
 * like the reference's gen_strtonum (lib/genlib.h:413-445, :507-530).  Output order = the reference's COO order: the file's
 * entries in file order, then the mirrors of the off-diagonal entries in file order (matrix_market_gen.c:60-125).
 *
 * The COO -> CSR step that follows (bench.cpp:208-229, coo_to_csr) runs on the GPU: b200spmv_coo_to_csr /
 * b200spmv_create_coo in b200spmv.h.
 */
#ifndef B200MTX_H
#define B200MTX_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum b200mtx_field { B200MTX_REAL = 0, B200MTX_INTEGER = 1, B200MTX_PATTERN = 2, B200MTX_COMPLEX = 3 };

typedef struct b200mtx_info {
	int64_t m, n;
	int64_t nnz;              /* entries returned (after symmetry expansion when requested)             */
	int64_t nnz_file;         /* entries in the file (Matrix_Market.nnz_sym)                            */
	int64_t nnz_diag;         /* file entries with row == col                                           */
	int64_t nnz_non_diag;
	int32_t symmetric;        /* 1 for symmetric / skew-symmetric / Hermitian files                     */
	int32_t skew_symmetry;
	int32_t hermitian_symmetry;
	int32_t field;            /* enum b200mtx_field of the FILE (the returned values are always double) */
} b200mtx_info;

/* Reads `path`.  On success returns 0 and hands out three malloc'd arrays of info->nnz entries (release each with
 * b200mtx_free).  expand_symmetry != 0: mirrored entries appended (what the harness does unless built with KEEP_SYMMETRY).
 * Errors (message in b200mtx_last_error): unreadable file, not "matrix coordinate", unknown field / symmetry, bad size line,
 * number of data lines != declared entries (matrix_market.c:237-238), an index outside [1, m] x [1, n], unparsable line. */
int b200mtx_read(const char * path, int32_t expand_symmetry, b200mtx_info * info, int32_t ** row, int32_t ** col, double ** val);

void b200mtx_free(void * p);

const char * b200mtx_last_error(void);

#ifdef __cplusplus
}
#endif

#endif

/*
 * mtx.c -- Matrix Market reader (include/b200mtx.h).
 *
 * What the reference does in mtx_read() / mtx_values_convert_to_real() (lib/storage_formats/matrix_market/matrix_market.c:
 * 150-325, 420-452; matrix_market_gen.c:60-200), written from scratch: the file is read in one piece, cut into lines, the
 * data lines are parsed in parallel with strtol / strtod (the reference's gen_strtonum ends in the same libc calls, so the
 * doubles are bit-identical), and for symmetric files the mirrors of the off-diagonal entries are appended in file order.
 */
#define _GNU_SOURCE
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <strings.h>
#include <stdarg.h>
#include <math.h>
#include <errno.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../../include/b200mtx.h"

static __thread char mtx_err[384];
const char * b200mtx_last_error(void) { return mtx_err; }
void b200mtx_free(void * p) { free(p); }

static int fail(const char * fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(mtx_err, sizeof(mtx_err), fmt, ap);
	va_end(ap);
	return 1;
}

/* whole file + a terminating NUL */
static char * slurp(const char * path, int64_t * size_out)
{
	FILE * f = fopen(path, "rb");
	if (!f)
		return NULL;
	if (fseeko(f, 0, SEEK_END)) {
		fclose(f);
		return NULL;
	}
	const int64_t size = ftello(f);
	rewind(f);
	char * buf = (char *) malloc((size_t) size + 2);
	if (!buf) {
		fclose(f);
		return NULL;
	}
	const size_t got = fread(buf, 1, (size_t) size, f);
	fclose(f);
	if ((int64_t) got != size) {
		free(buf);
		return NULL;
	}
	buf[size] = '\0';
	buf[size + 1] = '\0';
	*size_out = size;
	return buf;
}

static int is_blank_line(const char * s)
{
	for (; *s; s++)
		if (*s != ' ' && *s != '\t' && *s != '\r')
			return 0;
	return 1;
}

int b200mtx_read(const char * path, int32_t expand_symmetry, b200mtx_info * info, int32_t ** row_out, int32_t ** col_out, double ** val_out)
{
	if (!path || !info || !row_out || !col_out || !val_out)
		return fail("NULL argument");
	*row_out = NULL;
	*col_out = NULL;
	*val_out = NULL;
	memset(info, 0, sizeof(*info));
	int64_t size = 0;
	char * buf = slurp(path, &size);
	if (!buf)
		return fail("cannot read '%s': %s", path, strerror(errno));

	/* lines: NUL-terminated in place; trailing empty lines dropped (the reference's file_to_lines yields no atom for them) */
	int64_t nlines = 0, cap = 1024;
	char ** lines = (char **) malloc((size_t) cap * sizeof(char *));
	int rc = 1;
	int32_t * R = NULL, * C = NULL;
	double * V = NULL;
	if (!lines) {
		free(buf);
		return fail("out of memory");
	}
	for (char * p = buf; p < buf + size;) {
		char * e = (char *) memchr(p, '\n', (size_t) (buf + size - p));
		if (!e)
			e = buf + size;
		*e = '\0';
		if (nlines == cap) {
			cap *= 2;
			char ** nl = (char **) realloc(lines, (size_t) cap * sizeof(char *));
			if (!nl) {
				fail("out of memory");
				goto done;
			}
			lines = nl;
		}
		lines[nlines++] = p;
		p = e + 1;
	}
	while (nlines > 0 && is_blank_line(lines[nlines - 1]))
		nlines--;
	if (nlines == 0) {
		fail("'%s' is empty", path);
		goto done;
	}

	/* header (matrix_market.c:150-222): without the %%MatrixMarket banner the reference assumes "coordinate real general" */
	int64_t i = 0;
	int field = B200MTX_REAL, symmetric = 0, skew = 0, herm = 0;
	{
		char t0[64] = "", object[64] = "", format[64] = "", fld[64] = "", sym[64] = "";
		sscanf(lines[0], "%63s %63s %63s %63s %63s", t0, object, format, fld, sym);
		if (!strcmp(t0, "%%MatrixMarket")) {
			if (strcmp(object, "matrix")) {
				fail("'%s': only 'matrix' objects are supported (found '%s')", path, object);
				goto done;
			}
			if (strcmp(format, "coordinate")) {
				fail("'%s': only the sparse 'coordinate' format is read here (found '%s')", path, format);
				goto done;
			}
			if (!strcmp(fld, "real")) field = B200MTX_REAL;
			else if (!strcmp(fld, "integer")) field = B200MTX_INTEGER;
			else if (!strcmp(fld, "pattern")) field = B200MTX_PATTERN;
			else if (!strcmp(fld, "complex")) field = B200MTX_COMPLEX;
			else {
				fail("'%s': unrecognized field type: %s", path, fld);
				goto done;
			}
			if (!strcmp(sym, "general")) symmetric = 0;
			else if (!strcmp(sym, "symmetric")) symmetric = 1;
			else if (!strcmp(sym, "skew-symmetric")) symmetric = skew = 1;
			else if (!strcasecmp(sym, "hermitian")) symmetric = herm = 1;
			else {
				fail("'%s': unsupported symmetry type: %s", path, sym);
				goto done;
			}
			i = 1;
		}
	}
	while (i < nlines && lines[i][0] == '%')
		i++;
	long M = 0, N = 0, nnz_file = 0;
	if (i >= nlines || sscanf(lines[i], "%ld%ld%ld", &M, &N, &nnz_file) != 3 || M < 0 || N < 0 || nnz_file < 0) {
		fail("'%s': invalid/missing matrix sizes: %s", path, i < nlines ? lines[i] : "(end of file)");
		goto done;
	}
	i++;
	if (nnz_file != nlines - i) {
		fail("'%s': remaining number of file lines (%lld) don't match the number of non-zeros (%ld)", path, (long long) (nlines - i), nnz_file);
		goto done;
	}
	const int64_t cap_nnz = (symmetric && expand_symmetry) ? 2 * (int64_t) nnz_file : nnz_file;
	if (M >= 0x7fffffffL || N >= 0x7fffffffL || cap_nnz > 0x7fffffffLL) {
		fail("'%s': m, n and nnz must fit int32 indices", path);
		goto done;
	}
	R = (int32_t *) malloc((size_t) (cap_nnz > 0 ? cap_nnz : 1) * sizeof(int32_t));
	C = (int32_t *) malloc((size_t) (cap_nnz > 0 ? cap_nnz : 1) * sizeof(int32_t));
	V = (double *) malloc((size_t) (cap_nnz > 0 ? cap_nnz : 1) * sizeof(double));
	if (!R || !C || !V) {
		fail("out of memory");
		goto done;
	}

	/* data lines (matrix_market_gen.c:128-200); values converted as mtx_values_convert_to_real does (matrix_market.c:420-452) */
	int64_t non_diag = 0, bad_line = -1;
	char ** data = lines + i;
	#pragma omp parallel for schedule(static) reduction(+ : non_diag)
	for (int64_t e = 0; e < nnz_file; e++) {
		char * p = data[e], * q;
		const long r = strtol(p, &q, 10);
		int ok = (q != p);
		p = q;
		const long c = strtol(p, &q, 10);
		ok = ok && (q != p);
		p = q;
		double v = 1.0;                                   /* pattern: pattern_dummy_vals */
		if (ok && field == B200MTX_REAL) {
			v = strtod(p, &q);
			ok = (q != p);
		} else if (ok && field == B200MTX_INTEGER) {
			v = (double) (int) strtol(p, &q, 10);             /* parsed into an int, then widened (:426-431) */
			ok = (q != p);
		} else if (ok && field == B200MTX_COMPLEX) {
			const double re = strtod(p, &q);
			ok = (q != p);
			p = q;
			const double im = strtod(p, &q);
			ok = ok && (q != p);
			v = hypot(re, im);                                /* cabs (:433-439); the mirror's conjugate has the same modulus */
		}
		if (!ok || r < 1 || r > M || c < 1 || c > N) {
			#pragma omp critical
			if (bad_line < 0 || e < bad_line)
				bad_line = e;
			continue;
		}
		R[e] = (int32_t) (r - 1);
		C[e] = (int32_t) (c - 1);
		V[e] = v;
		non_diag += (r != c);
	}
	if (bad_line >= 0) {
		fail("'%s': badly formed entry or index out of range at data line %lld: %s", path, (long long) (bad_line + 1), data[bad_line]);
		goto done;
	}
	int64_t nnz = nnz_file;
	if (symmetric && expand_symmetry) {
		/* mirrors of the off-diagonal entries, appended in file order (matrix_market_gen.c:60-125) */
		int64_t j = nnz_file;
		for (int64_t e = 0; e < nnz_file; e++)
			if (C[e] != R[e]) {
				R[j] = C[e];
				C[j] = R[e];
				V[j] = (skew && field != B200MTX_COMPLEX) ? -V[e] : V[e];
				j++;
			}
		nnz = j;
	}
	info->m = M;
	info->n = N;
	info->nnz = nnz;
	info->nnz_file = nnz_file;
	info->nnz_non_diag = non_diag;
	info->nnz_diag = nnz_file - non_diag;
	info->symmetric = symmetric;
	info->skew_symmetry = skew;
	info->hermitian_symmetry = herm;
	info->field = field;
	*row_out = R;
	*col_out = C;
	*val_out = V;
	R = NULL;
	C = NULL;
	V = NULL;
	rc = 0;
done:
	free(R);
	free(C);
	free(V);
	free(lines);
	free(buf);
	return rc;
}

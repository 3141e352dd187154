/*
 * b200_plugin.cpp -- format plugin of the LiHaoxu/SpMV-Research benchmark harness backed by libb200spmv.
 *
 * Drop this file into benchmark_code/BENCH/src/spmv_kernels/ and add the two targets shown in
 * INTEGRATION.md to Makefile_in; it defines exactly the three symbols every plugin translation unit
 * defines (spmv_kernels/spmv_kernel.h:9-29, template spmv_kernels/spmv_kernel_template.cpp:21-93):
 *
 *     struct Matrix_Format * csr_to_format(INT_T * row_ptr, INT_T * col_ind, ValueTypeReference * values,
 *                                          long m, long n, long nnz, long symmetric, long symmetry_expanded);
 *     int statistics_print_labels(char * buf, long buf_n);
 *     struct B200_Format : Matrix_Format { spmv(x, y), statistics_start(), statistics_print_data(buf, n) }
 *
 * No CUDA in this file: every device operation goes through the C-ABI of include/b200spmv.h, so the
 * harness's ordinary C++ compiler builds it (no nvcc pass, no _nv suffix needed).
 *
 * Build-time switch:  -DB200_FORMAT_SELL  selects SELL-C-sigma (format name B200_SELL_C_SIGMA), -DB200_FORMAT_AUTO builds
 *                     both at csr_to_format time, times them on the device and keeps the faster (B200_AUTO_CSR /
 *                     B200_AUTO_SELL); default is the adaptive CSR kernel (B200_CSR_ADAPTIVE).
 * Run-time switches (environment, read once in csr_to_format):
 *     B200SPMV_TRANSFER = first | always   x upload / y download on the first spmv call only (the reference's
 *                         GPU-plugin convention, spmv_kernels/GPU_clean/cuda_csr.cu:353-430; default) or on every call
 *     B200SPMV_PIN      = 1 | 0   page-lock the harness's x / y buffers on first use (default 1)
 *     B200SPMV_DEVICE   = CUDA device ordinal (default: current device)
 *     B200SPMV_SELL_C / B200SPMV_SELL_SIGMA   slice height / sort window (defaults 32 / 16384, sell_c_s.cpp:58-60)
 * Errors follow the reference: message on stderr and exit(1) (lib/debug.h:135); there is no CPU fallback.
 *
 * The `_f` build axis (make.sh:215-216: -D'DOUBLE=0' -D'ValueType=float', ValueTypeReference stays double): the same file
 * compiles to the fp32 plugin over b200spmv_f32_* -- csr_to_format narrows the double values on the GPU, spmv takes float
 * vectors and accumulates in float like K/csr.cpp:334-350 built `_f`.  CSR only (B200_FORMAT_SELL / _AUTO are refused at
 * compile time); format name B200_CSR_ADAPTIVE_F32.  x and y move on every call (no flags in the f32 entry point).
 */
#include <stdlib.h>
#include <stdio.h>
#include <string.h>
#include <stdint.h>

#include "macros/cpp_defines.h"

#include "spmv_kernel.h"

#include "b200spmv.h"

#ifndef INT_T
	#error "INT_T must be defined by the harness build (make.sh:166)"
#endif

#if !defined(DOUBLE)
	#error "DOUBLE must be defined by the harness build (make.sh:212-216): 1 = _d, 0 = _f"
#endif
#if DOUBLE == 0 && (defined(B200_FORMAT_SELL) || defined(B200_FORMAT_AUTO))
	#error "the fp32 (_f) build of the B200 plugin is CSR only"
#endif

static void
b200_die(const char * what)
{
	fprintf(stderr, "b200spmv plugin: %s: %s\n", what, b200spmv_last_error());
	exit(1);
}


#if DOUBLE

struct B200_Format : Matrix_Format
{
	b200spmv_t * handle;
	int transfer_always;
	long calls;
	double convert_ms;
	double stats[7];

	B200_Format(INT_T * row_ptr, INT_T * col_ind, ValueTypeReference * values, long m, long n, long nnz) : Matrix_Format(m, n, nnz)
	{
		static_assert(sizeof(INT_T) == 4, "libb200spmv takes int32 indices (INT_T = int32_t)");
		static_assert(sizeof(ValueType) == 8 && sizeof(ValueTypeReference) == 8, "libb200spmv is the fp64 (_d) build");
		b200spmv_opts o;
		b200spmv_default_opts(&o);
		#if defined(B200_FORMAT_AUTO)
			o.format = B200SPMV_AUTO;
			format_name = (char *) "B200_AUTO";
		#elif defined(B200_FORMAT_SELL)
			o.format = B200SPMV_SELL_C_SIGMA;
			format_name = (char *) "B200_SELL_C_SIGMA";
		#else
			o.format = B200SPMV_CSR_ADAPTIVE;
			format_name = (char *) "B200_CSR_ADAPTIVE";
		#endif
		o.pin_host_vectors = 1;     /* x and y are the harness's aligned_alloc buffers, alive for the whole run (bench_spmv.cpp:598-606) */
		const char * e;
		if ((e = getenv("B200SPMV_PIN")) != NULL)
			o.pin_host_vectors = atoi(e);
		if ((e = getenv("B200SPMV_DEVICE")) != NULL)
			o.device = atoi(e);
		if ((e = getenv("B200SPMV_SELL_C")) != NULL)
			o.sell_C = atoi(e);
		if ((e = getenv("B200SPMV_SELL_SIGMA")) != NULL)
			o.sell_sigma = atoi(e);
		e = getenv("B200SPMV_TRANSFER");
		transfer_always = (e != NULL && strcmp(e, "always") == 0);
		calls = 0;
		handle = NULL;
		if (b200spmv_create(&handle, (const int32_t *) row_ptr, (const int32_t *) col_ind, (const double *) values, m, n, nnz, &o))
			b200_die("csr_to_format");
		if (b200spmv_stats(handle, stats, 7))
			b200_die("stats");
		mem_footprint = stats[0];
		convert_ms = stats[6];
		#if defined(B200_FORMAT_AUTO)
			format_name = (char *) (b200spmv_format(handle) == B200SPMV_SELL_C_SIGMA ? "B200_AUTO_SELL" : "B200_AUTO_CSR");
		#endif
	}

	~B200_Format()
	{
		b200spmv_destroy(handle);
	}

	void spmv(ValueType * x, ValueType * y);
	void statistics_start();
	int statistics_print_data(char * buf, long buf_n);
};


void
B200_Format::spmv(ValueType * x, ValueType * y)
{
	int32_t flags = 0;
	if (!transfer_always && calls > 0)
		flags = B200SPMV_SKIP_X_UPLOAD | B200SPMV_SKIP_Y_DOWNLOAD;
	if (b200spmv_spmv_host_ex(handle, (const double *) x, (double *) y, flags))    /* returns synchronised: the harness host-times the call */
		b200_die("spmv");
	calls++;
}


#else    /* ---- the `_f` build ---- */

struct B200_Format : Matrix_Format
{
	b200spmv_f32_t * handle;
	long calls;
	double stats[7];
	double convert_ms;

	B200_Format(INT_T * row_ptr, INT_T * col_ind, ValueTypeReference * values, long m, long n, long nnz) : Matrix_Format(m, n, nnz)
	{
		static_assert(sizeof(INT_T) == 4, "libb200spmv takes int32 indices (INT_T = int32_t)");
		static_assert(sizeof(ValueType) == 4 && sizeof(ValueTypeReference) == 8, "this is the fp32 (_f) build: float vectors, double matrix input");
		format_name = (char *) "B200_CSR_ADAPTIVE_F32";
		int device = -1;
		const char * e;
		if ((e = getenv("B200SPMV_DEVICE")) != NULL)
			device = atoi(e);
		calls = 0;
		handle = NULL;
		if (b200spmv_f32_create(&handle, (const int32_t *) row_ptr, (const int32_t *) col_ind, (const double *) values, m, n, nnz, device))
			b200_die("csr_to_format (f32)");
		const long ntiles = (nnz + 2047) / 2048;
		mem_footprint = (double) ntiles * 2048 * (sizeof(float) + sizeof(INT_T)) + (double) (m + 1) * sizeof(INT_T) + (double) ntiles * 12;    /* padded streams + row_ptr + tile_row / carry / tail */
		memset(stats, 0, sizeof(stats));
		stats[1] = ntiles;
		stats[3] = nnz;
		convert_ms = 0;
	}

	~B200_Format()
	{
		b200spmv_f32_destroy(handle);
	}

	void spmv(ValueType * x, ValueType * y);
	void statistics_start();
	int statistics_print_data(char * buf, long buf_n);
};


void
B200_Format::spmv(ValueType * x, ValueType * y)
{
	if (b200spmv_f32_spmv_host(handle, (const float *) x, (float *) y))    /* returns synchronised: the harness host-times the call */
		b200_die("spmv (f32)");
	calls++;
}


#endif


struct Matrix_Format *
csr_to_format(INT_T * row_ptr, INT_T * col_ind, ValueTypeReference * values, long m, long n, long nnz, long symmetric, long symmetry_expanded)
{
	if (symmetric && !symmetry_expanded)
	{
		fprintf(stderr, "symmetric matrices have to be expanded to be supported by this format\n");    /* spmv_kernels/csr.cpp:224-225 */
		exit(1);
	}
	return new B200_Format(row_ptr, col_ind, values, m, n, nnz);
}


//==========================================================================================================================================
//= Print Statistics
//==========================================================================================================================================


void
B200_Format::statistics_start()
{
	calls = 0;     /* called after the warm-up (bench_spmv.cpp:303): with transfer "first" the next call moves x / y again */
}


int
statistics_print_labels(char * buf, long buf_n)
{
	return snprintf(buf, buf_n, ",%s,%s,%s,%s", "b200_units", "b200_split", "b200_padded_or_nnz", "b200_convert_ms");
}


int
B200_Format::statistics_print_data(char * buf, long buf_n)
{
	return snprintf(buf, buf_n, ",%.0lf,%.0lf,%.0lf,%.3lf", stats[1], stats[2], stats[3], convert_ms);
}

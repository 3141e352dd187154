/*
 * coo_emul.cpp -- TEST INFRASTRUCTURE, not product code.
 *
 * The GPU COO -> CSR conversion (spmv_research_b200/csrc/coo.cu) restated step for step on the CPU around the very sort
 * networks the kernel uses (csrc/coo_sort.h compiled as plain C++): count -> exclusive scan -> scatter of
 * (column << 32 | entry) keys in a SCRAMBLED order (the device's atomics give an arbitrary one) -> per-row sort (32 emulated
 * lanes and the shuffle network for rows <= 32 keys, the same-direction bitonic network otherwise) -> emit + duplicate
 * count.  tests/test_coo_emul.py compares the result with the oracle's coo_to_csr.  Nothing in the product loads this.
 */
#define COO_HOST_EMULATION 1
#include <stdlib.h>
#include <vector>
#include "../../spmv_research_b200/csrc/coo_sort.h"

static void sort_row(uint64_t * k, int64_t n)
{
	if (n <= 32) {
		uint64_t v[32], w[32];
		for (int l = 0; l < 32; l++)
			v[l] = l < n ? k[l] : ~0ull;
		for (int kk = 2; kk <= 32; kk <<= 1)
			for (int j = kk >> 1; j > 0; j >>= 1) {
				for (int l = 0; l < 32; l++)
					w[l] = coo_bitonic32_step(v[l], v[l ^ j], l, kk, j);      /* v[l ^ j] = __shfl_xor_sync(v, j) */
				for (int l = 0; l < 32; l++)
					v[l] = w[l];
			}
		for (int l = 0; l < n; l++)
			k[l] = v[l];
		return;
	}
	int lp = 0;
	while (((int64_t) 1 << lp) < n)
		lp++;
	const int64_t half = ((int64_t) 1 << lp) >> 1;
	for (int lk = 1; lk <= lp; lk++) {
		for (int lane = 31; lane >= 0; lane--)                                /* lanes in any order within a phase */
			for (int64_t t = lane; t < half; t += 32) {
				int64_t lo, hi;
				coo_flip_pair(t, lk, &lo, &hi);
				coo_cmpx(k, lo, hi, n);
			}
		for (int ld = lk - 2; ld >= 0; ld--)
			for (int lane = 31; lane >= 0; lane--)
				for (int64_t t = lane; t < half; t += 32) {
					int64_t lo, hi;
					coo_half_pair(t, ld, &lo, &hi);
					coo_cmpx(k, lo, hi, n);
				}
	}
}

extern "C" int64_t coo_emul_to_csr(const int32_t * row, const int32_t * col, const double * val, int64_t m, int64_t n, int64_t nnz,
                                   int32_t * row_ptr, int32_t * col_out, double * val_out, uint32_t scramble)
{
	std::vector<int64_t> off(m + 1, 0);
	for (int64_t e = 0; e < nnz; e++) {
		if (row[e] < 0 || row[e] >= m || col[e] < 0 || col[e] >= n)
			return -1;
		off[row[e]]++;
	}
	int64_t run = 0;
	for (int64_t i = 0; i <= m; i++) {
		int64_t c = off[i];
		off[i] = run;
		run += c;
		row_ptr[i] = (int32_t) off[i];
	}
	std::vector<uint32_t> cursor(m > 0 ? m : 1, 0);
	std::vector<uint64_t> keys(nnz > 0 ? nnz : 1);
	/* entries visited in a scrambled order: e = (e0 * stride + shift) mod nnz with gcd(stride, nnz) = 1 */
	int64_t stride = 1;
	if (scramble && nnz > 2) {
		stride = (int64_t) (scramble % (uint32_t) nnz) | 1;
		for (;;) {
			int64_t a = stride, b = nnz;
			while (b) {
				int64_t t = a % b;
				a = b;
				b = t;
			}
			if (a == 1)
				break;
			stride += 2;
		}
	}
	for (int64_t q = 0; q < nnz; q++) {
		const int64_t e = (int64_t) (((__int128) q * stride + scramble) % nnz);
		const int64_t pos = off[row[e]] + cursor[row[e]]++;
		keys[pos] = ((uint64_t) (uint32_t) col[e] << 32) | (uint32_t) e;
	}
	int64_t dups = 0;
	for (int64_t i = 0; i < m; i++) {
		const int64_t a = row_ptr[i], len = row_ptr[i + 1] - a;
		sort_row(keys.data() + a, len);
		for (int64_t t = 0; t < len; t++) {
			const uint64_t v = keys[a + t];
			col_out[a + t] = (int32_t) (v >> 32);
			val_out[a + t] = val[(uint32_t) v];
			if (t > 0 && (v >> 32) == (keys[a + t - 1] >> 32))
				dups++;
		}
	}
	return dups;
}

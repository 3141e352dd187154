/*
 * coo_sort.h -- the per-row key sort of the GPU COO -> CSR conversion (coo.cu), written once
 * for the device and also compiled as plain C++ with emulated lanes (tests/emul/coo_emul.cpp -- test infrastructure) so
 * that the networks can be checked on the CPU.
 *
 * Keys are (column << 32) | entry index: unique, so any correct sort gives THE order "columns ascending, equal columns in
 * input order" -- the stable result the oracle's restatement of coo_to_csr (lib/storage_formats/csr/csr_gen.c:178-213)
 * produces.
 */
#pragma once
#include <stdint.h>

#if defined(__CUDACC__) && !defined(COO_HOST_EMULATION)
#define COO_FN __device__ __forceinline__
#else
#define COO_FN static inline
#endif

/* One compare-exchange step of the 32-key register bitonic sort (ascending): lane holds `self`, its partner lane ^ j holds
 * `other`; merge block size k (2 .. 32), distance j (k/2 .. 1). */
COO_FN uint64_t coo_bitonic32_step(uint64_t self, uint64_t other, int lane, int k, int j)
{
	const int up = (lane & k) == 0;             /* k = 32: every lane ascending */
	const int lower = (lane & j) == 0;
	const uint64_t lo = self < other ? self : other, hi = self < other ? other : self;
	return (lower == up) ? lo : hi;
}

/* comparator (lo, hi) of thread-index t in the FLIP stage of merge size k = 2^lk, and in a HALF-CLEANER of distance 2^ld:
 * both order (lower index, higher index) ascending, so virtual +inf keys past the end never move */
COO_FN void coo_flip_pair(int64_t t, int lk, int64_t * lo, int64_t * hi)
{
	const int64_t k = (int64_t) 1 << lk, hk = k >> 1;
	const int64_t blk = t >> (lk - 1), off = t & (hk - 1);
	*lo = blk * k + off;
	*hi = blk * k + k - 1 - off;
}

COO_FN void coo_half_pair(int64_t t, int ld, int64_t * lo, int64_t * hi)
{
	const int64_t d = (int64_t) 1 << ld;
	const int64_t blk = t >> ld, off = t & (d - 1);
	*lo = blk * 2 * d + off;
	*hi = *lo + d;
}

COO_FN void coo_cmpx(uint64_t * a, int64_t lo, int64_t hi, int64_t n)
{
	if (hi < n) {
		uint64_t x = a[lo], y = a[hi];
		if (x > y) {
			a[lo] = y;
			a[hi] = x;
		}
	}
}

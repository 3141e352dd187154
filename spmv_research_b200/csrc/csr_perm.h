/*
 * csr_perm.h -- EXPERIMENT (CSR_EXP_PERM, off in the shipped library): which thread of a tile sums which row.
 *
 * In the thread-per-row bin of csr_tile_kernel, lane l reads s_prod[a_l + j] with a_l = the row's first position in the
 * window.  The 16 lanes of a half-warp hit bank pair (a_l + j) mod 16 -- at random, ~3 wavefronts per half-warp instead of 1
 * (DESIGN.md 6).  If lane b of every half-warp only ever takes rows with a_l mod 16 == b, every step j is conflict-free.
 * So, per tile: bucket the rows the tile completes by (row start - window base) mod 16; half-warp h takes the h-th row of
 * each bucket; buckets shorter than the longest leave holes, which the tile's EMPTY rows (they read nothing) fill first.
 * Tiles whose table would exceed 2.5 slots per row keep the natural order (no table).
 * A tile then needs 16 * (longest bucket) slots instead of rows_full, and y is stored scattered over the tile's rows.
 * The sum of a row is the same sequential ascending-j sum whichever thread computes it: y is bit-identical to the shipped
 * kernel's.  Built on the host at create time (experiment only); plain C++ so that tests/ can check it without a GPU.
 *
 * tile_info[t] = (R0 | split flag in bit 31, row_ptr[R0]) as the tile plan stores it (csr_kernel.cu step 2).
 */
#pragma once
#include <stdint.h>
#include <vector>

#define CSR_PERM_HOLE 0xFFFFu
#define CSR_PERM_MIN_ROWS 128        /* = CSR_THREADS / 2: the thread-per-row bin */

/* rows the tile completes itself (its owned rows minus a tail-split last row), and its first row */
static inline int csr_perm_rows_full(const int32_t * tile_info, int64_t tile, int T, int ext_cap, int * R0_out)
{
	const int R0 = tile_info[2 * tile] & 0x7fffffff, R1 = tile_info[2 * (tile + 1)] & 0x7fffffff;
	const int nrows = R1 - R0;
	const int64_t end_own = tile_info[2 * (tile + 1) + 1];
	const int64_t over = end_own - ((int64_t) tile * T + T);
	const bool tail_split = nrows > 0 && over > ext_cap;
	*R0_out = R0;
	return tail_split ? nrows - 1 : nrows;
}

/* off: ntiles + 1 slot offsets (multiples of 16); perm: tile-local row of each slot or CSR_PERM_HOLE */
static inline void csr_perm_build(const int32_t * row_ptr, const int32_t * tile_info, int64_t ntiles, int T, int ext_cap,
                                  std::vector<int64_t> & off, std::vector<uint16_t> & perm)
{
	off.assign(ntiles + 1, 0);
	perm.clear();
	std::vector<int> bucket[16];
	std::vector<int> empties;
	for (int64_t t = 0; t < ntiles; t++) {
		int R0;
		const int rows_full = csr_perm_rows_full(tile_info, t, T, ext_cap, &R0);
		off[t] = (int64_t) perm.size();
		if (rows_full <= CSR_PERM_MIN_ROWS || rows_full >= (int) CSR_PERM_HOLE)
			continue;                                  /* other bins (or too many rows for 16-bit slots): no table, off[t+1] == off[t] */
		const int64_t base = t * (int64_t) T;
		for (int b = 0; b < 16; b++)
			bucket[b].clear();
		empties.clear();
		size_t longest = 0;
		for (int r = 0; r < rows_full; r++) {
			const int64_t a = row_ptr[R0 + r], e = row_ptr[R0 + r + 1];
			if (a == e) {
				empties.push_back(r);
				continue;
			}
			std::vector<int> & B = bucket[(a - base) & 15];
			B.push_back(r);
			longest = B.size() > longest ? B.size() : longest;
		}
		size_t ne = 0, holes = 0;
		for (int b = 0; b < 16; b++)
			holes += longest - bucket[b].size();
		const size_t extra = empties.size() > holes ? (empties.size() - holes + 15) / 16 : 0;       /* half-warps of empty rows only */
		/* lopsided buckets (e.g. every row 8 long: two residues only) would give a table of mostly holes: keep the natural
		 * order for such a tile (no table: the kernel falls back to row = thread) */
		if (2 * 16 * (longest + extra) > 5 * (size_t) rows_full)
			continue;
		for (size_t h = 0; h < longest + extra; h++)
			for (int b = 0; b < 16; b++) {
				uint16_t v = CSR_PERM_HOLE;
				if (h < bucket[b].size())
					v = (uint16_t) bucket[b][h];
				else if (ne < empties.size())
					v = (uint16_t) empties[ne++];
				perm.push_back(v);
			}
	}
	off[ntiles] = (int64_t) perm.size();
}

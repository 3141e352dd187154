/*
 * csr_perm.h -- EXPERIMENT (CSR_EXP_PERM, off in the shipped library): which thread of a tile sums which row.
 *
 * In the thread-per-row bin of csr_tile_kernel, lane l reads s_prod[a_l + j] with a_l = the row's first position in the
 * window.  The 16 lanes of a half-warp hit bank pair (a_l + j) mod 16 -- at random, ~3 wavefronts per half-warp instead of 1
 * (DESIGN.md 6).  If lane b of every half-warp only ever takes rows with a_l mod 16 == b, every step j is conflict-free.
 * So, per tile: bucket the rows the tile completes by (row start - window base) mod 16; half-warp h takes the h-th row of
 * each bucket; buckets shorter than the longest leave holes, which the tile's EMPTY rows (they read nothing) fill first.
 * Whether a tile gets a table, and in which of two orders, is decided by a wavefront model (csr_perm_cost below).
 * A tile then needs 16 * (longest bucket) slots instead of rows_full, and y is stored scattered over the tile's rows.
 * The sum of a row is the same sequential ascending-j sum whichever thread computes it: y is bit-identical to the shipped
 * kernel's.  Built on the host at create time (experiment only); plain C++ so that tests/ can check it without a GPU.
 *
 * tile_info[t] = (R0 | split flag in bit 31, row_ptr[R0]) as the tile plan stores it (csr_kernel.cu step 2).
 */
#pragma once
#include <stdint.h>
#include <algorithm>
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

/* Modelled L1 data-pipe wavefronts of one tile's row sums for a given slot order (slot -> tile-local row or HOLE):
 * per warp and per step j of its longest row, each half-warp costs the highest multiplicity of a bank pair
 * ((start + j) mod 16) among its active lanes (64-bit shared loads are served per half-warp); plus one wavefront per
 * distinct 128-byte line of y the warp's store touches. */
static inline int64_t csr_perm_cost(const int32_t * row_ptr, int R0, int64_t base, const uint16_t * slot, size_t nslots)
{
	int64_t cost = 0;
	for (size_t w = 0; w < nslots; w += 32) {
		const size_t we = w + 32 < nslots ? w + 32 : nslots;
		int64_t a[32], len[32], steps = 0, lines[32];
		int nl = 0;
		for (size_t s = w; s < we; s++) {
			const size_t l = s - w;
			if (slot[s] == CSR_PERM_HOLE) {
				a[l] = len[l] = 0;
				continue;
			}
			const int64_t i = R0 + (int64_t) slot[s];
			a[l] = row_ptr[i] - base;
			len[l] = row_ptr[i + 1] - row_ptr[i];
			steps = len[l] > steps ? len[l] : steps;
			const int64_t line = i >> 4;
			int k = 0;
			while (k < nl && lines[k] != line)
				k++;
			if (k == nl)
				lines[nl++] = line;
		}
		cost += nl;
		for (int64_t j = 0; j < steps; j++)
			for (size_t h = w; h < we; h += 16) {
				int cnt[16] = {0}, mx = 0;
				for (size_t s = h; s < h + 16 && s < we; s++)
					if (len[s - w] > j) {
						const int c = ++cnt[(a[s - w] + j) & 15];
						mx = c > mx ? c : mx;
					}
				cost += mx;
			}
	}
	return cost;
}

/* off: ntiles + 1 slot offsets (multiples of 16); perm: tile-local row of each slot or CSR_PERM_HOLE.
 * Per tile three orders are costed with csr_perm_cost -- the natural one (row = thread, no table), the bucket table with rows
 * in index order (y stores stay local) and with rows longest-first (half-warps of similar length) -- and the cheapest kept.
 * stats (optional, 4 entries): tiles of the bin, tiles given a table, modelled cost natural / chosen. */
static inline void csr_perm_build(const int32_t * row_ptr, const int32_t * tile_info, int64_t ntiles, int T, int ext_cap,
                                  std::vector<int64_t> & off, std::vector<uint16_t> & perm, int64_t * stats = nullptr)
{
	off.assign(ntiles + 1, 0);
	perm.clear();
	std::vector<int> bucket[16];
	std::vector<int> empties;
	std::vector<uint16_t> cand[3];
	int64_t st[4] = {0, 0, 0, 0};
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
		size_t holes = 0;
		for (int b = 0; b < 16; b++)
			holes += longest - bucket[b].size();
		const size_t extra = empties.size() > holes ? (empties.size() - holes + 15) / 16 : 0;       /* half-warps of empty rows only */
		cand[0].resize(rows_full);
		for (int r = 0; r < rows_full; r++)
			cand[0][r] = (uint16_t) r;
		const bool table_ok = 16 * (longest + extra) <= 4 * (size_t) rows_full;                     /* bound the table's size */
		for (int c = 1; c <= 2 && table_ok; c++) {
			if (c == 2)
				for (int b = 0; b < 16; b++)
					std::stable_sort(bucket[b].begin(), bucket[b].end(), [&](int p, int q) {
						return row_ptr[R0 + p + 1] - row_ptr[R0 + p] > row_ptr[R0 + q + 1] - row_ptr[R0 + q];
					});
			cand[c].clear();
			size_t ne = 0;
			for (size_t h = 0; h < longest + extra; h++)
				for (int b = 0; b < 16; b++) {
					uint16_t v = CSR_PERM_HOLE;
					if (h < bucket[b].size())
						v = (uint16_t) bucket[b][h];
					else if (ne < empties.size())
						v = (uint16_t) empties[ne++];
					cand[c].push_back(v);
				}
		}
		int best = 0;
		int64_t cost[3];
		cost[0] = csr_perm_cost(row_ptr, R0, base, cand[0].data(), cand[0].size());
		for (int c = 1; c <= 2 && table_ok; c++) {
			cost[c] = csr_perm_cost(row_ptr, R0, base, cand[c].data(), cand[c].size());
			if (cost[c] < cost[best])
				best = c;
		}
		st[0]++;
		st[2] += cost[0];
		st[3] += cost[best];
		if (best == 0)
			continue;                                  /* the natural order is the cheapest: no table for this tile */
		st[1]++;
		perm.insert(perm.end(), cand[best].begin(), cand[best].end());
	}
	off[ntiles] = (int64_t) perm.size();
	if (stats)
		for (int k = 0; k < 4; k++)
			stats[k] = st[k];
}

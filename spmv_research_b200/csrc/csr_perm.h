/*
 * csr_perm.h -- EXPERIMENT (CSR_EXP_PERM, off in the shipped library): which thread of a tile sums which row.
 *
 * In the thread-per-row bin of csr_tile_kernel, lane l reads s_prod[a_l + j] with a_l = the row's first position in the
 * window.  The 16 lanes of a half-warp hit bank pair (a_l + j) mod 16 -- at random, ~3 wavefronts per half-warp instead of 1
 * (DESIGN.md 6).  If lane b of every half-warp only ever takes rows with a_l mod 16 == b, every step j is conflict-free.
 * So, per tile: bucket the rows the tile completes by (row start - window base) mod 16; half-warp h takes the h-th row of
 * each bucket; buckets shorter than the longest leave holes, which the tile's EMPTY rows (they read nothing) fill first.
 * Whether a tile gets a table, and in which of two orders, is decided by a wavefront model (csr_perm_cost below).
 * The bins with 2 / 4 / 8 lanes per row get the same treatment with groups of lanes in place of lanes (see csr_perm_build).
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

/* lanes per row of a tile that completes rows_full rows: the kernel's binning rule (csr_kernel.cu step 5), 1 = thread per row */
static inline int csr_perm_lpr(int rows_full)
{
	if (rows_full > CSR_PERM_MIN_ROWS)
		return 1;
	int lpr = 2;
	while (lpr < 32 && rows_full * (lpr * 2) <= 256)
		lpr *= 2;
	return lpr;
}

/* Modelled L1 data-pipe wavefronts of one tile's row sums for a given order of its GROUP slots (a group = lpr adjacent
 * lanes summing one row, lane l reading positions start + l + k * lpr; slot -> tile-local row or HOLE): per warp and per
 * step k of its longest row, each half-warp costs the highest multiplicity of a bank pair (position mod 16) among its active
 * lanes (64-bit shared loads are served per half-warp); plus one wavefront per distinct 128-byte line of y the warp stores. */
static inline int64_t csr_perm_cost(const int32_t * row_ptr, int R0, int64_t base, const uint16_t * slot, size_t nslots, int lpr)
{
	const size_t gw = 32 / lpr, gh = 16 / lpr;              /* groups per warp / per half-warp */
	int64_t cost = 0;
	for (size_t w = 0; w < nslots; w += gw) {
		const size_t we = w + gw < nslots ? w + gw : nslots;
		int64_t a[32], len[32], steps = 0, lines[32];
		int nl = 0;
		for (size_t s = w; s < we; s++) {
			const size_t g = s - w;
			if (slot[s] == CSR_PERM_HOLE) {
				a[g] = len[g] = 0;
				continue;
			}
			const int64_t i = R0 + (int64_t) slot[s];
			a[g] = row_ptr[i] - base;
			len[g] = row_ptr[i + 1] - row_ptr[i];
			const int64_t st = (len[g] + lpr - 1) / lpr;
			steps = st > steps ? st : steps;
			const int64_t line = i >> 4;
			int k = 0;
			while (k < nl && lines[k] != line)
				k++;
			if (k == nl)
				lines[nl++] = line;
		}
		cost += nl;
		for (int64_t k = 0; k < steps; k++)
			for (size_t h = w; h < we; h += gh) {
				int cnt[16] = {0}, mx = 0;
				for (size_t s = h; s < h + gh && s < we; s++)
					for (int l = 0; l < lpr; l++) {
						const int64_t pos = l + k * lpr;
						if (pos < len[s - w]) {
							const int c = ++cnt[(a[s - w] + pos) & 15];
							mx = c > mx ? c : mx;
						}
					}
				cost += mx;
			}
	}
	return cost;
}

/* off: ntiles + 1 slot offsets; perm: tile-local row of each group slot or CSR_PERM_HOLE.
 * With lpr lanes per row (1, 2, 4 or 8; bins with 16 or 32 are conflict-free as they are) a half-warp sums G = 16 / lpr rows
 * and group windows of lpr adjacent bank pairs never collide when the G rows start at the same offset c = start mod lpr and
 * at G different values of q = (start / lpr) mod G.  So the rows are bucketed by (c, q) -- for lpr = 1 simply start mod 16 --
 * and half-warp h of class c takes the h-th row of each of its G buckets.
 * Per tile three orders are costed with csr_perm_cost -- the natural one (no table), the bucket table with rows in index
 * order (y stores stay local) and with rows longest-first (half-warps of similar length) -- and the cheapest kept.
 * stats (optional, 4 entries): tiles considered, tiles given a table, modelled cost natural / chosen.
 * lanes_bins_only: leave the thread-per-row bin alone (the build that rotates those rows instead, CSR_EXP_ROT). */
static inline void csr_perm_build(const int32_t * row_ptr, const int32_t * tile_info, int64_t ntiles, int T, int ext_cap,
                                  std::vector<int64_t> & off, std::vector<uint16_t> & perm, int64_t * stats = nullptr, bool lanes_bins_only = false)
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
		if (rows_full < 1 || rows_full >= (int) CSR_PERM_HOLE)
			continue;                                  /* nothing to order (or too many rows for 16-bit slots): off[t+1] == off[t] */
		const int lpr = csr_perm_lpr(rows_full);
		if (lpr > 8 || (lpr == 1 && lanes_bins_only))
			continue;                                  /* one row per half-warp or per warp: adjacent positions, no conflicts */
		const int G = 16 / lpr;
		const int64_t base = t * (int64_t) T;
		for (int b = 0; b < 16; b++)
			bucket[b].clear();
		empties.clear();
		for (int r = 0; r < rows_full; r++) {
			const int64_t a = row_ptr[R0 + r] - base, e = row_ptr[R0 + r + 1] - base;
			if (a == e) {
				empties.push_back(r);
				continue;
			}
			bucket[(a % lpr) * G + (a / lpr) % G].push_back(r);
		}
		size_t halfwarps = 0, holes = 0, H[8];
		for (int c = 0; c < lpr; c++) {
			H[c] = 0;
			for (int q = 0; q < G; q++)
				H[c] = bucket[c * G + q].size() > H[c] ? bucket[c * G + q].size() : H[c];
			for (int q = 0; q < G; q++)
				holes += H[c] - bucket[c * G + q].size();
			halfwarps += H[c];
		}
		const size_t extra = empties.size() > holes ? (empties.size() - holes + G - 1) / G : 0;     /* half-warps of empty rows only */
		cand[0].resize(rows_full);
		for (int r = 0; r < rows_full; r++)
			cand[0][r] = (uint16_t) r;
		const bool table_ok = G * (halfwarps + extra) <= 4 * (size_t) rows_full;                    /* bound the table's size */
		for (int v = 1; v <= 2 && table_ok; v++) {
			if (v == 2)
				for (int b = 0; b < 16; b++)
					std::stable_sort(bucket[b].begin(), bucket[b].end(), [&](int p, int q) {
						return row_ptr[R0 + p + 1] - row_ptr[R0 + p] > row_ptr[R0 + q + 1] - row_ptr[R0 + q];
					});
			cand[v].clear();
			size_t ne = 0;
			for (int c = 0; c <= lpr; c++) {               /* c == lpr: the extra half-warps (empty rows only) */
				const size_t Hc = c < lpr ? H[c] : extra;
				for (size_t h = 0; h < Hc; h++)
					for (int q = 0; q < G; q++) {
						uint16_t e = CSR_PERM_HOLE;
						if (c < lpr && h < bucket[c * G + q].size())
							e = (uint16_t) bucket[c * G + q][h];
						else if (ne < empties.size())
							e = (uint16_t) empties[ne++];
						cand[v].push_back(e);
					}
			}
		}
		int best = 0;
		int64_t cost[3];
		cost[0] = csr_perm_cost(row_ptr, R0, base, cand[0].data(), cand[0].size(), lpr);
		for (int v = 1; v <= 2 && table_ok; v++) {
			cost[v] = csr_perm_cost(row_ptr, R0, base, cand[v].data(), cand[v].size(), lpr);
			if (cost[v] < cost[best])
				best = v;
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

/* ---------------------------------------------------------------------------------------------------------------------
 * CSR_EXP_ROT -- second experiment on the same conflicts, thread-per-row bin only: rows keep their threads (y stays
 * coalesced, no holes), but thread l starts its row at element phi_l and wraps around: at step j it reads
 * s_prod[a_l + (j + phi_l) mod len_l].  phi is chosen per half-warp (16 consecutive tile-local rows), lane by lane, as the
 * rotation that collides least with the lanes already placed, counted over every step (greedy).  The row sum becomes a
 * cyclic rotation of the ascending-j sum: deterministic and inside the per-row bound, but no longer bit-equal to the
 * sequential order.  rot: one byte per row of the plan's row space (0 outside the bin).  stats: modelled row-sum
 * wavefronts natural / rotated.
 */
static inline void csr_rot_build(const int32_t * row_ptr, const int32_t * tile_info, int64_t ntiles, int64_t m, int T, int ext_cap,
                                 std::vector<uint8_t> & rot, int64_t * stats = nullptr)
{
	rot.assign((size_t) (m > 0 ? m : 1), 0);
	int64_t st[2] = {0, 0};
	std::vector<int> occ;
	for (int64_t t = 0; t < ntiles; t++) {
		int R0;
		const int rows_full = csr_perm_rows_full(tile_info, t, T, ext_cap, &R0);
		if (rows_full <= CSR_PERM_MIN_ROWS)
			continue;
		const int64_t base = t * (int64_t) T;
		for (int h = 0; h < rows_full; h += 16) {
			const int he = h + 16 < rows_full ? h + 16 : rows_full;
			int64_t steps = 0;
			for (int r = h; r < he; r++) {
				const int64_t len = row_ptr[R0 + r + 1] - row_ptr[R0 + r];
				steps = len > steps ? len : steps;
			}
			if (steps == 0)
				continue;
			/* occ[j * 16 + bank pair]: lanes placed so far that read this bank pair at step j; nat: the same with phi = 0 */
			occ.assign((size_t) steps * 32, 0);
			int * nat = occ.data() + steps * 16;
			for (int r = h; r < he; r++) {
				const int64_t a = row_ptr[R0 + r] - base, len = row_ptr[R0 + r + 1] - row_ptr[R0 + r];
				if (len == 0)
					continue;
				int best_phi = 0;
				int64_t best = -1;
				const int64_t tries = len < 16 ? len : 16;              /* rotations beyond 15 repeat the bank offsets */
				for (int64_t p = 0; p < tries; p++) {
					int64_t c = 0;
					for (int64_t j = 0; j < len; j++)
						c += occ[(size_t) j * 16 + ((a + (j + p) % len) & 15)];
					if (best < 0 || c < best) {
						best = c;
						best_phi = (int) p;
					}
				}
				rot[(size_t) R0 + r] = (uint8_t) best_phi;
				for (int64_t j = 0; j < len; j++) {
					occ[(size_t) j * 16 + ((a + (j + best_phi) % len) & 15)]++;
					nat[(size_t) j * 16 + ((a + j) & 15)]++;
				}
			}
			for (int64_t j = 0; j < steps; j++) {
				int m0 = 0, m1 = 0;
				for (int b = 0; b < 16; b++) {
					m0 = nat[j * 16 + b] > m0 ? nat[j * 16 + b] : m0;
					m1 = occ[(size_t) j * 16 + b] > m1 ? occ[(size_t) j * 16 + b] : m1;
				}
				st[0] += m0;
				st[1] += m1;
			}
		}
	}
	if (stats) {
		stats[0] = st[0];
		stats[1] = st[1];
	}
}

/*
 * f32_tile.h -- tile geometry and row classification of the fp32 CSR kernel (f32.cu; SURVEY.md section 8f row N4: the
 * reference's `_f` build axis, make.sh:215-216), written once for the device and also compiled as plain C++ by
 * tests/emul/f32_emul.cpp (test infrastructure) so that the bookkeeping -- which row a tile finishes, which pieces go to
 * the carry / tail partials, what the fix-up adds in which order -- is checked on the CPU.
 *
 * The nnz stream is cut into windows of T non-zeros, one CTA per window.  tile_row[t] = the row that contains position
 * t * T (the last r with row_ptr[r] <= t * T).  Window t touches the rows [r_first, r_last]; of a row r it holds the
 * positions [max(row_ptr[r], base), min(row_ptr[r + 1], lim)):
 *   - the row starts and ends inside the window          -> y[r] = sum                      (empty rows inside: 0)
 *   - it started in an earlier window (only r_first can) -> carry[t] = sum of the piece
 *   - it starts here and ends later (only r_last can)    -> tail[t]  = sum of the piece
 * and the fix-up completes every row that starts in t and ends later: y[r] = tail[t] + carry[t + 1] + carry[t + 2] + ...
 * up to the window in which the row ends, in window order (deterministic, no atomics, no memset of y).  Rows before
 * tile_row[0] are empty and are zeroed by window 0.
 */
#pragma once
#include <stdint.h>

#if defined(__CUDACC__) && !defined(F32_HOST_EMULATION)
#define F32_FN __device__ __forceinline__
#else
#define F32_FN static inline
#endif

typedef struct {
	int64_t base, lim;          /* window = positions [base, lim) */
	int64_t r_first, r_last;    /* rows touched */
} f32_geom;

F32_FN void f32_geometry(const int32_t * row_ptr, const int32_t * tile_row, int64_t tile, int64_t ntiles, int64_t m, int64_t nnz, int T,
                         f32_geom * g)
{
	g->base = tile * (int64_t) T;
	g->lim = g->base + T < nnz ? g->base + T : nnz;
	g->r_first = tile_row[tile];
	if (tile + 1 < ntiles) {
		const int64_t r_next = tile_row[tile + 1];          /* contains position lim */
		g->r_last = (row_ptr[r_next] == g->lim) ? r_next - 1 : r_next;
	} else
		g->r_last = m - 1;
}

/* tile_row of the window starting at position pos: the last r in [0, m) with row_ptr[r] <= pos (pos < nnz) */
F32_FN int32_t f32_tile_row_of(const int32_t * row_ptr, int64_t m, int64_t pos)
{
	int64_t lo = 0, hi = m + 1;                 /* first index in [0, m] with row_ptr[idx] > pos */
	while (lo < hi) {
		const int64_t mid = (lo + hi) >> 1;
		if (row_ptr[mid] > pos)
			hi = mid;
		else
			lo = mid + 1;
	}
	return (int32_t) (lo - 1);
}

/* piece of row r inside the window: positions [*a, *b) RELATIVE to base; returns 1 = whole row here (store y), 2 = carry
 * (the row started earlier), 3 = tail (starts here, ends later) */
F32_FN int f32_row_piece(const int32_t * row_ptr, const f32_geom * g, int64_t r, int * a, int * b)
{
	const int64_t s = row_ptr[r], e = row_ptr[r + 1];
	*a = (int) ((s > g->base ? s : g->base) - g->base);
	*b = (int) ((e < g->lim ? e : g->lim) - g->base);
	if (s < g->base)
		return 2;
	return e <= g->lim ? 1 : 3;
}

/* fix-up of window t: returns 1 and the row / its completed sum when a row starts in t and ends in a later window */
F32_FN int f32_fixup(const int32_t * row_ptr, const int32_t * tile_row, const float * carry, const float * tail, int64_t tile,
                     int64_t ntiles, int64_t m, int64_t nnz, int T, int64_t * row_out, float * sum_out)
{
	f32_geom g;
	f32_geometry(row_ptr, tile_row, tile, ntiles, m, nnz, T, &g);
	const int64_t r = g.r_last;
	if (r < g.r_first || row_ptr[r] < g.base || row_ptr[r + 1] <= g.lim)
		return 0;
	const int64_t e = row_ptr[r + 1];
	float acc = tail[tile];
	for (int64_t u = tile + 1; u < ntiles; u++) {
		acc += carry[u];
		const int64_t lim_u = (u + 1) * (int64_t) T < nnz ? (u + 1) * (int64_t) T : nnz;
		if (e <= lim_u)
			break;
	}
	*row_out = r;
	*sum_out = acc;
	return 1;
}

/* Row reduction bins.  >= F32_WARP_ROWS_BELOW rows in a window: one thread per row (sequential ascending-j sum).  Fewer: LPR
 * adjacent lanes per row, LPR = the largest power of two in [4, 32] with rows * LPR <= 256 (every row of the window is then
 * reduced in one pass of the CTA); lane l of a group sums the elements a + l, a + l + LPR, ... of its row, then the group
 * combines with the butterfly of __shfl_xor_sync (offsets LPR / 2, ..., 1): group lane 0 holds the sum. */
#define F32_WARP_ROWS_BELOW 64
F32_FN int f32_lanes_per_row(int64_t nrows)
{
	int lpr = 4;
	while (lpr < 32 && nrows * (lpr * 2) <= 256)
		lpr *= 2;
	return lpr;
}

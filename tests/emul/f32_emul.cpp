/*
 * f32_emul.cpp -- TEST INFRASTRUCTURE, not product code.
 *
 * The fp32 CSR product of spmv_research_b200/csrc/f32.cu replayed on the CPU with the kernel's own bookkeeping
 * (csrc/f32_tile.h compiled as plain C++): narrowing, tile_row search, one "CTA" per window with the 256 threads / 8 warps
 * executed in turn (products to the window buffer; thread-per-row or lanes-per-row reduction with the butterfly order of
 * __shfl_xor_sync; y / carry / tail stores), then the fix-up.  Every float operation happens in the device's order, so
 * the result is what the GPU must produce bit for bit.  writes[r] counts the stores to y[r] (must be exactly 1).
 * Build with -ffp-contract=off (the device code uses __fmul_rn / __fadd_rn: no FMA).
 */
#define F32_HOST_EMULATION 1
#include <stdlib.h>
#include <vector>
#include "../../spmv_research_b200/csrc/f32_tile.h"

static const int THREADS = 256, NPT = 8, T = THREADS * NPT;

extern "C" int f32_emul_spmv(const int32_t * row_ptr, const int32_t * col_ind, const double * values, int64_t m, int64_t n, int64_t nnz,
                             const float * x, float * y, int32_t * writes)
{
	(void) n;
	const int64_t ntiles = (nnz + T - 1) / T;
	for (int64_t r = 0; r < m; r++)
		writes[r] = 0;
	if (m == 0)
		return 0;
	if (ntiles == 0) {
		for (int64_t r = 0; r < m; r++) {
			y[r] = 0.0f;
			writes[r]++;
		}
		return 0;
	}
	std::vector<int32_t> col((size_t) ntiles * T, 0), tile_row(ntiles);
	std::vector<float> val((size_t) ntiles * T, 0.0f), carry(ntiles, -7.0f), tail(ntiles, -7.0f), s_prod(T);
	for (int64_t j = 0; j < nnz; j++) {
		col[j] = col_ind[j];
		val[j] = (float) values[j];
	}
	for (int64_t t = 0; t < ntiles; t++)
		tile_row[t] = f32_tile_row_of(row_ptr, m, t * (int64_t) T);
	for (int64_t tile = 0; tile < ntiles; tile++) {
		f32_geom g;
		f32_geometry(row_ptr, tile_row.data(), tile, ntiles, m, nnz, T, &g);
		for (int tid = 0; tid < THREADS; tid++)
			for (int k = 0; k < NPT; k++) {
				const int64_t idx = g.base + k * THREADS + tid;
				s_prod[k * THREADS + tid] = val[idx] * x[col[idx]];
			}
		if (tile == 0)
			for (int64_t r = 0; r < g.r_first; r++) {
				y[r] = 0.0f;
				writes[r]++;
			}
		const int64_t nrows = g.r_last - g.r_first + 1;
		if (nrows >= F32_WARP_ROWS_BELOW) {
			for (int tid = 0; tid < THREADS; tid++)
				for (int64_t r = g.r_first + tid; r <= g.r_last; r += THREADS) {
					int a, b;
					const int kind = f32_row_piece(row_ptr, &g, r, &a, &b);
					float acc = 0.0f;
					for (int j = a; j < b; j++)
						acc = acc + s_prod[j];
					if (kind == 1) {
						y[r] = acc;
						writes[r]++;
					} else if (kind == 2)
						carry[tile] = acc;
					else
						tail[tile] = acc;
				}
		} else {
			const int lpr = f32_lanes_per_row(nrows);
			const int groups = THREADS / lpr;
			for (int64_t rb = g.r_first; rb <= g.r_last; rb += groups)
				for (int grp = 0; grp < groups; grp++) {
					const int64_t r = rb + grp;
					if (r > g.r_last)
						continue;
					int a, b;
					const int kind = f32_row_piece(row_ptr, &g, r, &a, &b);
					float acc[32], nxt[32];
					for (int lane = 0; lane < lpr; lane++) {
						acc[lane] = 0.0f;
						for (int j = a + lane; j < b; j += lpr)
							acc[lane] = acc[lane] + s_prod[j];
					}
					for (int o = lpr >> 1; o > 0; o >>= 1) {
						for (int lane = 0; lane < lpr; lane++)
							nxt[lane] = acc[lane] + acc[lane ^ o];
						for (int lane = 0; lane < lpr; lane++)
							acc[lane] = nxt[lane];
					}
					if (kind == 1) {
						y[r] = acc[0];
						writes[r]++;
					} else if (kind == 2)
						carry[tile] = acc[0];
					else
						tail[tile] = acc[0];
				}
		}
	}
	for (int64_t t = 0; t < ntiles; t++) {
		int64_t r;
		float s;
		if (f32_fixup(row_ptr, tile_row.data(), carry.data(), tail.data(), t, ntiles, m, nnz, T, &r, &s)) {
			y[r] = s;
			writes[r]++;
		}
	}
	return 0;
}

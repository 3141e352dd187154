// sell_kernel.cu -- fp64 SELL-C-sigma SpMV for sm_100a.
//
// Semantics of the reference kernel (K/sell-C-s/RISC-V/sellcs_mv_kernels_epi.c:178-257): per
// slice, acc[lane] += val[k*C+lane] * x[col[k*C+lane]] for k < slice_width, then
// y[row_order[slice*C+lane]] = acc[lane]; the ragged last slice stores min(C, m - s*C) lanes.
//
// B200 mapping: the unit of work is an ITEM = (slice, group of 32 lanes, chunk of SELL_KCH slice
// columns), one warp per item, lane = sorted row.  Each k step of a warp is one 256-byte values
// load and one 128-byte col_ind load (column-major slices are coalesced by construction), eight
// steps in flight per lane before the first x gather; no reduction is needed because a lane is a
// row.  Cutting slices along k bounds the sequential depth of a lane: a slice holding one of the
// generator's giant skew rows (width 500 100 at config 4) becomes thousands of items instead of
// one 500 100-step loop.  Slices cut into several items write per-item partials that the last-arriving
// warp of the cut adds in chunk order (deterministic, no atomics on y; no second launch).  Padding entries are
// (col 0, val 0.0) exactly as the reference writes them and are multiplied like real ones.
#include "common.cuh"

template <bool HINT, int U, int MINB>
__global__ void __launch_bounds__(SELL_THREADS, MINB)
sell_kernel(const int4 * __restrict__ items, int64_t nitems, const int32_t * __restrict__ row_order,
            const int32_t * __restrict__ scol, const double * __restrict__ sval, const double * __restrict__ x,
            double * __restrict__ partial, int64_t m, int C, YOut out, const int32_t * __restrict__ slot_cut,
            const int4 * __restrict__ cuts, unsigned int * __restrict__ cut_cnt, int64_t norm_cut_base)
{
	__shared__ double s_red[SELL_THREADS / 32];
	const uint64_t pol = make_policy_evict_last();
	const uint64_t pol_stream = make_policy_evict_first();
	const int lane = threadIdx.x & 31;
	const int64_t it = (int64_t) blockIdx.x * (SELL_THREADS / 32) + (threadIdx.x >> 5);
	double nrm = 0.0;
	if (it < nitems) {
		const int4 item = items[it];                 // packed by sell_item_fill_kernel: offset, column count, slot, first sorted row
		int kc = (int) ((uint32_t) item.y >> 16);
		if (lane >= C)
			kc = 0;                                  // C < 32: the upper lanes of the warp idle
		int64_t p = ((int64_t) (item.y & 0xffff) << 32) + (int64_t) (uint32_t) item.x + lane;
		const int64_t r = (int64_t) item.w + lane;
		// the destination row travels with the first stream loads instead of after the last gather
		int dst_row = 0;
		if (item.z < 0 && r < m && lane < C)
			dst_row = row_order[r];
		double acc = 0.0;
		int k = 0;
		for (; k + U <= kc; k += U) {
			double v[U];
			int c[U];
			#pragma unroll
			for (int u = 0; u < U; u++) {
				v[u] = ld_stream_f64(sval + p + (int64_t) u * C, pol_stream);
				c[u] = ld_stream_i32(scol + p + (int64_t) u * C, pol_stream);
			}
			#pragma unroll
			for (int u = 0; u < U; u++)
				acc = fma(v[u], ld_x<HINT>(x + c[u], pol), acc);
			p += U * (int64_t) C;
		}
		if (k < kc) {
			double v[U];
			int c[U];
			#pragma unroll
			for (int u = 0; u < U; u++) {
				const bool on = (k + u < kc);
				v[u] = on ? ld_stream_f64(sval + p + (int64_t) u * C, pol_stream) : 0.0;
				c[u] = on ? ld_stream_i32(scol + p + (int64_t) u * C, pol_stream) : 0;
			}
			#pragma unroll
			for (int u = 0; u < U; u++)
				if (k + u < kc)
					acc = fma(v[u], ld_x<HINT>(x + c[u], pol), acc);
		}
		if (item.z >= 0) {
			// a cut slice: publish this item's partial; the warp that arrives LAST at the cut adds all of them in chunk order
			// (a fixed order: y does not depend on which warp that was -- deterministic, no atomics on y) and finishes the rows.
			// Round 1 did this in a second launch (sell_fixup_kernel, ~5 us + a launch gap per product).
			partial[(int64_t) item.z * 32 + lane] = acc;
			const int q = slot_cut[item.z];
			const int4 cut = cuts[q];                    // x: slice, y: first partial slot, z: number of items, w: lane group
			__threadfence();
			__syncwarp();                                // every lane's partial is fenced before lane 0 counts the item in
			unsigned int old = 0;
			if (lane == 0)
				old = atomicAdd(&cut_cnt[q], 1u);
			old = __shfl_sync(0xffffffffu, old, 0);
			if (old + 1u == (unsigned int) cut.z) {
				__threadfence();
				double a = 0.0;
				for (int j = 0; j < cut.z; j++)
					a += __ldcg(partial + ((int64_t) cut.y + j) * 32 + lane);
				const int64_t rr = (int64_t) cut.x * C + cut.w * 32 + lane;
				double sq = 0.0;
				if (rr < m && lane < C) {
					const double scale = out.scale ? *out.scale : 1.0;
					a = y_finish(out, row_order[rr], a, scale);
					sq = a * a;
				}
				if (out.norm_partial) {
					sq = warp_sum(sq);
					if (lane == 0)
						out.norm_partial[norm_cut_base + q] = sq;      // one slot per cut: the norm's grouping is fixed too
				}
				if (lane == 0)
					cut_cnt[q] = 0u;
			}
		} else if (r < m && lane < C) {
			const double scale = out.scale ? *out.scale : 1.0;
			acc = y_finish(out, dst_row, acc, scale);
			nrm = acc * acc;
		}
	}
	if (out.norm_partial) {
		nrm = warp_sum(nrm);
		if (lane == 0)
			s_red[threadIdx.x >> 5] = nrm;
		__syncthreads();
		if (threadIdx.x == 0) {
			double tot = 0.0;
			#pragma unroll
			for (int wv = 0; wv < SELL_THREADS / 32; wv++)
				tot += s_red[wv];
			out.norm_partial[blockIdx.x] = tot;
		}
	}
}

int launch_sell_spmv(b200spmv * h, const double * dx, double * dy, cudaStream_t st, const b200spmv_launch * l)
{
	YOut out;
	memset(&out, 0, sizeof(out));
	out.y = dy;
	if (l) {
		if (l->n_peer < 0 || l->n_peer > 8)
			B200_FAIL("n_peer must be in [0, 8]");
		out.n_peer = l->n_peer;
		for (int d = 0; d < l->n_peer; d++)
			out.peer[d] = l->y_peer[d];
		out.peer_lo = 0;
		out.peer_hi = h->m;
		if (l->peer_row_lo != 0 || l->peer_row_hi != 0) {
			out.peer_lo = l->peer_row_lo;
			out.peer_hi = l->peer_row_hi;
		}
		out.scale = l->scale;
		out.norm_partial = l->norm2 ? h->d_norm_partial : nullptr;
	}
	out.accum = h->acc_mode & 1;
	out.partial = (h->acc_mode >> 1) & 1;
	constexpr int WPC = SELL_THREADS / 32;
	const int64_t first = h->range_count >= 0 ? h->range_first : 0;
	const int64_t nit = h->range_count >= 0 ? h->range_count : h->sell_nitems;
	const int64_t ctas = (nit + WPC - 1) / WPC;
	const int64_t ncuts = h->sell_ncuts;
	if (ctas > 0) {
		const int4 * items = reinterpret_cast<const int4 *>(h->d_sell_items) + first;
		#define SELL_LAUNCH(HINT, U, MINB)                                                                                          \
			sell_kernel<HINT, U, MINB><<<(unsigned) ctas, SELL_THREADS, 0, st>>>(items, nit, h->d_row_order, h->d_scol, \
			                                                                     h->d_sval, dx, h->d_sell_partial, h->m, h->C, out, h->d_sell_slot_cut, \
			                                                                     reinterpret_cast<const int4 *>(h->d_sell_cuts), h->d_sell_cut_cnt, ctas)
		const bool hint = (h->x_hint == 0);
		// shape 0: short slices are latency-bound (more resident warps win), wide ones want registers for loads in flight
		// (measured, profiles/r01f_sweep.txt: 8 x 48 wins below ~32 columns per slice, 8 x 32 above)
		int shape = h->sell_variant;
		if (shape == 0)
			shape = (h->m > 0 && (double) h->padded / (double) h->m < 32.0) ? 2 : 1;
		switch (shape) {
		case 2:  if (hint) SELL_LAUNCH(true, 8, 12); else SELL_LAUNCH(false, 8, 12); break;    // 48 warps / SM, 8 steps in flight
		case 3:  if (hint) SELL_LAUNCH(true, 4, 16); else SELL_LAUNCH(false, 4, 16); break;    // 64 warps / SM, 4 steps in flight
		case 4:  if (hint) SELL_LAUNCH(true, 16, 4); else SELL_LAUNCH(false, 16, 4); break;    // 16 warps / SM, 16 steps in flight
		default: if (hint) SELL_LAUNCH(true, 8, 8); else SELL_LAUNCH(false, 8, 8); break;      // 32 warps / SM, 8 steps in flight
		}
		#undef SELL_LAUNCH
		CU_LAUNCH();
	}
	if (l && l->norm2)
		return launch_norm_reduce(h, ctas + ncuts, l->norm2, st);
	return 0;
}

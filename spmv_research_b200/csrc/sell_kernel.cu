// sell_kernel.cu -- fp64 SELL-C-sigma SpMV for sm_100a.
//
// Semantics of the reference kernel (K/sell-C-s/RISC-V/sellcs_mv_kernels_epi.c:178-257): per
// slice, acc[lane] += val[k*C+lane] * x[col[k*C+lane]] for k < slice_width, then
// y[row_order[slice*C+lane]] = acc[lane]; the ragged last slice stores min(C, m - s*C) lanes.
//
// B200 mapping: one thread owns TWO adjacent lanes (sorted rows 2q, 2q+1) of a slice, so each
// step is one 128-bit values load + one 64-bit col_ind load, contiguous across the threads of a
// slice (column-major slices are coalesced by construction); four steps are in flight per thread
// before the first x gather; no reduction is needed because a lane is a row.  Padding entries
// are (col 0, val 0.0) exactly as the reference writes them and are multiplied like real ones.
#include "common.cuh"

template <bool HINT>
__global__ void __launch_bounds__(SELL_THREADS, 8)
sell_kernel(const int32_t * __restrict__ row_order, const int32_t * __restrict__ slice_width,
            const int64_t * __restrict__ slice_ptr, const int32_t * __restrict__ scol, const double * __restrict__ sval,
            const double * __restrict__ x, int64_t m, int C, int c_shift, YOut out)
{
	__shared__ double s_red[SELL_THREADS / 32];
	const uint64_t pol = make_policy_evict_last();
	const uint64_t pol_stream = make_policy_evict_first();
	const int64_t q = (int64_t) blockIdx.x * SELL_THREADS + threadIdx.x;
	const int64_t r0 = 2 * q;
	double nrm = 0.0;
	if (r0 < m) {
		const int64_t s = r0 >> c_shift;
		const int w = slice_width[s];
		int64_t p = slice_ptr[s] + (r0 & (C - 1));
		double acc0 = 0.0, acc1 = 0.0;
		int k = 0;
		for (; k + 4 <= w; k += 4) {
			double2 v[4];
			int2 c[4];
			#pragma unroll
			for (int u = 0; u < 4; u++) {
				v[u] = ld_stream_f64x2(sval + p + (int64_t) u * C, pol_stream);
				c[u] = ld_stream_i32x2(scol + p + (int64_t) u * C, pol_stream);
			}
			#pragma unroll
			for (int u = 0; u < 4; u++) {
				acc0 = fma(v[u].x, ld_x<HINT>(x + c[u].x, pol), acc0);
				acc1 = fma(v[u].y, ld_x<HINT>(x + c[u].y, pol), acc1);
			}
			p += 4 * (int64_t) C;
		}
		for (; k < w; k++) {
			const double2 v = ld_stream_f64x2(sval + p, pol_stream);
			const int2 c = ld_stream_i32x2(scol + p, pol_stream);
			acc0 = fma(v.x, ld_x<HINT>(x + c.x, pol), acc0);
			acc1 = fma(v.y, ld_x<HINT>(x + c.y, pol), acc1);
			p += C;
		}
		const double scale = out.scale ? *out.scale : 1.0;
		acc0 *= scale;
		acc1 *= scale;
		y_store(out, row_order[r0], acc0);
		nrm = acc0 * acc0;
		if (r0 + 1 < m) {
			y_store(out, row_order[r0 + 1], acc1);
			nrm += acc1 * acc1;
		}
	}
	if (out.norm_partial) {
		nrm = warp_sum(nrm);
		if ((threadIdx.x & 31) == 0)
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
		out.scale = l->scale;
		out.norm_partial = l->norm2 ? h->d_norm_partial : nullptr;
	}
	const int64_t ctas = (h->m + SELL_ROWS_PER_CTA - 1) / SELL_ROWS_PER_CTA;
	if (ctas == 0)
		return 0;
	int shift = 0;
	while ((1 << shift) < h->C)
		shift++;
	if (h->x_hint == 0)
		sell_kernel<true><<<(unsigned) ctas, SELL_THREADS, 0, st>>>(h->d_row_order, h->d_slice_width, h->d_slice_ptr, h->d_scol,
		                                                           h->d_sval, dx, h->m, h->C, shift, out);
	else
		sell_kernel<false><<<(unsigned) ctas, SELL_THREADS, 0, st>>>(h->d_row_order, h->d_slice_width, h->d_slice_ptr, h->d_scol,
		                                                            h->d_sval, dx, h->m, h->C, shift, out);
	CU_LAUNCH();
	if (l && l->norm2)
		return launch_norm_reduce(h, ctas, l->norm2, st);
	return 0;
}

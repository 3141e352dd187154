// amg_device.cu -- pass 2 of the synthetic matrix source on the GPU.
//
// b200spmv_amg_fill_device() is the device counterpart of b200amg_fill() (csrc/amg.c), i.e. of the second half of the
// reference's artificial_matrix_generation() (artificial-matrix-generator/artificial_matrix_generation.c:366-523): given
// row_ptr (pass 1, host: libm-dependent, m draws) it writes col_ind / values of the rows [i_s, i_e) straight into HBM,
// bit-identical to the reference.  The generator logic is amg_fill.h (one warp per anchor segment, see there); this file
// plans the launch: which segments the row range touches, the duplicate-set storage each one gets, scratch for the (at
// most two) segments that stick out of the range.
#include <vector>
#include "common.cuh"
#include "../../include/b200amg.h"
#include "amg_fill.h"
#include "amg_fill_plan.h"

// longest row of each segment [seg_first + blockIdx.x]
__global__ void __launch_bounds__(256) amgd_seg_maxdeg_kernel(const int32_t * row_ptr, int64_t m, int64_t period, int64_t seg_first,
                                                               int32_t * maxdeg)
{
	__shared__ int red[8];
	const int64_t i0 = (seg_first + blockIdx.x) * period;
	const int64_t i1 = i0 + period > m ? m : i0 + period;
	int mx = 0;
	for (int64_t i = i0 + threadIdx.x; i < i1; i += blockDim.x)
		mx = max(mx, row_ptr[i + 1] - row_ptr[i]);
	for (int o = 16; o > 0; o >>= 1)
		mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
	if ((threadIdx.x & 31) == 0)
		red[threadIdx.x >> 5] = mx;
	__syncthreads();
	if (threadIdx.x == 0) {
		for (int w = 1; w < 8; w++)
			mx = max(mx, red[w]);
		maxdeg[blockIdx.x] = mx;
	}
}

// One warp (= one CTA) per anchor segment seg_lo + blockIdx.x.  Up to 32 CTAs are resident per SM: the <= 4096 segments of
// a matrix are all in flight at once, each a single dependent instruction stream (lane 0) that the SM interleaves with
// the others; the row sort and the copy out of the staging buffer use all 32 lanes.
__global__ void __launch_bounds__(32) amgd_fill_kernel(amgd_model M, const int32_t * row_ptr, int64_t m, int64_t seg_lo, int64_t seg_first,
                                                        int64_t base, int32_t * col, double * val, uint64_t * slots,
                                                        const int64_t * slot_off, unsigned long long * err)
{
	__shared__ int32_t rng_words[AMGD_DEG + 1];
	__shared__ int32_t stage[AMGD_STAGE];
	const int64_t seg = seg_lo + blockIdx.x;
	const int64_t i0 = seg * M.period;
	const int64_t i1 = i0 + M.period > m ? m : i0 + M.period;
	const int64_t off = slot_off[seg - seg_first];
	const uint64_t mask = (uint64_t) (slot_off[seg - seg_first + 1] - off) - 1;
	const int64_t rc = amgd_segment(&M, row_ptr, i0, i1, base, col, val, rng_words, stage, slots + off, mask);
	if (rc && (threadIdx.x & 31) == 0)
		atomicMax(err, (unsigned long long) rc);
}

namespace {
struct scratch {
	void * p[8] = {};
	int n = 0;
	~scratch()
	{
		for (int i = 0; i < n; i++)
			cudaFree(p[i]);
	}
	template <class T> cudaError_t get(T ** out, size_t bytes)
	{
		void * q = nullptr;
		cudaError_t e = cudaMalloc(&q, bytes ? bytes : 1);
		if (e == cudaSuccess)
			p[n++] = q;
		*out = (T *) q;
		return e;
	}
};
}        // namespace

extern "C" int b200spmv_amg_fill_device(const b200amg_params * p, const int32_t * d_row_ptr, double bw_corrected, int64_t i_s, int64_t i_e,
                                        int32_t * d_col_ind, double * d_values, void * stream)
{
	B200Range nvtx_range_("b200spmv_amg_fill_device");
	if (!p || !d_row_ptr)
		B200_FAIL("amg_fill_device: null argument");
	const int64_t m = p->nr_rows;
	if (amgd_period(m) == 0 || p->nr_cols < 1)
		B200_FAIL("generator needs nr_rows >= 100 (the reference divides by nr_rows / 100) and nr_cols >= 1");
	if (p->placement != 'r' && p->placement != 'd' && p->placement != 's')
		B200_FAIL("placement must be 'r' (random), 'd' (diagonal) or 's' (simple)");
	if (i_s < 0 || i_e > m || i_s > i_e)
		B200_FAIL("bad row range");
	if (i_s == i_e)
		return 0;
	if (!d_col_ind || !d_values)
		B200_FAIL("amg_fill_device: null output");
	{
		int ndev = 0, cur = 0, major = 0;
		cudaError_t e = cudaGetDeviceCount(&ndev);
		if (e != cudaSuccess || ndev < 1)
			B200_FAIL("no CUDA device: libb200spmv has no CPU fallback (%s)", cudaGetErrorString(e));
		CU(cudaGetDevice(&cur));
		CU(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, cur));
		if (major != 10)
			B200_FAIL("device %d has compute capability %d.x; this library carries sm_100a code only", cur, major);
	}
	cudaStream_t st = (cudaStream_t) stream;
	scratch S;

	amgd_plan P;
	amgd_plan_segments(&P, m, i_s, i_e);
	const int64_t nseg = P.seg_last - P.seg_first + 1;
	int32_t * d_maxdeg;
	CU(S.get(&d_maxdeg, nseg * sizeof(int32_t)));
	amgd_seg_maxdeg_kernel<<<(unsigned) nseg, 256, 0, st>>>(d_row_ptr, m, P.period, P.seg_first, d_maxdeg);
	CU_LAUNCH();
	std::vector<int32_t> maxdeg(nseg);
	CU(cudaMemcpyAsync(maxdeg.data(), d_maxdeg, nseg * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
	// the handful of row_ptr entries the host needs: the range's ends and the ends of the segments sticking out of it
	const int64_t first_i0 = P.seg_first * P.period;
	const int64_t last_i0 = P.seg_last * P.period, last_i1 = last_i0 + P.period > m ? m : last_i0 + P.period;
	const int64_t want[6] = {i_s, i_e, first_i0, first_i0 + P.period > m ? m : first_i0 + P.period, last_i0, last_i1};
	int32_t rp[6];
	for (int q = 0; q < 6; q++)
		CU(cudaMemcpyAsync(&rp[q], d_row_ptr + want[q], sizeof(int32_t), cudaMemcpyDeviceToHost, st));
	CU(cudaStreamSynchronize(st));
	amgd_make_plan(&P, m, i_s, i_e, maxdeg.data());

	uint64_t * d_slots;
	int64_t * d_slot_off;
	unsigned long long * d_err;
	CU(S.get(&d_slots, (size_t) P.slot_off[nseg] * sizeof(uint64_t)));
	CU(S.get(&d_slot_off, (nseg + 1) * sizeof(int64_t)));
	CU(S.get(&d_err, sizeof(unsigned long long)));
	CU(cudaMemsetAsync(d_slots, 0, (size_t) P.slot_off[nseg] * sizeof(uint64_t), st));
	CU(cudaMemsetAsync(d_err, 0, sizeof(unsigned long long), st));
	CU(cudaMemcpyAsync(d_slot_off, P.slot_off.data(), (nseg + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, st));

	amgd_model M;
	M.nr_cols = p->nr_cols;
	M.period = P.period;
	M.bw_corrected = bw_corrected;
	M.p_right = p->avg_num_neighbours / 2;
	M.crs = p->cross_row_similarity;
	M.seed = p->seed;
	M.placement = p->placement;

	const int64_t base = rp[0];
	if (P.full_first <= P.full_last) {
		amgd_fill_kernel<<<(unsigned) (P.full_last - P.full_first + 1), 32, 0, st>>>(M, d_row_ptr, m, P.full_first, P.seg_first, base, d_col_ind,
		                                                                              d_values, d_slots, d_slot_off, d_err);
		CU_LAUNCH();
	}
	// segments that stick out of [i_s, i_e): generated whole into scratch (a row depends on the stream since its segment
	// start), the overlap copied
	for (int side = 0; side < 2; side++) {
		const int64_t s = side == 0 ? P.seg_first : P.seg_last;
		if (s >= P.full_first && s <= P.full_last)
			continue;
		if (side == 1 && P.seg_last == P.seg_first)
			continue;
		const int64_t i0 = s * P.period, i1 = i0 + P.period > m ? m : i0 + P.period;
		const int64_t sb = side == 0 ? rp[2] : rp[4], sn = (side == 0 ? rp[3] : rp[5]) - sb;
		const int64_t a = i0 > i_s ? i0 : i_s, b = i1 < i_e ? i1 : i_e;
		// row_ptr[a], row_ptr[b]: a is i_s or i0, b is i_e or i1 -- all among the entries fetched above
		const int64_t rp_a = (a == i_s) ? rp[0] : sb, rp_b = (b == i_e) ? rp[1] : sb + sn;
		int32_t * tc;
		double * tv;
		CU(S.get(&tc, (size_t) sn * sizeof(int32_t)));
		CU(S.get(&tv, (size_t) sn * sizeof(double)));
		amgd_fill_kernel<<<1, 32, 0, st>>>(M, d_row_ptr, m, s, P.seg_first, sb, tc, tv, d_slots, d_slot_off, d_err);
		CU_LAUNCH();
		CU(cudaMemcpyAsync(d_col_ind + (rp_a - base), tc + (rp_a - sb), (size_t) (rp_b - rp_a) * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
		CU(cudaMemcpyAsync(d_values + (rp_a - base), tv + (rp_a - sb), (size_t) (rp_b - rp_a) * sizeof(double), cudaMemcpyDeviceToDevice, st));
	}
	unsigned long long err = 0;
	CU(cudaMemcpyAsync(&err, d_err, sizeof(err), cudaMemcpyDeviceToHost, st));
	CU(cudaStreamSynchronize(st));
	if (err)
		B200_FAIL("row %lld: similarity copy hit a duplicate column (the reference aborts here too)", (long long) err - 1);
	return 0;
}

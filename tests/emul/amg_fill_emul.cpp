/*
 * amg_fill_emul.cpp -- TEST INFRASTRUCTURE, not product code.
 *
 * Compiles the device generator's source (spmv_research_b200/csrc/amg_fill.h + amg_fill_plan.h) as plain C++ with a
 * one-lane "warp", so that tests/test_amg_device_emul.py can compare its output bit for bit with the host generator
 * (csrc/amg.c) and the reference generator without a GPU.  Mirrors the driver of csrc/amg_device.cu: same planning, same
 * scratch handling for segments that stick out of the requested row range.  Nothing in the product loads this library.
 */
#define AMGD_HOST_EMULATION 1
#include <stdlib.h>
#include <string.h>
#include "../../spmv_research_b200/csrc/amg_fill.h"
#include "../../spmv_research_b200/csrc/amg_fill_plan.h"

extern "C" int64_t amgd_emul_fill(int64_t nr_rows, int64_t nr_cols, char placement, double neigh, double crs, uint32_t seed,
                                  const int32_t * row_ptr, double bw_corrected, int64_t i_s, int64_t i_e, int32_t * col, double * val)
{
	if (amgd_period(nr_rows) == 0 || i_s < 0 || i_e > nr_rows || i_s > i_e)
		return -1;
	if (i_s == i_e)
		return 0;
	amgd_plan P;
	amgd_plan_segments(&P, nr_rows, i_s, i_e);
	const int64_t nseg = P.seg_last - P.seg_first + 1;
	std::vector<int32_t> maxdeg(nseg, 0);
	for (int64_t s = 0; s < nseg; s++) {
		int64_t i0 = (P.seg_first + s) * P.period, i1 = i0 + P.period > nr_rows ? nr_rows : i0 + P.period;
		for (int64_t i = i0; i < i1; i++)
			if (row_ptr[i + 1] - row_ptr[i] > maxdeg[s])
				maxdeg[s] = row_ptr[i + 1] - row_ptr[i];
	}
	amgd_make_plan(&P, nr_rows, i_s, i_e, maxdeg.data());
	std::vector<uint64_t> slots(P.slot_off[nseg], 0);
	amgd_model M;
	M.nr_cols = nr_cols;
	M.period = P.period;
	M.bw_corrected = bw_corrected;
	M.p_right = neigh / 2;
	M.crs = crs;
	M.seed = seed;
	M.placement = placement;
	const int64_t base = row_ptr[i_s];
	int32_t rng_words[AMGD_DEG + 1], stage[AMGD_STAGE];
	for (int64_t s = P.seg_first; s <= P.seg_last; s++) {
		int64_t i0 = s * P.period, i1 = i0 + P.period > nr_rows ? nr_rows : i0 + P.period;
		uint64_t * sl = slots.data() + P.slot_off[s - P.seg_first];
		uint64_t mask = (uint64_t) (P.slot_off[s - P.seg_first + 1] - P.slot_off[s - P.seg_first]) - 1;
		int64_t rc;
		if (s >= P.full_first && s <= P.full_last) {
			rc = amgd_segment(&M, row_ptr, i0, i1, base, col, val, rng_words, stage, sl, mask);
		} else {
			int64_t sb = row_ptr[i0], sn = row_ptr[i1] - sb;
			std::vector<int32_t> tc(sn > 0 ? sn : 1);
			std::vector<double> tv(sn > 0 ? sn : 1);
			rc = amgd_segment(&M, row_ptr, i0, i1, sb, tc.data(), tv.data(), rng_words, stage, sl, mask);
			int64_t a = i0 > i_s ? i0 : i_s, b = i1 < i_e ? i1 : i_e;
			int64_t off = row_ptr[a] - sb, cnt = row_ptr[b] - row_ptr[a];
			memcpy(col + (row_ptr[a] - base), tc.data() + off, cnt * sizeof(int32_t));
			memcpy(val + (row_ptr[a] - base), tv.data() + off, cnt * sizeof(double));
		}
		if (rc)
			return rc;
	}
	return 0;
}

/* the warp sort alone, for lengths around the power-of-two and staging boundaries */
extern "C" void amgd_emul_sort(int32_t * a, int64_t n) { amgd_sort(a, n); }

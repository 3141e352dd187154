/*
 * amg_fill_plan.h -- host-side planning shared by the device pass 2 (amg_device.cu) and its CPU emulation
 * (tests/emul/amg_fill_emul.cpp): the anchor period (artificial_matrix_generation.c:259-262), the segments a row range
 * touches, and the duplicate-set storage each segment gets (sized by ITS longest row, so a matrix with one giant skew row
 * does not multiply that row's table by 4096 segments).
 */
#pragma once
#include <stdint.h>
#include <vector>

/* rows per anchor segment; 0 when the reference itself cannot run (it divides by nr_rows / 100) */
static inline int64_t amgd_period(int64_t nr_rows)
{
	if (nr_rows < 100)
		return 0;
	int64_t anchors = nr_rows / 100;
	if (anchors > (1 << 12))
		anchors = 1 << 12;
	return nr_rows / anchors;
}

/* power of two >= 2 * max_degree + 2 (at most half full: probing always ends), at least 64 */
static inline uint64_t amgd_set_capacity(int64_t max_degree)
{
	uint64_t cap = 64;
	while (cap < (uint64_t) max_degree * 2 + 2)
		cap <<= 1;
	return cap;
}

struct amgd_plan {
	int64_t period, seg_first, seg_last;        /* segments [seg_first, seg_last] overlap rows [i_s, i_e) */
	int64_t full_first, full_last;              /* the ones lying wholly inside (full_first > full_last: none) */
	std::vector<int64_t> slot_off;              /* seg_last - seg_first + 2 offsets into the set storage */
};

/* max_degree[s - seg_first] = longest row of segment s */
static inline void amgd_make_plan(amgd_plan * P, int64_t m, int64_t i_s, int64_t i_e, const int32_t * max_degree)
{
	const int64_t nseg = P->seg_last - P->seg_first + 1;
	P->slot_off.assign(nseg + 1, 0);
	for (int64_t s = 0; s < nseg; s++)
		P->slot_off[s + 1] = P->slot_off[s] + (int64_t) amgd_set_capacity(max_degree[s]);
	P->full_first = (i_s + P->period - 1) / P->period;
	P->full_last = P->seg_last;
	int64_t end_last = (P->seg_last + 1) * P->period;
	if (end_last > m)
		end_last = m;
	if (end_last > i_e)
		P->full_last = P->seg_last - 1;
}

static inline void amgd_plan_segments(amgd_plan * P, int64_t m, int64_t i_s, int64_t i_e)
{
	P->period = amgd_period(m);
	P->seg_first = i_s / P->period;
	P->seg_last = (i_e - 1) / P->period;
}

// TEST INFRASTRUCTURE: csrc/csr_perm.h (the CSR_EXP_PERM experiment's row -> thread table) compiled as plain C++ so that
// tests/test_perm_emul.py can check it without a GPU.
#include "../../spmv_research_b200/csrc/csr_perm.h"
#include <string.h>

extern "C" int64_t perm_emul_build(const int32_t * row_ptr, const int32_t * tile_info, int64_t ntiles, int T, int ext, int64_t * off_out,
                                   uint16_t * perm_out, int64_t perm_cap, int64_t * stats4)
{
	std::vector<int64_t> off;
	std::vector<uint16_t> perm;
	csr_perm_build(row_ptr, tile_info, ntiles, T, ext, off, perm, stats4);
	memcpy(off_out, off.data(), off.size() * sizeof(int64_t));
	if ((int64_t) perm.size() <= perm_cap && !perm.empty())
		memcpy(perm_out, perm.data(), perm.size() * sizeof(uint16_t));
	return (int64_t) perm.size();
}

extern "C" void rot_emul_build(const int32_t * row_ptr, const int32_t * tile_info, int64_t ntiles, int64_t m, int T, int ext, uint8_t * rot_out,
                               int64_t * stats2)
{
	std::vector<uint8_t> rot;
	csr_rot_build(row_ptr, tile_info, ntiles, m, T, ext, rot, stats2);
	memcpy(rot_out, rot.data(), (size_t) (m > 0 ? m : 1));
}

// kexp.cu -- kernel experiments on generator matrices (one B200): floors of the CSR SpMV data path.
//   stream      : values + col_ind only                      (12 B/nnz HBM stream floor in this code shape)
//   gather_ldg  : stream + x gather through LSU loads, no row reduction
//   gather_tex  : stream + x gather through the texture pipe (tex1Dfetch<int2>)
//   colgather_* : col_ind + gather only (no values)
// Each reports ms and GB/s over the matrix's compulsory bytes so numbers compare with bench.py.
// Build: see tools/build_kexp.sh.   Run: tools/kexp "<param line>" ["<param line>" ...]
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>
#include "../include/b200amg.h"
#include "../include/b200spmv.h"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

constexpr int TH = 256;

__device__ __forceinline__ uint64_t pol_first() { uint64_t p; asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ uint64_t pol_last() { uint64_t p; asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ double lds_f64(const double * p, uint64_t pol) { double r; asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(r) : "l"(p), "l"(pol)); return r; }
__device__ __forceinline__ int lds_i32(const int * p, uint64_t pol) { int r; asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.s32 %0, [%1], %2;" : "=r"(r) : "l"(p), "l"(pol)); return r; }
__device__ __forceinline__ double ldx(const double * p, uint64_t pol) { double r; asm volatile("ld.global.nc.L1::evict_last.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(r) : "l"(p), "l"(pol)); return r; }

// MODE: 0 stream only, 1 LDG gather, 2 TEX gather;  VAL: load values or not
template <int NPT, int MODE, bool VAL>
__global__ void __launch_bounds__(TH) k_floor(const int * __restrict__ col, const double * __restrict__ val, const double * __restrict__ x,
                                              cudaTextureObject_t tx, double * __restrict__ out)
{
	const int tid = threadIdx.x;
	const int64_t base = (int64_t) blockIdx.x * (TH * NPT);
	const uint64_t pf = pol_first(), pl = pol_last();
	int c[NPT];
	double v[NPT];
	#pragma unroll
	for (int k = 0; k < NPT; k++) {
		const int64_t idx = base + k * TH + tid;
		c[k] = lds_i32(col + idx, pf);
		v[k] = VAL ? lds_f64(val + idx, pf) : 1.0;
	}
	double acc = 0.0;
	#pragma unroll
	for (int k = 0; k < NPT; k++) {
		if (MODE == 0)
			acc += v[k] * (double) c[k];
		else if (MODE == 1)
			acc += v[k] * ldx(x + c[k], pl);
		else {
			const int2 t = tex1Dfetch<int2>(tx, c[k]);
			acc += v[k] * __hiloint2double(t.y, t.x);
		}
	}
	out[(int64_t) blockIdx.x * TH + tid] = acc;
}

// stream + LDG gather + NSH dependent 64-bit shuffles per group of 32 nnz (cost of a register segmented scan)
template <int NPT, int NSH>
__global__ void __launch_bounds__(TH) k_shfl(const int * __restrict__ col, const double * __restrict__ val, const double * __restrict__ x, double * __restrict__ out)
{
	const int tid = threadIdx.x;
	const int64_t base = (int64_t) blockIdx.x * (TH * NPT);
	const uint64_t pf = pol_first(), pl = pol_last();
	int c[NPT];
	double v[NPT];
	#pragma unroll
	for (int k = 0; k < NPT; k++) {
		const int64_t idx = base + k * TH + tid;
		c[k] = lds_i32(col + idx, pf);
		v[k] = lds_f64(val + idx, pf);
	}
	double acc = 0.0;
	#pragma unroll
	for (int k = 0; k < NPT; k++) {
		double p = v[k] * ldx(x + c[k], pl);
		#pragma unroll
		for (int s = 0; s < NSH; s++) {
			const double t = __shfl_up_sync(0xffffffffu, p, 1 << (s % 5));
			if ((tid & 31) >= (1 << (s % 5)))
				p += t;
		}
		acc += p;
	}
	out[(int64_t) blockIdx.x * TH + tid] = acc;
}

// stream + LDG gather + products through shared memory and back (thread-per-"row" of RL consecutive products)
template <int NPT, int RL>
__global__ void __launch_bounds__(TH) k_smem(const int * __restrict__ col, const double * __restrict__ val, const double * __restrict__ x, double * __restrict__ out)
{
	__shared__ double s_prod[TH * NPT];
	const int tid = threadIdx.x;
	const int64_t base = (int64_t) blockIdx.x * (TH * NPT);
	const uint64_t pf = pol_first(), pl = pol_last();
	int c[NPT];
	double v[NPT];
	#pragma unroll
	for (int k = 0; k < NPT; k++) {
		const int64_t idx = base + k * TH + tid;
		c[k] = lds_i32(col + idx, pf);
		v[k] = lds_f64(val + idx, pf);
	}
	#pragma unroll
	for (int k = 0; k < NPT; k++)
		s_prod[k * TH + tid] = v[k] * ldx(x + c[k], pl);
	__syncthreads();
	constexpr int ROWS = TH * NPT / RL;
	for (int r = tid; r < ROWS; r += TH) {
		double acc = 0.0;
		#pragma unroll
		for (int j = 0; j < RL; j++)
			acc += s_prod[r * RL + j];
		out[(int64_t) blockIdx.x * ROWS + r] = acc;
	}
}

// ---- persistent CTA, bulk-async (TMA 1D) ring of STAGES tiles, one producer warp + 8 consumer warps ----
__device__ __forceinline__ uint32_t s32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t * b, unsigned n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(s32(b)), "r"(n)); }
__device__ __forceinline__ void mbar_expect(uint64_t * b, unsigned bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(s32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t * b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(s32(b)) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t * b, unsigned phase)
{
	unsigned ok;
	do {
		asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(s32(b)), "r"(phase) : "memory");
	} while (!ok);
}
__device__ __forceinline__ void bulk_g2s(void * dst, const void * src, unsigned bytes, uint64_t * bar, uint64_t pol)
{
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
	             :: "r"(s32(dst)), "l"(src), "r"(bytes), "r"(s32(bar)), "l"(pol) : "memory");
}

template <int NPT, int STAGES, int MODE>       // MODE 0: stream only, 1: + gather
__global__ void __launch_bounds__(TH + 32) k_ring(const int * __restrict__ col, const double * __restrict__ val, const double * __restrict__ x,
                                                  int64_t ntiles, double * __restrict__ out)
{
	constexpr int T = TH * NPT;
	extern __shared__ __align__(128) unsigned char smem[];
	double * s_val = reinterpret_cast<double *>(smem);                           // STAGES * T
	int * s_col = reinterpret_cast<int *>(smem + (size_t) STAGES * T * 8);       // STAGES * T
	__shared__ uint64_t full[STAGES], empty[STAGES];
	const int tid = threadIdx.x;
	if (tid == 0) {
		for (int s = 0; s < STAGES; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], TH / 32); }
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncthreads();
	if (tid >= TH) {
		if (tid == TH) {
			const uint64_t pf = pol_first();
			int s = 0; unsigned ph = 0;
			for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
				mbar_wait(&empty[s], ph ^ 1);
				mbar_expect(&full[s], T * 12);
				bulk_g2s(s_val + (size_t) s * T, val + t * T, T * 8, &full[s], pf);
				bulk_g2s(s_col + (size_t) s * T, col + t * T, T * 4, &full[s], pf);
				if (++s == STAGES) { s = 0; ph ^= 1; }
			}
		}
		return;
	}
	const uint64_t pl = pol_last();
	double acc = 0.0;
	int s = 0; unsigned ph = 0;
	for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
		mbar_wait(&full[s], ph);
		const double * sv = s_val + (size_t) s * T;
		const int * sc = s_col + (size_t) s * T;
		int c[NPT];
		double v[NPT];
		#pragma unroll
		for (int k = 0; k < NPT; k++) { c[k] = sc[k * TH + tid]; v[k] = sv[k * TH + tid]; }
		#pragma unroll
		for (int k = 0; k < NPT; k++)
			acc += (MODE == 0) ? v[k] * (double) c[k] : v[k] * ldx(x + c[k], pl);
		__syncwarp();
		if ((tid & 31) == 0)
			mbar_arrive(&empty[s]);
		if (++s == STAGES) { s = 0; ph ^= 1; }
	}
	out[(int64_t) blockIdx.x * TH + tid] = acc;
}

// simple row-per-thread reference (serial ascending-j sum per row) and per-row bound sum |a x|
__global__ void k_ref(const int * __restrict__ rp, const int * __restrict__ col, const double * __restrict__ val, const double * __restrict__ x,
                      int64_t m, double * __restrict__ y, double * __restrict__ bound)
{
	const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= m) return;
	double s = 0, b = 0;
	for (int j = rp[r]; j < rp[r + 1]; j++) { const double p = val[j] * x[col[j]]; s += p; b += fabs(p); }
	y[r] = s; bound[r] = b;
}
__global__ void k_cmp(const double * __restrict__ y, const double * __restrict__ yref, const double * __restrict__ bound, int64_t m, unsigned long long * nbad, double * worst)
{
	const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= m) return;
	const double e = fabs(y[r] - yref[r]);
	if (!(e <= 1e-12 * bound[r])) atomicAdd(nbad, 1ULL);
	const double q = e / fmax(bound[r], 1e-300);
	if (q > 1e-13) atomicMax((unsigned long long *) worst, (unsigned long long) __double_as_longlong(q));
}

struct Mat { int64_t m, n, nnz; std::vector<int> rp, col; std::vector<double> val; };

static bool parse_line(const char * s, b200amg_params * p)
{
	char dist[32], plc[32];
	long r, c;
	unsigned seed;
	memset(p, 0, sizeof(*p));
	if (sscanf(s, "%ld %ld %lf %lf %31s %31s %lf %lf %lf %lf %u", &r, &c, &p->avg_nnz_per_row, &p->std_nnz_per_row, dist, plc, &p->bw_scaled, &p->skew,
	           &p->avg_num_neighbours, &p->cross_row_similarity, &seed) != 11)
		return false;
	p->nr_rows = r; p->nr_cols = c; p->distribution = dist[0]; p->placement = plc[0]; p->seed = seed;
	return true;
}

template <typename F> static float time_ms(F f, int iters = 10)
{
	cudaEvent_t a, b;
	cudaEventCreate(&a); cudaEventCreate(&b);
	for (int i = 0; i < 3; i++) f();
	std::vector<float> t(iters);
	for (int i = 0; i < iters; i++) {
		cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
		cudaEventElapsedTime(&t[i], a, b);
	}
	CK(cudaGetLastError());
	std::sort(t.begin(), t.end());
	return t[iters / 2];
}

int main(int argc, char ** argv)
{
	for (int a = 1; a < argc; a++) {
		b200amg_params p;
		if (!parse_line(argv[a], &p)) { printf("bad line %s\n", argv[a]); return 1; }
		Mat M;
		M.m = p.nr_rows; M.n = p.nr_cols;
		M.rp.resize(M.m + 1);
		double bw;
		M.nnz = b200amg_row_ptr(&p, M.rp.data(), &bw);
		if (M.nnz < 0) { printf("gen failed: %s\n", b200amg_last_error()); return 1; }
		M.col.resize(M.nnz); M.val.resize(M.nnz);
		if (b200amg_fill(&p, M.rp.data(), bw, 0, M.m, M.col.data(), M.val.data())) { printf("fill failed\n"); return 1; }
		// distinct 128-byte x lines / 32-byte sectors per group of 32 consecutive nnz (a warp-wide gather instruction)
		double lines = 0, sectors = 0;
		int64_t groups = 0;
		for (int64_t g = 0; g + 32 <= M.nnz; g += 32 * 7) {    // sample every 7th group
			int l[32], s[32];
			for (int i = 0; i < 32; i++) { l[i] = M.col[g + i] >> 4; s[i] = M.col[g + i] >> 2; }
			std::sort(l, l + 32); std::sort(s, s + 32);
			lines += std::unique(l, l + 32) - l; sectors += std::unique(s, s + 32) - s;
			groups++;
		}
		const double bc = 12.0 * M.nnz + 12.0 * M.m + 8.0 * M.n + 4;
		printf("== %s\n   nnz %ld  B_c %.1f MB  distinct x lines per 32 nnz %.2f, sectors %.2f\n", argv[a], (long) M.nnz, bc * 1e-6, lines / groups, sectors / groups);
		constexpr int NPT = 8;
		const int64_t T = TH * NPT, ntiles = (M.nnz + T - 1) / T, padded = ntiles * T;
		int * d_col; double * d_val, * d_x, * d_out;
		CK(cudaMalloc(&d_col, padded * 4)); CK(cudaMalloc(&d_val, padded * 8)); CK(cudaMalloc(&d_x, M.n * 8)); CK(cudaMalloc(&d_out, (ntiles + 1024) * TH * 8));
		CK(cudaMemset(d_col, 0, padded * 4)); CK(cudaMemset(d_val, 0, padded * 8));
		CK(cudaMemcpy(d_col, M.col.data(), M.nnz * 4, cudaMemcpyHostToDevice));
		CK(cudaMemcpy(d_val, M.val.data(), M.nnz * 8, cudaMemcpyHostToDevice));
		std::vector<double> hx(M.n, 1.0);
		CK(cudaMemcpy(d_x, hx.data(), M.n * 8, cudaMemcpyHostToDevice));
		cudaResourceDesc rd; memset(&rd, 0, sizeof(rd));
		rd.resType = cudaResourceTypeLinear; rd.res.linear.devPtr = d_x; rd.res.linear.desc = cudaCreateChannelDesc<int2>(); rd.res.linear.sizeInBytes = M.n * 8;
		cudaTextureDesc td; memset(&td, 0, sizeof(td)); td.readMode = cudaReadModeElementType;
		cudaTextureObject_t tx = 0;
		CK(cudaCreateTextureObject(&tx, &rd, &td, nullptr));
		const unsigned G = (unsigned) ntiles;
		std::vector<std::pair<const char *, float>> r;
		if (getenv("KEXP_FLOORS")) {
		r.push_back({ "stream (col+val)", time_ms([&] { k_floor<NPT, 0, true><<<G, TH>>>(d_col, d_val, d_x, tx, d_out); }) });
		r.push_back({ "stream + LDG gather", time_ms([&] { k_floor<NPT, 1, true><<<G, TH>>>(d_col, d_val, d_x, tx, d_out); }) });
		r.push_back({ "  + 5 shfl64/group", time_ms([&] { k_shfl<NPT, 5><<<G, TH>>>(d_col, d_val, d_x, d_out); }) });
		r.push_back({ "  + 10 shfl64/group", time_ms([&] { k_shfl<NPT, 10><<<G, TH>>>(d_col, d_val, d_x, d_out); }) });
		r.push_back({ "  + smem rt, rows of 8", time_ms([&] { k_smem<NPT, 8><<<G, TH>>>(d_col, d_val, d_x, d_out); }) });
		r.push_back({ "  + smem rt, rows of 9", time_ms([&] { k_smem<NPT, 9><<<G, TH>>>(d_col, d_val, d_x, d_out); }) });
		r.push_back({ "  + smem rt, rows of 16", time_ms([&] { k_smem<NPT, 16><<<G, TH>>>(d_col, d_val, d_x, d_out); }) });
		{
			auto ring = [&](auto kern, int stages, int ctas_per_sm, const char * name) {
				const size_t sm = (size_t) stages * T * 12;
				CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sm));
				const unsigned g = 148 * ctas_per_sm;
				r.push_back({ name, time_ms([&] { kern<<<g, TH + 32, sm>>>(d_col, d_val, d_x, ntiles, d_out); }) });
			};
			ring(k_ring<NPT, 4, 0>, 4, 2, "ring4x2 stream");
			ring(k_ring<NPT, 4, 1>, 4, 2, "ring4x2 + gather");
			ring(k_ring<NPT, 3, 0>, 3, 3, "ring3x3 stream");
			ring(k_ring<NPT, 3, 1>, 3, 3, "ring3x3 + gather");
			ring(k_ring<NPT, 2, 0>, 2, 4, "ring2x4 stream");
			ring(k_ring<NPT, 2, 1>, 2, 4, "ring2x4 + gather");
			ring(k_ring<NPT, 8, 1>, 8, 1, "ring8x1 + gather");
		}
		}
		for (auto & e : r)
			printf("   %-24s %8.4f ms  %7.0f GB/s(B_c)  %6.1f Gnnz/s\n", e.first, e.second, bc / e.second * 1e-6, M.nnz / e.second * 1e-6);
		// ---- the library's kernels through the C-ABI, checked against k_ref under the per-row bound ----
		{
			int * d_rp; double * d_y, * d_yref, * d_bound; unsigned long long * d_nbad; double * d_worst;
			CK(cudaMalloc(&d_rp, (M.m + 1) * 4)); CK(cudaMalloc(&d_y, M.m * 8)); CK(cudaMalloc(&d_yref, M.m * 8)); CK(cudaMalloc(&d_bound, M.m * 8));
			CK(cudaMalloc(&d_nbad, 8)); CK(cudaMalloc(&d_worst, 8));
			CK(cudaMemcpy(d_rp, M.rp.data(), (M.m + 1) * 4, cudaMemcpyHostToDevice));
			std::vector<double> hx2(M.n);
			uint64_t st = 88172645463325252ULL;
			for (auto & v : hx2) { st ^= st << 13; st ^= st >> 7; st ^= st << 17; v = (double) (st >> 11) / 9007199254740992.0 * 2.0 - 1.0; }
			double * d_x2; CK(cudaMalloc(&d_x2, M.n * 8)); CK(cudaMemcpy(d_x2, hx2.data(), M.n * 8, cudaMemcpyHostToDevice));
			k_ref<<<(unsigned) ((M.m + 255) / 256), 256>>>(d_rp, d_col, d_val, d_x2, M.m, d_yref, d_bound);
			CK(cudaDeviceSynchronize());
			struct V { const char * name; int fmt, npt, variant, ext; } vs[] = {
				{ "lib csr auto", B200SPMV_CSR_ADAPTIVE, 0, 0, 0 }, { "lib csr npt8 ext256", B200SPMV_CSR_ADAPTIVE, 8, 0, 256 },
				{ "lib csr npt16 ext2048", B200SPMV_CSR_ADAPTIVE, 16, 0, 2048 },
				{ "lib sell C32", B200SPMV_SELL_C_SIGMA, 0, 0, 0 } };
			for (auto & v : vs) {
				if (getenv("KEXP_VARIANTS") && !strstr(getenv("KEXP_VARIANTS"), v.name)) continue;
				b200spmv_opts o; b200spmv_default_opts(&o);
				o.format = v.fmt; o.csr_nnz_per_thread = v.npt; o.csr_ext = v.ext;
				b200spmv_t * h = nullptr;
				if (b200spmv_create_device(&h, d_rp, d_col, d_val, M.m, M.n, M.nnz, &o)) { printf("   %-24s create failed: %s\n", v.name, b200spmv_last_error()); continue; }
				CK(cudaMemset(d_y, 0xFF, M.m * 8)); CK(cudaMemset(d_nbad, 0, 8)); CK(cudaMemset(d_worst, 0, 8));
				if (b200spmv_spmv_device(h, d_x2, d_y, nullptr)) { printf("   %-24s spmv failed: %s\n", v.name, b200spmv_last_error()); b200spmv_destroy(h); continue; }
				k_cmp<<<(unsigned) ((M.m + 255) / 256), 256>>>(d_y, d_yref, d_bound, M.m, d_nbad, d_worst);
				unsigned long long nbad; double worst;
				CK(cudaMemcpy(&nbad, d_nbad, 8, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(&worst, d_worst, 8, cudaMemcpyDeviceToHost));
				std::vector<float> ms(13);
				if (b200spmv_time_device(h, d_x, d_y, 13, 0, ms.data())) { printf("   time failed: %s\n", b200spmv_last_error()); b200spmv_destroy(h); continue; }
				std::sort(ms.begin() + 3, ms.end());
				const float med = ms[8];
				printf("   %-24s %8.4f ms  %7.0f GB/s(B_c)  %6.1f Gnnz/s   bad rows %llu  max err/bound %.2e\n", v.name, med, bc / med * 1e-6, M.nnz / med * 1e-6, nbad, worst);
				b200spmv_destroy(h);
			}
			cudaFree(d_rp); cudaFree(d_y); cudaFree(d_yref); cudaFree(d_bound); cudaFree(d_nbad); cudaFree(d_worst); cudaFree(d_x2);
		}
		cudaDestroyTextureObject(tx);
		cudaFree(d_col); cudaFree(d_val); cudaFree(d_x); cudaFree(d_out);
	}
	return 0;
}

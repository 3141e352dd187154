// pcg.cu -- Jacobi-preconditioned CG on the device-resident SpMV.
//
// Follows preconditioned_cg() of the reference harness (benchmark_code/BENCH/src/bench_cg.cpp:93-320) step by step:
// x0 = 0, K = diag(A), eps = 1e-15 ||b||, explicit residual + x_best every 100 iterations with the restart rule of
// :222-236, stop on ||r|| < eps.  The reference's GPU plugins are wrong under this caller (they cache x on the first
// spmv call); here every vector stays in HBM, the products go through b200spmv_spmv_device and the vector
// work of one iteration is three fused passes:
//   pcg_dot2_kernel    zr = z.r, pAp = p.Ap            (per-CTA partials, fixed-order final sum: deterministic)
//   pcg_update_kernel  x += ak p; r -= ak Ap; z = r / K; partials of z.r and r.r
//   pcg_dir_kernel     p = z + bk p
// with the scalars (ak, bk, ||r||) kept on the device; the host reads one double per iteration for the stop test.
#include "common.cuh"

constexpr int PCG_THREADS = 256;
constexpr int PCG_CTAS = 148 * 4;

// First entry of each row with j == i (bench_cg.cpp:121-131); rows without one are counted.
__global__ void pcg_diag_kernel(const int32_t * __restrict__ row_ptr, const int32_t * __restrict__ col, const double * __restrict__ val,
                                int64_t m, double * __restrict__ K, int * __restrict__ n_bad)
{
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= m)
		return;
	double d = 0.0;
	for (int64_t j = row_ptr[i]; j < row_ptr[i + 1]; j++)
		if (col[j] == i) {
			d = val[j];
			break;
		}
	K[i] = d;
	if (d == 0.0)
		atomicAdd(n_bad, 1);
}

int pcg_extract_diag(b200spmv * h)
{
	if (h->m != h->n)
		B200_FAIL("keep_diag needs a square matrix");
	int * d_bad;
	int bad = 0;
	CU(cudaMalloc(&h->d_diag, (h->m > 0 ? h->m : 1) * sizeof(double)));
	CU(cudaMalloc(&d_bad, sizeof(int)));
	CU(cudaMemsetAsync(d_bad, 0, sizeof(int), h->stream));
	if (h->m > 0) {
		pcg_diag_kernel<<<(unsigned) ((h->m + 255) / 256), 256, 0, h->stream>>>(h->d_row_ptr, h->d_col, h->d_val, h->m, h->d_diag, d_bad);
		CU_LAUNCH();
	}
	CU(cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
	CU(cudaStreamSynchronize(h->stream));
	CU(cudaFree(d_bad));
	h->diag_bad = bad;
	return 0;
}

__device__ __forceinline__ void block_store_partials(double a, double b, double * __restrict__ pa, double * __restrict__ pb)
{
	__shared__ double sa[PCG_THREADS / 32], sb[PCG_THREADS / 32];
	a = warp_sum(a);
	b = warp_sum(b);
	if ((threadIdx.x & 31) == 0) {
		sa[threadIdx.x >> 5] = a;
		sb[threadIdx.x >> 5] = b;
	}
	__syncthreads();
	if (threadIdx.x == 0) {
		double ta = 0.0, tb = 0.0;
		#pragma unroll
		for (int w = 0; w < PCG_THREADS / 32; w++) {
			ta += sa[w];
			tb += sb[w];
		}
		pa[blockIdx.x] = ta;
		pb[blockIdx.x] = tb;
	}
}

// partial[0..G) = a1.b1, partial[G..2G) = a2.b2
__global__ void __launch_bounds__(PCG_THREADS)
pcg_dot2_kernel(const double * __restrict__ a1, const double * __restrict__ b1, const double * __restrict__ a2, const double * __restrict__ b2,
                int64_t m, double * __restrict__ partial)
{
	double s1 = 0.0, s2 = 0.0;
	for (int64_t i = (int64_t) blockIdx.x * PCG_THREADS + threadIdx.x; i < m; i += (int64_t) gridDim.x * PCG_THREADS) {
		s1 += a1[i] * b1[i];
		s2 += a2[i] * b2[i];
	}
	block_store_partials(s1, s2, partial, partial + gridDim.x);
}

// scal[] layout: 0 zr_old, 1 ak, 2 bk, 3 err (= ||r||), 4 scratch
enum { S_ZR = 0, S_AK = 1, S_BK = 2, S_ERR = 3, S_TMP = 4 };

// mode 0: zr = sum A, ak = zr / sum B                  (after pcg_dot2_kernel(z, r, p, Ap))
// mode 1: bk = sum A / zr_old, err = sqrt(sum B)       (after pcg_update_kernel)
// mode 2: err = sqrt(sum B)                            (norms only)
__global__ void pcg_final_kernel(const double * __restrict__ partial, int G, int mode, double * __restrict__ scal)
{
	__shared__ double sa[32], sb[32];
	double a = 0.0, b = 0.0;
	for (int i = threadIdx.x; i < G; i += blockDim.x) {
		a += partial[i];
		b += partial[G + i];
	}
	a = warp_sum(a);
	b = warp_sum(b);
	if ((threadIdx.x & 31) == 0) {
		sa[threadIdx.x >> 5] = a;
		sb[threadIdx.x >> 5] = b;
	}
	__syncthreads();
	if (threadIdx.x == 0) {
		double ta = 0.0, tb = 0.0;
		for (int w = 0; w < (int) (blockDim.x >> 5); w++) {
			ta += sa[w];
			tb += sb[w];
		}
		if (mode == 0) {
			scal[S_ZR] = ta;
			scal[S_AK] = ta / tb;
		} else if (mode == 1) {
			scal[S_BK] = ta / scal[S_ZR];
			scal[S_ERR] = sqrt(tb);
		} else {
			scal[S_TMP] = sqrt(ta);
			scal[S_ERR] = sqrt(tb);
		}
	}
}

// x += ak p;  r -= ak Ap;  z = r / K;  partials of z.r and r.r          (bench_cg.cpp:262-277)
__global__ void __launch_bounds__(PCG_THREADS)
pcg_update_kernel(double * __restrict__ x, double * __restrict__ r, double * __restrict__ z, const double * __restrict__ p,
                  const double * __restrict__ Ap, const double * __restrict__ K, const double * __restrict__ scal, int64_t m,
                  double * __restrict__ partial)
{
	const double ak = scal[S_AK];
	double s1 = 0.0, s2 = 0.0;
	for (int64_t i = (int64_t) blockIdx.x * PCG_THREADS + threadIdx.x; i < m; i += (int64_t) gridDim.x * PCG_THREADS) {
		x[i] = x[i] + ak * p[i];
		const double ri = r[i] + -ak * Ap[i];
		const double zi = ri / K[i];
		r[i] = ri;
		z[i] = zi;
		s1 += zi * ri;
		s2 += ri * ri;
	}
	block_store_partials(s1, s2, partial, partial + gridDim.x);
}

// p = z + bk p                                                           (bench_cg.cpp:284)
__global__ void pcg_dir_kernel(double * __restrict__ p, const double * __restrict__ z, const double * __restrict__ scal, int64_t m)
{
	const double bk = scal[S_BK];
	for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t) gridDim.x * blockDim.x)
		p[i] = z[i] + bk * p[i];
}

// out = b - Ax; when z != nullptr also z = out / K and p = z (start / restart); partials of out.out twice
__global__ void __launch_bounds__(PCG_THREADS)
pcg_residual_kernel(double * __restrict__ out, const double * __restrict__ b, const double * __restrict__ Ax, double * __restrict__ z,
                    double * __restrict__ p, const double * __restrict__ K, int64_t m, double * __restrict__ partial)
{
	double s = 0.0;
	for (int64_t i = (int64_t) blockIdx.x * PCG_THREADS + threadIdx.x; i < m; i += (int64_t) gridDim.x * PCG_THREADS) {
		const double ri = b[i] + -1.0 * Ax[i];
		out[i] = ri;
		if (z) {
			const double zi = ri / K[i];
			z[i] = zi;
			p[i] = zi;
		}
		s += ri * ri;
	}
	block_store_partials(s, s, partial, partial + gridDim.x);
}

// r = re;  z = r / K;  p = z                                              (restart, bench_cg.cpp:224-235)
__global__ void pcg_restart_kernel(double * __restrict__ r, const double * __restrict__ re, double * __restrict__ z, double * __restrict__ p,
                                   const double * __restrict__ K, int64_t m)
{
	for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t) gridDim.x * blockDim.x) {
		const double ri = re[i];
		const double zi = ri / K[i];
		r[i] = ri;
		z[i] = zi;
		p[i] = zi;
	}
}

static int read_scal(b200spmv * h, const double * d_scal, int idx, double * out)
{
	CU(cudaMemcpyAsync(out, d_scal + idx, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
	CU(cudaStreamSynchronize(h->stream));
	return 0;
}

extern "C" int b200spmv_pcg(b200spmv_t * h, const double * b_host, double * x_host, int64_t max_iterations, b200spmv_pcg_result * res)
{
	if (!h || !b_host || !x_host || !res)
		B200_FAIL("NULL argument");
	if (!h->d_diag)
		B200_FAIL("b200spmv_pcg: the handle was created without keep_diag");
	if (h->diag_bad > 0)
		B200_FAIL("bad K, zero in diagonal (%d rows)", h->diag_bad);          // bench_cg.cpp:132-133
	CU(cudaSetDevice(h->device));
	const int64_t m = h->m;
	memset(res, 0, sizeof(*res));
	if (m == 0)
		return 0;
	cudaStream_t st = h->stream;
	const int G = (int) ((m + PCG_THREADS - 1) / PCG_THREADS < PCG_CTAS ? (m + PCG_THREADS - 1) / PCG_THREADS : PCG_CTAS);
	double * mem = nullptr;
	CU(cudaMalloc(&mem, (size_t) (8 * m + 2 * PCG_CTAS + 8) * sizeof(double)));
	double * r = mem, * re = mem + m, * p = mem + 2 * m, * z = mem + 3 * m, * Ap = mem + 4 * m, * x = mem + 5 * m, * xb = mem + 6 * m,
	       * b = mem + 7 * m, * partial = mem + 8 * m, * scal = partial + 2 * PCG_CTAS;
	int rc = 1;
	double err = 0, bn = 0, err_explicit = 0, err_best = 0, eps = 0, eps_counter = 0;
	int64_t k = 0, products = 0, restarts = 0;
	const int64_t restart_k = 100;
	#define PCG_CHECK(call) do { if ((call) != cudaSuccess) goto done; } while (0)
	#define PCG_RC(call) do { if (call) goto done_msg; } while (0)
	PCG_CHECK(cudaMemcpyAsync(b, b_host, m * sizeof(double), cudaMemcpyHostToDevice, st));
	PCG_CHECK(cudaMemsetAsync(x, 0, 2 * m * sizeof(double), st));                      // x = x_best = 0 (:141-147)
	// r0 = b - A x0, z0 = r0 / K, p0 = z0, err = ||r0||, ||b||                        (:149-175)
	PCG_RC(launch_spmv_any(h, x, Ap, st));
	products++;
	pcg_residual_kernel<<<G, PCG_THREADS, 0, st>>>(r, b, Ap, z, p, h->d_diag, m, partial);
	pcg_final_kernel<<<1, 1024, 0, st>>>(partial, G, 2, scal);
	PCG_RC(read_scal(h, scal, S_ERR, &err));
	pcg_dot2_kernel<<<G, PCG_THREADS, 0, st>>>(b, b, b, b, m, partial);
	pcg_final_kernel<<<1, 1024, 0, st>>>(partial, G, 2, scal);
	PCG_RC(read_scal(h, scal, S_ERR, &bn));
	eps = 1.0e-15 * bn;
	eps_counter = 1.0e-7 * bn;
	err_explicit = err;
	err_best = err;
	while (k < max_iterations) {
		const bool check = (k > 0) && !(k % restart_k);
		if (check) {                                                                   // explicit residual (:189-211)
			PCG_RC(launch_spmv_any(h, x, Ap, st));
			products++;
			pcg_residual_kernel<<<G, PCG_THREADS, 0, st>>>(re, b, Ap, nullptr, nullptr, h->d_diag, m, partial);
			pcg_final_kernel<<<1, 1024, 0, st>>>(partial, G, 2, scal);
			PCG_RC(read_scal(h, scal, S_ERR, &err_explicit));
			if (err_explicit < err_best) {
				PCG_CHECK(cudaMemcpyAsync(xb, x, m * sizeof(double), cudaMemcpyDeviceToDevice, st));
				err_best = err_explicit;
			}
			if ((err_best > eps_counter) && (err_explicit / err > 1e3)) {              // restart (:222-236); err = ||r_k||, known
				pcg_restart_kernel<<<G, PCG_THREADS, 0, st>>>(r, re, z, p, h->d_diag, m);
				restarts++;
			}
		}
		if (err < eps)
			break;
		PCG_RC(launch_spmv_any(h, p, Ap, st));                                         // A pk (:248)
		products++;
		pcg_dot2_kernel<<<G, PCG_THREADS, 0, st>>>(z, r, p, Ap, m, partial);
		pcg_final_kernel<<<1, 1024, 0, st>>>(partial, G, 0, scal);
		pcg_update_kernel<<<G, PCG_THREADS, 0, st>>>(x, r, z, p, Ap, h->d_diag, scal, m, partial);
		pcg_final_kernel<<<1, 1024, 0, st>>>(partial, G, 1, scal);
		pcg_dir_kernel<<<G, PCG_THREADS, 0, st>>>(p, z, scal, m);
		PCG_RC(read_scal(h, scal, S_ERR, &err));                                       // ||r_{k+1}||: next iteration's stop test (:213-218, :238)
		k++;
	}
	// final explicit residual, x_best (:290-312)
	PCG_RC(launch_spmv_any(h, x, Ap, st));
	products++;
	pcg_residual_kernel<<<G, PCG_THREADS, 0, st>>>(re, b, Ap, nullptr, nullptr, h->d_diag, m, partial);
	pcg_final_kernel<<<1, 1024, 0, st>>>(partial, G, 2, scal);
	PCG_RC(read_scal(h, scal, S_ERR, &err_explicit));
	if (err_explicit < err_best) {
		PCG_CHECK(cudaMemcpyAsync(xb, x, m * sizeof(double), cudaMemcpyDeviceToDevice, st));
		err_best = err_explicit;
	}
	PCG_CHECK(cudaMemcpyAsync(x_host, xb, m * sizeof(double), cudaMemcpyDeviceToHost, st));
	PCG_CHECK(cudaStreamSynchronize(st));
	PCG_CHECK(cudaGetLastError());
	res->iterations = k;
	res->products = products;
	res->restarts = restarts;
	res->err = err;
	res->err_explicit = err_explicit;
	res->err_best = err_best;
	res->eps = eps;
	res->converged = (err < eps) ? 1 : 0;
	rc = 0;
done:
	if (rc)
		snprintf(g_b200_err, sizeof(g_b200_err), "b200spmv_pcg: CUDA failure: %s", cudaGetErrorString(cudaGetLastError()));
done_msg:
	cudaFree(mem);
	return rc;
	#undef PCG_CHECK
	#undef PCG_RC
}

// =====================================================================================================================
// Jacobi-preconditioned BiCGSTAB on the device-resident SpMV: preconditioned_bicgstab() of the reference harness
// (benchmark_code/BENCH/src/bench_bicg.cpp:150-456) step by step.  Like the reference there is NO stop test (:333-334 are
// commented out): exactly max_iterations iterations, x_best = the iterate with the smallest explicit residual at the
// checkpoints (every 100 iterations, :307-331) and at the end (:427-446).  Nothing in an iteration depends on the host, so
// the loop is launched without a single synchronisation between checkpoints; the scalars live in bscal[] on the device:
//   bi_y_kernel       y = p / K                                                        (:349-355)
//   SpMV              v = A y                                                          (:358)
//   pcg_dot2_kernel   (r0, v)          -> bi_final mode 0: s_a = s_pk_p / (r0, v)      (:368)
//   bi_hs_kernel      h = x + s_a y;  s = r - s_a v;  z = s / K                        (:375-386)
//   SpMV              t = A z                                                          (:392)
//   bi_w_kernel       (t/K, s/K), (t/K, t/K) -> bi_final mode 1: s_w                   (:399-416)
//   bi_update_kernel  r = s - s_w t;  x = h + s_w z;  (r0, r), (r, r)
//                                     -> bi_final mode 2: s_b, s_pk_p, ||r||           (:418-447)
//   bi_dir_kernel     p = r + s_b (p - s_w v)                                          (:437-440)
enum { B_PKP = 0, B_A = 1, B_W = 2, B_B = 3, B_ERR = 4 };

__global__ void bi_final_kernel(const double * __restrict__ partial, int G, int mode, double * __restrict__ bscal, double * __restrict__ trace_slot)
{
	__shared__ double sa[32], sb[32];
	double a = 0.0, b = 0.0;
	for (int i = threadIdx.x; i < G; i += blockDim.x) {
		a += partial[i];
		b += partial[G + i];
	}
	a = warp_sum(a);
	b = warp_sum(b);
	if ((threadIdx.x & 31) == 0) {
		sa[threadIdx.x >> 5] = a;
		sb[threadIdx.x >> 5] = b;
	}
	__syncthreads();
	if (threadIdx.x == 0) {
		double ta = 0.0, tb = 0.0;
		for (int w = 0; w < (int) (blockDim.x >> 5); w++) {
			ta += sa[w];
			tb += sb[w];
		}
		if (mode == 0) {
			bscal[B_A] = bscal[B_PKP] / ta;
		} else if (mode == 1) {
			bscal[B_W] = ta / tb;
		} else if (mode == 2) {
			const double s_pk = ta;
			bscal[B_B] = (s_pk / bscal[B_PKP]) * (bscal[B_A] / bscal[B_W]);
			bscal[B_PKP] = s_pk;
			bscal[B_ERR] = sqrt(tb);
			if (trace_slot)
				*trace_slot = sqrt(tb);
		} else {                                        // start: s_pk_p = (r0, r) = sum A, err = sqrt(sum B)
			bscal[B_PKP] = ta;
			bscal[B_ERR] = sqrt(tb);
			if (trace_slot)
				*trace_slot = sqrt(tb);
		}
	}
}

// r = b - Ax; r0 = r; p = r; partials of r.r twice                               (bench_bicg.cpp:218-250)
__global__ void __launch_bounds__(PCG_THREADS)
bi_start_kernel(double * __restrict__ r, double * __restrict__ r0, double * __restrict__ p, const double * __restrict__ b,
                const double * __restrict__ Ax, int64_t m, double * __restrict__ partial)
{
	double s = 0.0;
	for (int64_t i = (int64_t) blockIdx.x * PCG_THREADS + threadIdx.x; i < m; i += (int64_t) gridDim.x * PCG_THREADS) {
		const double ri = b[i] + -1.0 * Ax[i];
		r[i] = ri;
		r0[i] = ri;
		p[i] = ri;
		s += ri * ri;
	}
	block_store_partials(s, s, partial, partial + gridDim.x);
}

__global__ void bi_y_kernel(double * __restrict__ y, const double * __restrict__ p, const double * __restrict__ K, int64_t m)
{
	for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t) gridDim.x * blockDim.x)
		y[i] = p[i] / K[i];
}

__global__ void bi_hs_kernel(double * __restrict__ h, double * __restrict__ s, double * __restrict__ z, const double * __restrict__ x,
                             const double * __restrict__ y, const double * __restrict__ r, const double * __restrict__ v,
                             const double * __restrict__ K, const double * __restrict__ bscal, int64_t m)
{
	const double s_a = bscal[B_A];
	for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t) gridDim.x * blockDim.x) {
		h[i] = x[i] + s_a * y[i];
		const double si = r[i] + -s_a * v[i];
		s[i] = si;
		z[i] = si / K[i];
	}
}

__global__ void __launch_bounds__(PCG_THREADS)
bi_w_kernel(const double * __restrict__ t, const double * __restrict__ s, const double * __restrict__ K, int64_t m, double * __restrict__ partial)
{
	double p1 = 0.0, p2 = 0.0;
	for (int64_t i = (int64_t) blockIdx.x * PCG_THREADS + threadIdx.x; i < m; i += (int64_t) gridDim.x * PCG_THREADS) {
		const double v1 = t[i] / K[i], v2 = s[i] / K[i];
		p1 += v1 * v2;
		p2 += v1 * v1;
	}
	block_store_partials(p1, p2, partial, partial + gridDim.x);
}

__global__ void __launch_bounds__(PCG_THREADS)
bi_update_kernel(double * __restrict__ r, double * __restrict__ x, const double * __restrict__ s, const double * __restrict__ t,
                 const double * __restrict__ h, const double * __restrict__ z, const double * __restrict__ r0,
                 const double * __restrict__ bscal, int64_t m, double * __restrict__ partial)
{
	const double s_w = bscal[B_W];
	double p1 = 0.0, p2 = 0.0;
	for (int64_t i = (int64_t) blockIdx.x * PCG_THREADS + threadIdx.x; i < m; i += (int64_t) gridDim.x * PCG_THREADS) {
		const double ri = s[i] + -s_w * t[i];
		r[i] = ri;
		x[i] = h[i] + s_w * z[i];
		p1 += r0[i] * ri;
		p2 += ri * ri;
	}
	block_store_partials(p1, p2, partial, partial + gridDim.x);
}

__global__ void bi_dir_kernel(double * __restrict__ p, const double * __restrict__ r, const double * __restrict__ v,
                              const double * __restrict__ bscal, int64_t m)
{
	const double s_b = bscal[B_B], s_w = bscal[B_W];
	for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t) gridDim.x * blockDim.x)
		p[i] = r[i] + s_b * (p[i] - s_w * v[i]);
}

extern "C" int b200spmv_bicgstab(b200spmv_t * h, const double * b_host, double * x_host, int64_t max_iterations, double * trace_host,
                                 b200spmv_pcg_result * res)
{
	if (!h || !b_host || !x_host || !res)
		B200_FAIL("NULL argument");
	if (max_iterations < 0)
		B200_FAIL("b200spmv_bicgstab: max_iterations < 0");
	if (!h->d_diag)
		B200_FAIL("b200spmv_bicgstab: the handle was created without keep_diag");
	if (h->diag_bad > 0)
		B200_FAIL("bad K, zero in diagonal (%d rows)", h->diag_bad);          // bench_bicg.cpp:209-210
	CU(cudaSetDevice(h->device));
	const int64_t m = h->m;
	memset(res, 0, sizeof(*res));
	if (m == 0)
		return 0;
	cudaStream_t st = h->stream;
	const int G = (int) ((m + PCG_THREADS - 1) / PCG_THREADS < PCG_CTAS ? (m + PCG_THREADS - 1) / PCG_THREADS : PCG_CTAS);
	const int64_t ntrace = trace_host ? max_iterations + 1 : 0;
	double * mem = nullptr;
	CU(cudaMalloc(&mem, (size_t) (12 * m + 2 * PCG_CTAS + 16 + ntrace) * sizeof(double)));
	double * r0 = mem, * r = mem + m, * re = mem + 2 * m, * p = mem + 3 * m, * z = mem + 4 * m, * hh = mem + 5 * m, * s = mem + 6 * m,
	       * v = mem + 7 * m, * buf = mem + 8 * m, * x = mem + 9 * m, * xb = mem + 10 * m, * b = mem + 11 * m, * partial = mem + 12 * m,
	       * bscal = partial + 2 * PCG_CTAS, * nscal = bscal + 8, * trace = bscal + 16;   // nscal: pcg_final_kernel's slots
	int rc = 1;
	double err = 0, bn = 0, err_explicit = 0, err_best = 0, eps = 0;
	int64_t k = 0, products = 0;
	const int64_t restart_k = 100;
	#define PCG_CHECK(call) do { if ((call) != cudaSuccess) goto done; } while (0)
	#define PCG_RC(call) do { if (call) goto done_msg; } while (0)
	PCG_CHECK(cudaMemcpyAsync(b, b_host, m * sizeof(double), cudaMemcpyHostToDevice, st));
	PCG_CHECK(cudaMemsetAsync(x, 0, 2 * m * sizeof(double), st));                      // x = x_best = 0 (:203-216)
	PCG_CHECK(cudaMemsetAsync(bscal, 0, 16 * sizeof(double), st));
	PCG_RC(launch_spmv_any(h, x, buf, st));
	products++;
	bi_start_kernel<<<G, PCG_THREADS, 0, st>>>(r, r0, p, b, buf, m, partial);
	bi_final_kernel<<<1, 1024, 0, st>>>(partial, G, 3, bscal, trace_host ? trace : nullptr);
	PCG_RC(read_scal(h, bscal, B_ERR, &err));
	pcg_dot2_kernel<<<G, PCG_THREADS, 0, st>>>(b, b, b, b, m, partial);
	pcg_final_kernel<<<1, 1024, 0, st>>>(partial, G, 2, nscal);
	PCG_RC(read_scal(h, nscal, S_ERR, &bn));
	eps = 1.0e-15 * bn;
	err_explicit = err;
	err_best = err;
	while (k < max_iterations) {
		if ((k > 0) && !(k % restart_k)) {                                             // explicit residual, x_best (:307-331)
			PCG_RC(launch_spmv_any(h, x, buf, st));
			products++;
			pcg_residual_kernel<<<G, PCG_THREADS, 0, st>>>(re, b, buf, nullptr, nullptr, h->d_diag, m, partial);
			pcg_final_kernel<<<1, 1024, 0, st>>>(partial, G, 2, nscal);
			PCG_RC(read_scal(h, nscal, S_ERR, &err_explicit));
			if (err_explicit < err_best) {
				PCG_CHECK(cudaMemcpyAsync(xb, x, m * sizeof(double), cudaMemcpyDeviceToDevice, st));
				err_best = err_explicit;
			}
		}
		bi_y_kernel<<<G, PCG_THREADS, 0, st>>>(buf, p, h->d_diag, m);
		PCG_RC(launch_spmv_any(h, buf, v, st));
		pcg_dot2_kernel<<<G, PCG_THREADS, 0, st>>>(r0, v, r0, v, m, partial);
		bi_final_kernel<<<1, 1024, 0, st>>>(partial, G, 0, bscal, nullptr);
		bi_hs_kernel<<<G, PCG_THREADS, 0, st>>>(hh, s, z, x, buf, r, v, h->d_diag, bscal, m);
		PCG_RC(launch_spmv_any(h, z, buf, st));
		products += 2;
		bi_w_kernel<<<G, PCG_THREADS, 0, st>>>(buf, s, h->d_diag, m, partial);
		bi_final_kernel<<<1, 1024, 0, st>>>(partial, G, 1, bscal, nullptr);
		bi_update_kernel<<<G, PCG_THREADS, 0, st>>>(r, x, s, buf, hh, z, r0, bscal, m, partial);
		bi_final_kernel<<<1, 1024, 0, st>>>(partial, G, 2, bscal, trace_host ? trace + k + 1 : nullptr);
		bi_dir_kernel<<<G, PCG_THREADS, 0, st>>>(p, r, v, bscal, m);
		k++;
	}
	PCG_CHECK(cudaGetLastError());
	PCG_RC(read_scal(h, bscal, B_ERR, &err));                                          // ||r_k|| after the last iteration
	PCG_RC(launch_spmv_any(h, x, buf, st));                                            // final explicit residual (:427-446)
	products++;
	pcg_residual_kernel<<<G, PCG_THREADS, 0, st>>>(re, b, buf, nullptr, nullptr, h->d_diag, m, partial);
	pcg_final_kernel<<<1, 1024, 0, st>>>(partial, G, 2, nscal);
	PCG_RC(read_scal(h, nscal, S_ERR, &err_explicit));
	if (err_explicit < err_best) {
		PCG_CHECK(cudaMemcpyAsync(xb, x, m * sizeof(double), cudaMemcpyDeviceToDevice, st));
		err_best = err_explicit;
	}
	PCG_CHECK(cudaMemcpyAsync(x_host, xb, m * sizeof(double), cudaMemcpyDeviceToHost, st));
	if (trace_host)
		PCG_CHECK(cudaMemcpyAsync(trace_host, trace, (size_t) (max_iterations + 1) * sizeof(double), cudaMemcpyDeviceToHost, st));
	PCG_CHECK(cudaStreamSynchronize(st));
	PCG_CHECK(cudaGetLastError());
	res->iterations = k;
	res->products = products;
	res->restarts = 0;
	res->err = err;
	res->err_explicit = err_explicit;
	res->err_best = err_best;
	res->eps = eps;
	res->converged = (err_best < eps) ? 1 : 0;
	rc = 0;
done:
	if (rc)
		snprintf(g_b200_err, sizeof(g_b200_err), "b200spmv_bicgstab: CUDA failure: %s", cudaGetErrorString(cudaGetLastError()));
done_msg:
	cudaFree(mem);
	return rc;
	#undef PCG_CHECK
	#undef PCG_RC
}

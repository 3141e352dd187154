// hbm_probe.cu -- two ceilings that bound SpMV on one B200, measured with CUDA events:
//   (1) read-only streaming bandwidth with the 256-bit evict-first loads the SpMV kernels use;
//   (2) x-gather throughput (8-byte gathers through L1) as a function of the gathered window
//       (L2-resident or not) and of the run length of consecutive columns ("cluster size").
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/hbm_probe tools/hbm_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__global__ void stream256(const double * __restrict__ p, int64_t n4, double * out)
{
	double acc = 0;
	for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t) gridDim.x * blockDim.x) {
		double a, b, c, d;
		asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v4.b64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p + 4 * i));
		acc += a + b + c + d;
	}
	if (acc == 123.456) *out = acc;
}

__global__ void stream128(const double2 * __restrict__ p, int64_t n2, double * out)
{
	double acc = 0;
	for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += (int64_t) gridDim.x * blockDim.x) {
		double2 v = __ldcs(p + i);
		acc += v.x + v.y;
	}
	if (acc == 123.456) *out = acc;
}

// every thread: 8 consecutive indices (one 256-bit load), 8 gathers
__global__ void gather8(const int * __restrict__ idx, const double * __restrict__ x, int64_t n8, double * out)
{
	double acc = 0;
	for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (int64_t) gridDim.x * blockDim.x) {
		int c[8];
		asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
		             : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3]), "=r"(c[4]), "=r"(c[5]), "=r"(c[6]), "=r"(c[7]) : "l"(idx + 8 * i));
		#pragma unroll
		for (int e = 0; e < 8; e++)
			acc += __ldg(x + c[e]);
	}
	if (acc == 123.456) *out = acc;
}

// lane-consecutive indices: thread t of the grid gathers idx[t], idx[t + stride] ... (one index per lane per load)
__global__ void gather1(const int * __restrict__ idx, const double * __restrict__ x, int64_t n, double * out)
{
	double acc = 0;
	const int64_t stride = (int64_t) gridDim.x * blockDim.x;
	int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	for (; i + 3 * stride < n; i += 4 * stride) {
		int c0 = idx[i], c1 = idx[i + stride], c2 = idx[i + 2 * stride], c3 = idx[i + 3 * stride];
		acc += __ldg(x + c0) + __ldg(x + c1) + __ldg(x + c2) + __ldg(x + c3);
	}
	if (acc == 123.456) *out = acc;
}

static float time_it(void (*launch)(void *), void * arg, int reps)
{
	cudaEvent_t a, b;
	cudaEventCreate(&a); cudaEventCreate(&b);
	launch(arg); launch(arg);
	CK(cudaDeviceSynchronize());
	float best = 1e30f;
	for (int r = 0; r < reps; r++) {
		cudaEventRecord(a); launch(arg); cudaEventRecord(b);
		CK(cudaEventSynchronize(b));
		float ms; cudaEventElapsedTime(&ms, a, b);
		if (ms < best) best = ms;
	}
	return best;
}

struct Args { const void * p; const double * x; int64_t n; double * out; int grid; };
static void l_stream256(void * a) { Args * g = (Args *) a; stream256<<<g->grid, 256>>>((const double *) g->p, g->n / 4, g->out); }
static void l_stream128(void * a) { Args * g = (Args *) a; stream128<<<g->grid, 256>>>((const double2 *) g->p, g->n / 2, g->out); }
static void l_gather8(void * a) { Args * g = (Args *) a; gather8<<<g->grid, 256>>>((const int *) g->p, g->x, g->n / 8, g->out); }
static void l_gather1(void * a) { Args * g = (Args *) a; gather1<<<g->grid, 256>>>((const int *) g->p, g->x, g->n, g->out); }

int main()
{
	cudaDeviceProp pr; CK(cudaGetDeviceProperties(&pr, 0));
	int l2 = 0, persist = 0;
	cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, 0);
	cudaDeviceGetAttribute(&persist, cudaDevAttrMaxPersistingL2CacheSize, 0);
	printf("device %s, %d SMs, L2 %d MB, max persisting L2 %d MB, mem %.0f GB\n", pr.name, pr.multiProcessorCount, l2 >> 20, persist >> 20, pr.totalGlobalMem / 1e9);
	double * out; CK(cudaMalloc(&out, 8));
	// (1) streaming
	const int64_t n = 1LL << 29;   // 4 GiB of doubles
	double * buf; CK(cudaMalloc(&buf, n * 8)); CK(cudaMemset(buf, 0, n * 8));
	for (int occ : {2, 4, 8}) {
		Args a = { buf, nullptr, n, out, pr.multiProcessorCount * occ };
		float t1 = time_it(l_stream256, &a, 5), t2 = time_it(l_stream128, &a, 5);
		printf("stream read 4 GiB, %d CTAs/SM: 256-bit evict_first %.1f GB/s | 128-bit ldcs %.1f GB/s\n", occ, n * 8 / t1 * 1e-6, n * 8 / t2 * 1e-6);
	}
	CK(cudaFree(buf));
	// (2) gathers
	const int64_t ng = 1LL << 27;  // 128 M gathers
	std::vector<int> h(ng);
	int * idx; CK(cudaMalloc(&idx, ng * 4));
	const int64_t xmax = 1LL << 28;   // 2 GiB of x at most
	double * x; CK(cudaMalloc(&x, xmax * 8)); CK(cudaMemset(x, 0, xmax * 8));
	uint64_t s = 88172645463325252ULL;
	auto rnd = [&]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; };
	for (int64_t win : { 300000LL, 5000000LL, 17000000LL, 200000000LL }) {          // x window in doubles: 2.4 MB, 40 MB, 136 MB, 1.6 GB
		for (int run : { 1, 2, 4, 16 }) {
			for (int64_t i = 0; i < ng; ) {
				int64_t c = rnd() % (win - run);
				for (int r = 0; r < run && i < ng; r++) h[i++] = (int) (c + r);
			}
			CK(cudaMemcpy(idx, h.data(), ng * 4, cudaMemcpyHostToDevice));
			Args a = { idx, x, ng, out, pr.multiProcessorCount * 8 };
			float t8 = time_it(l_gather8, &a, 3), t1 = time_it(l_gather1, &a, 3);
			printf("gather window %6.1f MB run %2d: 8-per-thread %.1f Ggather/s (= %.0f GB/s of 12 B/nnz) | lane-consecutive %.1f Ggather/s (= %.0f GB/s)\n",
			       win * 8 / 1e6, run, ng / t8 * 1e-6, ng / t8 * 1e-6 * 12, ng / t1 * 1e-6, ng / t1 * 1e-6 * 12);
		}
	}
	return 0;
}

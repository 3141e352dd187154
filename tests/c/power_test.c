/* power_test.c -- a plain C host program driving the multi-GPU power iteration through the C-ABI alone (include/b200spmv.h,
 * include/b200amg.h): no Python, no torch, no CUDA runtime of its own.  It forks one process per rank (one GPU each), hands
 * the 128-byte NCCL id from rank 0 to the others through a file, builds each rank's row shard of a generator matrix
 * (reference partitioner on the GPU), runs both exchanges, and checks EVERY step against the serial CSR kernel of the
 * oracle (oracle/liboracle.so: subkernel_csr_scalar, K/csr.cpp:334-350) under the per-row bound
 * |y - y_ref| <= 1e-12 * sum_j |a_ij x_j|.  Test infrastructure: the oracle is linked here as the checker only.
 *
 *   make -C tests/c && tests/c/power_test [world = 2] [rows = 400000]          exit code 0 = every rank passed
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/wait.h>
#include <unistd.h>
#include "b200spmv.h"

void oracle_csr_spmv_scalar(const int32_t *row_ptr, const int32_t *col_ind, const double *val, const double *x, double *y, int64_t i_s, int64_t i_e);
void oracle_row_abs_bound(const int32_t *row_ptr, const int32_t *col_ind, const double *val, const double *x, double *bound, int64_t m);

#define CHECK(call)                                                                           \
	do {                                                                                  \
		if (call) {                                                                   \
			fprintf(stderr, "rank %d: %s failed: %s\n", rank, #call, b200spmv_last_error()); \
			return 1;                                                             \
		}                                                                             \
	} while (0)

static int run_rank(int rank, int world, int64_t rows, const char * id_path)
{
	b200amg_params p;
	memset(&p, 0, sizeof(p));
	p.nr_rows = p.nr_cols = rows;
	p.avg_nnz_per_row = 30; p.std_nnz_per_row = 10; p.distribution = 'n'; p.placement = 'r';
	p.bw_scaled = 0.3; p.skew = 0; p.avg_num_neighbours = 0.5; p.cross_row_similarity = 0.5; p.seed = 14;
	CHECK(b200spmv_set_device(rank));
	/* pass 1 on every rank (row lengths), then only this rank's rows */
	int32_t * rp = malloc((size_t) (rows + 1) * sizeof(int32_t));
	double bw = 0.0;
	const int64_t nnz = b200amg_row_ptr(&p, rp, &bw);
	if (nnz < 0) { fprintf(stderr, "rank %d: generator: %s\n", rank, b200amg_last_error()); return 1; }
	int64_t bounds[16];
	CHECK(b200spmv_partition(rp, rows, nnz, world, bounds));
	const int64_t i_s = bounds[2 * rank], i_e = bounds[2 * rank + 1], m = i_e - i_s, nnz_loc = rp[i_e] - rp[i_s];
	int32_t * col = malloc((size_t) (nnz_loc > 0 ? nnz_loc : 1) * sizeof(int32_t));
	double * val = malloc((size_t) (nnz_loc > 0 ? nnz_loc : 1) * sizeof(double));
	if (b200amg_fill(&p, rp, bw, i_s, i_e, col, val)) { fprintf(stderr, "rank %d: fill: %s\n", rank, b200amg_last_error()); return 1; }
	int32_t * rp_loc = malloc((size_t) (m + 1) * sizeof(int32_t));
	for (int64_t i = 0; i <= m; i++)
		rp_loc[i] = rp[i_s + i] - rp[i_s];

	unsigned char id[128];
	b200spmv_comm_t * comm = NULL;
	if (world > 1) {
		if (rank == 0) {
			CHECK(b200spmv_comm_unique_id(id));
			char tmp[512];
			snprintf(tmp, sizeof(tmp), "%s.tmp", id_path);
			FILE * f = fopen(tmp, "wb");
			fwrite(id, 1, 128, f);
			fclose(f);
			rename(tmp, id_path);
		} else {
			FILE * f = NULL;
			for (int t = 0; t < 2400 && !(f = fopen(id_path, "rb")); t++)
				usleep(50000);
			if (!f || fread(id, 1, 128, f) != 128) { fprintf(stderr, "rank %d: no NCCL id\n", rank); return 1; }
			fclose(f);
		}
		CHECK(b200spmv_comm_create(&comm, id, rank, world, rank));
	}

	double * x = malloc((size_t) rows * sizeof(double)), * xn = malloc((size_t) rows * sizeof(double));
	double * xr = malloc((size_t) rows * sizeof(double)), * y_ref = malloc((size_t) (m + 1) * sizeof(double)), * bound = malloc((size_t) (m + 1) * sizeof(double));
	int fails = 0;
	double eig_by_mode[4] = { 0, 0, 0, 0 };
	/* mode 0: NCCL all-gather, 4 chunks; 1: fused stores + ncclAllReduce; 2: fused stores + mailbox barrier, 2 column panels
	 * interleaved over 4 row chunks; 3: as 2 under a CUDA graph */
	for (int mode = 0; mode < 4; mode++) {
		b200spmv_opts o;
		b200spmv_default_opts(&o);
		o.device = rank;
		o.col_panels = (mode >= 2) ? 2 : 1;
		b200spmv_t * h = NULL;
		CHECK(b200spmv_create(&h, rp_loc, col, val, m, rows, nnz_loc, &o));
		int64_t c_lo, c_hi;
		CHECK(b200spmv_col_range(h, &c_lo, &c_hi));
		b200spmv_power_opts po;
		memset(&po, 0, sizeof(po));
		po.exchange = (mode == 0) ? B200SPMV_X_NCCL : B200SPMV_X_FUSED;
		po.chunks = (mode == 1) ? 0 : 4;
		po.scalar_sync = (mode >= 2) ? 1 : 0;
		po.use_graph = (mode == 3);
		b200spmv_power_t * pw = NULL;
		CHECK(b200spmv_power_create(&pw, h, comm, bounds, rows, &po));
		for (int64_t i = 0; i < rows; i++)
			x[i] = 0.5 + (double) ((i * 2654435761u) & 1023) / 1024.0;
		CHECK(b200spmv_power_set_start(pw, x));
		double eig = 0.0, worst = 0.0;
		if (mode == 3)
			CHECK(b200spmv_power_iterate(pw, 6, &eig, NULL));       /* captures the graph, replays it */
		for (int step = 0; step < 3; step++) {
			double s = 0.0;
			CHECK(b200spmv_power_get_x(pw, x, &s));
			memset(xr, 0, (size_t) rows * sizeof(double));
			for (int64_t j = c_lo; j < c_hi; j++)
				xr[j] = x[j];
			oracle_csr_spmv_scalar(rp_loc, col, val, xr, y_ref, 0, m);
			oracle_row_abs_bound(rp_loc, col, val, xr, bound, m);
			CHECK(b200spmv_power_iterate(pw, 1, &eig, NULL));
			CHECK(b200spmv_power_get_x(pw, xn, NULL));
			for (int64_t i = 0; i < m; i++) {
				const double e = fabs(xn[i_s + i] - s * y_ref[i]), b = 1e-12 * fabs(s) * bound[i];
				if (!(e <= b)) {
					if (fails < 5)
						fprintf(stderr, "rank %d mode %d step %d row %lld: %.17g vs %.17g (bound %.3g)\n", rank, mode, step, (long long) (i_s + i), xn[i_s + i], s * y_ref[i], b);
					fails++;
				}
				if (b > 0 && e / b > worst)
					worst = e / b;
			}
		}
		int64_t info[7];
		CHECK(b200spmv_power_info(pw, info, 7));
		eig_by_mode[mode] = eig;
		printf("rank %d mode %d: chunks %lld panels %lld band [%lld, %lld) graph %lld launches/iter %lld nccl/iter %lld  worst err/bound %.3g  eig %.15g\n",
		       rank, mode, (long long) info[0], (long long) info[1], (long long) info[2], (long long) info[3], (long long) info[4],
		       (long long) info[5], (long long) info[6], worst, eig);
		b200spmv_power_destroy(pw);
		b200spmv_destroy(h);
	}
	/* modes 0-2 ran the same number of iterations from the same start: same eigenvalue estimate up to the summation order */
	for (int mode = 1; mode < 3; mode++)
		if (!(fabs(eig_by_mode[mode] - eig_by_mode[0]) <= 1e-12 * fabs(eig_by_mode[0]))) {
			fprintf(stderr, "rank %d: eigenvalue of mode %d (%.17g) differs from mode 0 (%.17g)\n", rank, mode, eig_by_mode[mode], eig_by_mode[0]);
			fails++;
		}
	b200spmv_comm_destroy(comm);
	return fails ? 1 : 0;
}

int main(int argc, char ** argv)
{
	const int world = argc > 1 ? atoi(argv[1]) : 2;
	const int64_t rows = argc > 2 ? atoll(argv[2]) : 400000;
	int32_t ndev = 0;
	if (b200spmv_device_count(&ndev) || ndev < world) {
		fprintf(stderr, "power_test: %d GPUs needed, %d present\n", world, ndev);
		return 77;
	}
	char id_path[256];
	snprintf(id_path, sizeof(id_path), "/tmp/b200spmv_power_test_%d.id", (int) getpid());
	/* NOTE: the device count query above initialised CUDA in this process; the ranks must not inherit that context, so they
	 * re-exec this binary instead of continuing after fork() */
	if (argc > 3) {
		const int rank = atoi(argv[3]);
		return run_rank(rank, world, rows, argv[4]);
	}
	pid_t pid[8];
	for (int r = 0; r < world; r++) {
		pid[r] = fork();
		if (pid[r] == 0) {
			char a1[16], a2[32], a3[16];
			snprintf(a1, sizeof(a1), "%d", world);
			snprintf(a2, sizeof(a2), "%lld", (long long) rows);
			snprintf(a3, sizeof(a3), "%d", r);
			execl("/proc/self/exe", argv[0], a1, a2, a3, id_path, (char *) NULL);
			_exit(127);
		}
	}
	int bad = 0;
	for (int r = 0; r < world; r++) {
		int st = 0;
		waitpid(pid[r], &st, 0);
		if (!WIFEXITED(st) || WEXITSTATUS(st) != 0)
			bad++;
	}
	unlink(id_path);
	printf("power_test: world %d, %s\n", world, bad ? "FAILED" : "all ranks passed");
	return bad ? 1 : 0;
}

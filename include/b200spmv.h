/*
 * b200spmv.h -- C-ABI of the B200-native fp64 SpMV engine (libb200spmv.so, sm_100a only).
 *
 * Drop-in boundary.  These entry points are what a format plugin of the reference harness
 * binds (reference paths relative to /root/reference, K/ = benchmark_code/BENCH/src/spmv_kernels/):
 *
 *   b200spmv_create()        <- csr_to_format(INT_T* row_ptr, INT_T* col_ind, ValueTypeReference* values,
 *                                             long m, long n, long nnz, long symmetric, long symmetry_expanded)
 *                               K/spmv_kernel.h:28 ; CSR ctor K/csr.cpp:58-167,221-240 ;
 *                               SELL-C-sigma ctor K/sell_c_s.cpp:58-75 -> K/sell-C-s/RISC-V/sellcs_format.c:137-224
 *   b200spmv_spmv_host()     <- Matrix_Format::spmv(ValueType* x, ValueType* y)   K/spmv_kernel.h:16
 *                               (reference GPU-plugin convention: K/GPU_clean/cuda_csr.cu:353-430)
 *   b200spmv_spmv_device()   <- the same product with x / y already resident in HBM (iterated SpMV;
 *                               nearest reference caller benchmark_code/BENCH/src/bench_cg.cpp:93-292)
 *   b200spmv_power_*()       <- net-new multi-GPU layer (row shards by the reference partitioner, see below)
 *   b200spmv_partition*()    <- loop_partitioner_balance_prefix_sums()            lib/parallel_util.h:156-184
 *   b200spmv_merge_coords()  <- MergePathSearch()                                 K/merge.cpp:230-251
 *   b200spmv_export()        <- read-back of the converted layout for bit-exact tests
 *                               (sellcs_matrix_t fields, K/sell-C-s/RISC-V/sellcs-spmv.h:28-47)
 *   b200spmv_last_error()    <- error() / gpuCudaErrorCheck print-and-exit        lib/debug.h:135, lib/cuda/cuda_util.h:5-14
 *
 * Conventions: plain pointers and sizes, no C++ / torch types; every function returns 0 on
 * success and non-zero on failure (message via b200spmv_last_error()); nothing throws.  A handle
 * is not thread-safe (the reference calls csr_to_format / spmv from one thread).  There is no CPU
 * fallback: without a CUDA device every compute entry point fails.
 * INT_T = int32_t, ValueType = double (the reference's `_d` build, make.sh:166,212-213).
 */
#ifndef B200SPMV_H
#define B200SPMV_H

#include <stdint.h>
#include "b200amg.h"      /* b200amg_params (the generator's 11 parameters) */

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200spmv b200spmv_t;

enum {
	B200SPMV_CSR_ADAPTIVE = 0,   /* nnz-tiled adaptive CSR (thread / sub-warp / warp / block per row) */
	B200SPMV_SELL_C_SIGMA = 1,   /* SELL-C-sigma, converted on the GPU                               */
	B200SPMV_AUTO         = 2    /* both are built at create time, each timed on the device for a few
	                                products, the faster one is kept (b200spmv_format() says which)  */
};

/* b200spmv_spmv_host_ex flags */
enum {
	B200SPMV_SKIP_X_UPLOAD   = 1,   /* x already uploaded by an earlier call (reference convention)  */
	B200SPMV_SKIP_Y_DOWNLOAD = 2    /* leave y in HBM                                                  */
};

/* b200spmv_export selectors */
enum {
	B200SPMV_X_ROW_ORDER   = 1,    /* int32[m]            sellcs row_order                            */
	B200SPMV_X_SLICE_WIDTH = 2,    /* int32[nslices]      sellcs slice_widths                         */
	B200SPMV_X_SLICE_PTR   = 3,    /* int64[nslices+1]    sellcs slice_pointers                       */
	B200SPMV_X_SELL_COL    = 4,    /* int32[padded]       sellcs column_indices                       */
	B200SPMV_X_SELL_VAL    = 5,    /* double[padded]      sellcs values                               */
	B200SPMV_X_TILE_ROW    = 6,    /* int32[ntiles+1]     first row owned by each CSR tile (with compressed
	                                                       rows: first NON-EMPTY row starting in the tile)   */
	B200SPMV_X_SPLIT_ROW   = 7,    /* int32[nsplit]       rows finished by the fix-up kernel          */
	B200SPMV_X_SPLIT_TILE  = 8     /* int32[nsplit]       owner tile of each split row                */
};

typedef struct b200spmv_opts {
	int32_t format;           /* B200SPMV_CSR_ADAPTIVE | B200SPMV_SELL_C_SIGMA                          */
	int32_t device;           /* CUDA device ordinal, -1 = current                                      */
	int32_t sell_C;           /* slice height: power of two in [2, 256]; default 32                     */
	int32_t sell_sigma;       /* sort window, multiple of sell_C; default 16384 (K/sell_c_s.cpp:59)     */
	int32_t csr_nnz_per_thread; /* 4 | 8 | 16 | 32 non-zeros per thread of a 256-thread tile; 0 = auto  */
	int32_t csr_ext;          /* tail-extension capacity in nnz (multiple of 64 in [64, 8192]); 0 = auto               */
	int32_t x_hint;           /* 0: L2 evict_last hint on x gathers (default), 1: plain loads           */
	int32_t sell_sigma_alt;   /* AUTO only: second sort window tried (0 = default 256, -1 = none)       */
	int32_t sell_variant;     /* SELL kernel shape (k steps in flight x resident warps per SM): 0 = chosen from
	                             the average slice width (AUTO: timed), 1 = 8 x 32, 2 = 8 x 48, 3 = 4 x 64,
	                             4 = 16 x 16                                                             */
	int32_t csr_compress_rows; /* CSR: plan and run on the non-empty rows only, zero the empty ones in a separate balanced
	                             launch: 0 = when a tile would own more than 2 tiles' worth of rows, 1 = never, 2 = always */
	int32_t reserved0;        /* must be 0 (keeps the layout of earlier builds)                                       */
	int32_t keep_diag;        /* 1: extract diag(A) at create time (square matrices; needed by b200spmv_pcg)            */
	int32_t pin_host_vectors; /* 1: b200spmv_spmv_host page-locks (cudaHostRegister) each distinct x / y buffer it is handed
	                             the first time it sees it (the harness's aligned_alloc vectors live for the whole run,
	                             bench_spmv.cpp:598-606) and unregisters them in b200spmv_destroy                    */
	int32_t col_panels;       /* column panels for an x band larger than L2: 0 = auto (band > 72 MB at >= 24 nnz / row, > 112 MB
	                             otherwise: max(2, ceil(band / 88 MB)) panels), 1 = none, 2..8 = that many; each panel is a sub-handle of the
	                             requested format, launched in ascending column order, y accumulated                  */
	int32_t host_pipeline;    /* b200spmv_spmv_host: 1 = upload x range by range (the columns each unit range of rows references,
	                             found at create time) on a third stream, so that the H2D copy of x overlaps the D2H copy of y
	                             (PCIe is full duplex) as well as the kernels; 0 (default) = upload the referenced band of x
	                             first, then overlap kernels and D2H only.  Same results either way.                */
	int32_t reserved[1];
} b200spmv_opts;

/* extra outputs / inputs of one device-resident product (all pointers are DEVICE pointers) */
typedef struct b200spmv_launch {
	const double * scale;     /* optional scalar s: y = s * (A x)        (power iteration, A(s x) = s A x) */
	double * norm2;           /* optional: receives sum_i y_i^2 of this handle's rows                    */
	double * y_peer[8];       /* optional extra destinations written like y (peer GPUs' next-x buffers,
	                             or one NVSwitch multicast address covering all of them)                 */
	int32_t n_peer;
	int32_t pad_;
	int64_t peer_row_lo;      /* only rows [peer_row_lo, peer_row_hi) of this handle are stored to y_peer  */
	int64_t peer_row_hi;      /* (the rows some rank's columns reference); both 0 = every row             */
} b200spmv_launch;

void b200spmv_default_opts(b200spmv_opts * opts);

/* Deep-copies / uploads the CSR arrays (HOST pointers, borrowed for the call only) and builds the format. */
int b200spmv_create(b200spmv_t ** out, const int32_t * row_ptr, const int32_t * col_ind, const double * values,
                    int64_t m, int64_t n, int64_t nnz, const b200spmv_opts * opts);

/* Same, from DEVICE pointers (copied device-to-device).  row_ptr may start at a non-zero offset
 * (a row shard of a larger matrix): it is rebased; col_ind / values point at the shard's first entry. */
int b200spmv_create_device(b200spmv_t ** out, const int32_t * d_row_ptr, const int32_t * d_col_ind, const double * d_values,
                           int64_t m, int64_t n, int64_t nnz, const b200spmv_opts * opts);

void b200spmv_destroy(b200spmv_t * h);

/* y = A x with HOST vectors: H2D of x (the referenced column band, see b200spmv_col_range), kernel(s), D2H of y,
 * stream synchronised on return. */
int b200spmv_spmv_host(b200spmv_t * h, const double * x, double * y);
int b200spmv_spmv_host_ex(b200spmv_t * h, const double * x, double * y, int32_t flags);

/* y = A x with DEVICE vectors on `stream` (a cudaStream_t; NULL = the CUDA legacy default stream); asynchronous. */
int b200spmv_spmv_device(b200spmv_t * h, const double * dx, double * dy, void * stream);
int b200spmv_spmv_device_ex(b200spmv_t * h, const double * dx, double * dy, void * stream, const b200spmv_launch * launch);

/* The format that runs: B200SPMV_CSR_ADAPTIVE or B200SPMV_SELL_C_SIGMA (what AUTO resolved to). */
int32_t b200spmv_format(const b200spmv_t * h);

/* Columns referenced by the matrix lie in [*lo, *hi) (found on the GPU at create time).  The products never read x
 * outside that range, so b200spmv_spmv_host uploads only x[lo..hi) and a multi-GPU exchange only needs those rows. */
int b200spmv_col_range(const b200spmv_t * h, int64_t * lo, int64_t * hi);

/* Number of kernel launches one b200spmv_spmv_device call issues for this handle (CSR: 1, + 1 with compressed rows; SELL: 1, + 1 with cut slices; panels: the sum). */
int32_t b200spmv_launches_per_spmv(const b200spmv_t * h);

/* Format statistics: out[0]=mem_footprint bytes, [1]=ntiles or nslices, [2]=nsplit rows,
 * [3]=padded entries (SELL) or nnz, [4]=tile nnz (CSR) or C (SELL), [5]=ext capacity or sigma, [6]=convert ms,
 * [7..9]=AUTO trial ms per product of CSR, SELL(sell_sigma), SELL(sell_sigma_alt); 0 = not tried,
 * [10]=1 when the CSR plan runs on compressed (non-empty) rows, [11]=number of empty rows zeroed separately,
 * [12]=number of column panels (1 = none; with panels [0] and [6] are totals, the rest describes panel 0). */
int b200spmv_stats(const b200spmv_t * h, double * out, int32_t n_out);

int b200spmv_export(const b200spmv_t * h, int32_t what, void * dst_host, int64_t capacity_bytes, int64_t * count_out);

/* Reference partitioner on the GPU: bounds_out[2p], bounds_out[2p+1] = [i_s, i_e) of worker p of W.
 * sums has N+1 entries of which sums[N] is never read (as in the reference); total = sums[N]-sums[0]. */
int b200spmv_partition_device(const int32_t * d_sums, int64_t N, int64_t total, int32_t W, int64_t * bounds_out_host);
int b200spmv_partition(const int32_t * sums_host, int64_t N, int64_t total, int32_t W, int64_t * bounds_out_host);

/* The same partition as b200spmv_partition_device for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  No counterpart in the reference.
 * Allocates nothing (no cudaMalloc / cudaFree / allocating library routine): the 2 W bounds are staged in a
 * file-scope static device array of B200SPMV_NOALLOC_MAX_WORKERS workers (64 KB, held by every GPU that loads the library
 * until exit); W beyond that is refused with an error.  d_sums: DEVICE pointer; bounds_out_host: HOST, 2 W entries; the
 * kernel and the copy run on `stream` (NULL = legacy default stream), which is synchronised before returning.  Every call
 * on a device shares the one array (it keeps the last call's contents; every entry that is read back is overwritten first):
 * calls that overlap on the GPU -- other streams, other threads -- race; calls on one stream are safe. */
#define B200SPMV_NOALLOC_MAX_WORKERS 4096
int b200spmv_partition_device_noalloc(const int32_t * d_sums, int64_t N, int64_t total, int32_t W, int64_t * bounds_out_host, void * stream);

/* Merge-path coordinates on the GPU for W workers: coords_out[2t], [2t+1] = (row, nz) where worker t
 * starts, t = 0..W (W+1 pairs), items per worker = ceil((m+nnz)/W) as K/merge.cpp:277-283. */
int b200spmv_merge_coords(const b200spmv_t * h, int32_t W, int64_t * coords_out_host);

/* Device memory helpers so that a host program needs no CUDA runtime of its own. */
int b200spmv_device_count(int32_t * count);
int b200spmv_set_device(int32_t device);
int b200spmv_malloc(void ** dptr, int64_t bytes);
int b200spmv_free(void * dptr);
int b200spmv_memcpy_h2d(void * dst, const void * src, int64_t bytes);
int b200spmv_memcpy_d2h(void * dst, const void * src, int64_t bytes);
int b200spmv_malloc_host(void ** hptr, int64_t bytes);      /* pinned */
int b200spmv_free_host(void * hptr);
int b200spmv_synchronize(void);
/* CUDA IPC: export a device allocation to / import one from another process of the same box. */
int b200spmv_ipc_export(void * dptr, unsigned char handle_out[64]);
int b200spmv_ipc_import(const unsigned char handle[64], void ** dptr_out);
int b200spmv_ipc_close(void * dptr);

/* Device timing of `iters` back-to-back products (CUDA events on the handle's stream); when
 * flush_bytes > 0 a scratch buffer of that size is rewritten before every product (outside the
 * per-product event pair) so that nothing of the matrix survives in L2.  ms_out[iters]. */
int b200spmv_time_device(b200spmv_t * h, const double * dx, double * dy, int32_t iters, int64_t flush_bytes, float * ms_out);

/* Jacobi-preconditioned CG entirely on the device: the reference harness's preconditioned_cg()
 * (benchmark_code/BENCH/src/bench_cg.cpp:93-320: x0 = 0, K = diag(A), eps = 1e-15 ||b||, explicit residual / x_best /
 * restart every 100 iterations) over b200spmv_spmv_device.  b, x: HOST vectors of m entries; the handle must have been
 * created with keep_diag = 1.  Fails with "bad K, zero in diagonal" like the reference when a diagonal entry is missing. */
typedef struct b200spmv_pcg_result {
	int64_t iterations;       /* k at exit (num_loops_out, bench_cg.cpp:318)                            */
	int64_t products;         /* SpMV products issued                                                   */
	int64_t restarts;
	double err;               /* ||r_k|| (recurrence)                                                   */
	double err_explicit;      /* ||b - A x|| at exit                                                    */
	double err_best;          /* of the returned x_best                                                 */
	double eps;               /* 1e-15 ||b||                                                            */
	int32_t converged;        /* err < eps                                                              */
	int32_t pad_;
} b200spmv_pcg_result;
int b200spmv_pcg(b200spmv_t * h, const double * b_host, double * x_host, int64_t max_iterations, b200spmv_pcg_result * result);

/* Jacobi-preconditioned BiCGSTAB entirely on the device: the reference harness's preconditioned_bicgstab()
 * (benchmark_code/BENCH/src/bench_bicg.cpp:150-456).  Like the reference it has NO stop test (:333-334 are commented out):
 * exactly max_iterations iterations; the returned x is x_best, the iterate with the smallest explicit residual seen at the
 * checkpoints (every 100 iterations) and at the end -- so, as in the reference, a cap far beyond convergence can return an
 * older iterate (or x0) once the recurrence degenerates to 0/0.  Two products per iteration, no host synchronisation
 * between checkpoints.  trace_host: NULL, or max_iterations + 1 doubles receiving ||r_k|| (recurrence) for k = 0 ..
 * max_iterations (the harness's "k = .. error = .." line, :343).  result: iterations = max_iterations, restarts = 0,
 * err = ||r_k|| after the last iteration, converged = (err_best < eps).  Needs keep_diag = 1. */
int b200spmv_bicgstab(b200spmv_t * h, const double * b_host, double * x_host, int64_t max_iterations, double * trace_host,
                      b200spmv_pcg_result * result);

/* COO -> CSR on the GPU: the reference's coo_to_csr() (lib/storage_formats/csr/csr_gen.c:178-213, called by import_file(),
 * benchmark_code/BENCH/src/bench.cpp:208-229) -- entries bucketed by row, columns ascending inside every row, equal
 * (row, col) pairs kept in input order.  All pointers are DEVICE pointers; indices base 0; d_row_ptr has m + 1 entries,
 * d_col_out / d_val_out nnz.  *duplicates_out (HOST, optional) = number of adjacent equal (row, col) pairs in the result.
 * Fails on an index outside [0, m) x [0, n).  Synchronises `stream` before returning. */
int b200spmv_coo_to_csr(const int32_t * d_row, const int32_t * d_col, const double * d_val, int64_t m, int64_t n, int64_t nnz,
                        int32_t * d_row_ptr, int32_t * d_col_out, double * d_val_out, int64_t * duplicates_out, void * stream);

/* import_file() + csr_to_format() for a matrix that arrives as COO (HOST arrays, e.g. from b200mtx_read() with the symmetry
 * expanded, include/b200mtx.h): upload, COO -> CSR on the GPU, format built from the device arrays.  Duplicate (row, col)
 * entries are refused (the reference leaves their order to an unstable sort; there is nothing bit-exact to reproduce). */
int b200spmv_create_coo(b200spmv_t ** out, const int32_t * row, const int32_t * col, const double * val, int64_t m, int64_t n,
                        int64_t nnz, const b200spmv_opts * opts);

/* Pass 2 of the synthetic matrix source on the GPU: the device counterpart of b200amg_fill() (include/b200amg.h), i.e. of
 * the non-zero placement half of the reference's artificial_matrix_generation()
 * (artificial-matrix-generator/artificial_matrix_generation.c:366-523; called from benchmark_code/BENCH/src/bench.cpp:250).
 * d_row_ptr: the WHOLE matrix's row_ptr (nr_rows + 1 entries, from b200amg_row_ptr -- pass 1 stays on the host because it
 * is the only part that depends on glibc's libm) in device memory; bw_corrected: pass 1's second output.  Writes the
 * non-zeros of rows [i_s, i_e) to d_col_ind / d_values (DEVICE pointers; entry j of the whole matrix lands at
 * j - row_ptr[i_s]), bit-identical to the reference generator, ready for b200spmv_create_device().  One warp per anchor
 * segment (<= ~4100 independent random streams); synchronises `stream` before returning. */
int b200spmv_amg_fill_device(const b200amg_params * p, const int32_t * d_row_ptr, double bw_corrected, int64_t i_s, int64_t i_e,
                             int32_t * d_col_ind, double * d_values, void * stream);

/* ---- fp32: the reference's `_f` build axis (benchmark_code/BENCH/src/make.sh:215-216: ValueType = float,
 * ValueTypeReference = double).  csr_to_format still receives DOUBLE values and narrows them (K/csr.cpp:58-167; here on
 * the GPU, round to nearest even like the C cast); Matrix_Format::spmv(ValueType * x, ValueType * y) takes float vectors
 * and accumulates in float (K/csr.cpp:334-350).  CSR only; a separate handle type, the fp64 path is untouched. */
typedef struct b200spmv_f32 b200spmv_f32_t;
int b200spmv_f32_create(b200spmv_f32_t ** out, const int32_t * row_ptr, const int32_t * col_ind, const double * values, int64_t m,
                        int64_t n, int64_t nnz, int32_t device /* -1 = current */);
int b200spmv_f32_spmv_host(b200spmv_f32_t * h, const float * x, float * y);                       /* host vectors, synchronised */
int b200spmv_f32_spmv_device(b200spmv_f32_t * h, const float * dx, float * dy, void * stream);    /* device vectors, asynchronous */
int b200spmv_f32_time_device(b200spmv_f32_t * h, const float * dx, float * dy, int32_t iters, int64_t flush_bytes, float * ms_out);
void b200spmv_f32_destroy(b200spmv_f32_t * h);


/* ---- multi-GPU iterated SpMV (power iteration) over a row-sharded matrix: one process per GPU (net-new, the
 * reference has no multi-GPU code -- what it contributes is the row partition: rank p owns the rows that
 * loop_partitioner_balance_prefix_sums(), lib/parallel_util.h:156-184, gives worker p of P, see b200spmv_partition*).
 * The iteration loop, the exchange of y (the next x) and the scalar arithmetic all run under this C-ABI; a caller
 * needs nothing but a way to hand a 128-byte id from rank 0 to the other ranks (a file, MPI, a torch store).
 *   exchange B200SPMV_X_NCCL : y slots all-gathered over NVLink with NCCL (grouped ncclSend / ncclRecv, unequal counts, in place in
 *                              the next-x buffer), band-limited (only rows some rank's columns reference), in `chunks` row
 *                              chunks so that chunk c travels under the kernel of chunk c + 1; ncclAllReduce of the norm.
 *   exchange B200SPMV_X_FUSED: the SpMV epilogue stores every finished row of the band straight into all peers' next-x
 *                              buffers (CUDA IPC mappings, NVLink P2P stores); the only synchronisation left is the scalar:
 *                              ncclAllReduce (scalar_sync 0) or a mailbox barrier over peer memory (scalar_sync 1).
 * With a column-panel shard (opts.col_panels) the launches are interleaved by row chunk -- chunk 0 of panel 0, chunk 0 of
 * panel 1, chunk 1 of panel 0, ... -- so rows become final (and leave over NVLink) progressively while each panel's part
 * of x stays L2-resident across a chunk.  NCCL is bound at run time (dlopen of libnccl.so.2, or $B200SPMV_NCCL_LIB). */
typedef struct b200spmv_comm b200spmv_comm_t;
typedef struct b200spmv_power b200spmv_power_t;
enum { B200SPMV_X_NCCL = 0, B200SPMV_X_FUSED = 1 };

typedef struct b200spmv_power_opts {
	int32_t exchange;         /* B200SPMV_X_NCCL | B200SPMV_X_FUSED                                                   */
	int32_t chunks;           /* row chunks per iteration, 1..16; 0 = auto (4 when the shard can be launched by tile ranges)  */
	int32_t full_exchange;    /* 1 = exchange every row, 0 (default) = only the union of the shards' referenced column bands */
	int32_t use_graph;        /* 1 = two iterations (one per buffer parity) are captured into a CUDA graph and replayed     */
	int32_t scalar_sync;      /* FUSED: 0 = ncclAllReduce of the norm, 1 = mailbox barrier over peer memory (no NCCL call)  */
	int32_t reserved[3];
} b200spmv_power_opts;

int b200spmv_comm_unique_id(unsigned char id_out[128]);                  /* ncclGetUniqueId: call on ONE rank, hand to all */
int b200spmv_comm_create(b200spmv_comm_t ** out, const unsigned char id[128], int32_t rank, int32_t world, int32_t device);
void b200spmv_comm_destroy(b200spmv_comm_t * c);

/* shard: this rank's rows [bounds[2 rank], bounds[2 rank + 1]) of the n x n matrix as a handle with m = that many rows and
 * n columns (CSR-adaptive, optionally column panels; SELL shards run unchunked).  bounds: 2 * world row bounds tiling
 * [0, n).  Collective over the communicator.  The shard handle stays owned by the caller and must outlive the object. */
int b200spmv_power_create(b200spmv_power_t ** out, b200spmv_t * shard, b200spmv_comm_t * comm, const int64_t * bounds, int64_t n,
                          const b200spmv_power_opts * opts);
/* x0: n doubles (HOST), the same on every rank; the iterate is kept un-normalised with its scale 1 / ||x|| beside it */
int b200spmv_power_set_start(b200spmv_power_t * pw, const double * x0_host);
/* `iters` iterations u' = s A u, s' = 1 / ||u'||.  eigenvalue_out (optional): ||A x|| of the last iteration with ||x|| = 1;
 * ms_out (optional): device time of the iterations (CUDA events on the launching stream).  Collective. */
int b200spmv_power_iterate(b200spmv_power_t * pw, int32_t iters, double * eigenvalue_out, float * ms_out);
/* The current un-normalised iterate (HOST, n doubles; rows outside the exchanged band that this rank does not own are stale
 * by design: nothing reads them) and its scale -- for per-step parity checks against the serial reference kernel. */
int b200spmv_power_get_x(b200spmv_power_t * pw, double * x_host, double * scale_out);
/* info_out[0] = row chunks used, [1] = column panels of the shard, [2] = band lo, [3] = band hi (global rows exchanged),
 * [4] = 1 when a CUDA graph is active, [5] = kernel launches per iteration, [6] = NCCL calls per iteration */
int b200spmv_power_info(const b200spmv_power_t * pw, int64_t * info_out, int32_t n_out);
void b200spmv_power_destroy(b200spmv_power_t * pw);

/* The row-chunk x column-panel launch order used above, as a pure host function (testable without a GPU): tile_row[q] =
 * B200SPMV_X_TILE_ROW of panel q (ntiles[q] + 1 entries, last = m).  unit_out[q * (K + 1) + c] = first tile of panel q in chunk
 * c; final_row_out[c] = rows [0, final_row_out[c]) are final once chunks < c of every panel have run.  Invariant: the rows a
 * panel finishes in chunk c were finished by every earlier panel in chunks <= c. */
int b200spmv_chunk_schedule(int32_t P, int32_t K, int64_t m, const int32_t * const * tile_row, const int64_t * ntiles,
                            int64_t * unit_out, int64_t * final_row_out);

const char * b200spmv_last_error(void);
const char * b200spmv_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B200SPMV_H */

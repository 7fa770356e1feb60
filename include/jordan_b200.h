/* jordan_b200 — B200-native distributed dense matrix inversion (block Gauss-Jordan with block-row
 * pivoting).  Public C ABI of libjordan_host.so (CPU-only part) and libjordan_b200.so (CUDA part).
 *
 * Citations "main.cpp:N" point into the reference program yusupov1alik/MPI-Jordan-crazy-acceleration
 * whose capabilities this library reproduces.  All sizes are 64-bit (the reference overflows
 * `int len = RpP*m*n`, main.cpp:366). */
#ifndef JORDAN_B200_H
#define JORDAN_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ error codes */
#define JB_OK 0
#define JB_ERR_CUDA (-1)
#define JB_ERR_SINGULAR (-2) /* Jordan() return value, main.cpp:1082 */
#define JB_ERR_NOMEM (-3)    /* main.cpp:989 */
#define JB_ERR_ARG (-4)
#define JB_ERR_IO (-5)
#define JB_ERR_TIMEOUT (-6) /* a cross-GPU wait gave up: a peer died or fell out of step */
const char* jb_last_error(void);

/* ------------------------------------------------------------------ block-cyclic layout math
 * (libjordan_host) — rows_p_process main.cpp:95-116, local_to_global 118-123,
 * num_block_rows 124-127, find_sender 526-532. */
int64_t jb_num_block_rows(int64_t n, int64_t m);
int64_t jb_find_sender(int64_t Nr, int64_t p);
int64_t jb_rows_per_process(int64_t Nr, int64_t p, int64_t k);
int64_t jb_local_to_global_row(int64_t local_row, int64_t m, int64_t p, int64_t k);
int64_t jb_global_to_local_row(int64_t global_row, int64_t m, int64_t p);
int64_t jb_owner_of_row(int64_t global_row, int64_t m, int64_t p);
int64_t jb_padded_ld(int64_t n, int64_t m); /* leading dimension used by the engine */

/* ------------------------------------------------------------------ generators (f, f_i: main.cpp:47-64)
 * ids: 1 |i-j| (default), 2 Hilbert, 3 anti-diagonal, 4 Lehmer, 5 diagonally dominant,
 *      6 row-permuted |i-j| (exercises off-diagonal block pivots). */
double jb_formula(int id, int64_t i, int64_t j, int64_t n);

/* ------------------------------------------------------------------ text matrix I/O (libjordan_host)
 * read_matrix main.cpp:209-282: n*n whitespace separated numbers, row-major, "%lf".
 * Returns 0, -1 cannot open, -2 cannot read (short / garbled file). */
int jb_read_matrix_rows(const char* path, int64_t n, int64_t row0, int64_t nrows, double* out);
int jb_read_matrix_full(const char* path, int64_t n, double* out);
/* Streaming form (one scan of the file, `count` values per call, e.g. one block row = h*n): what
 * read_matrix does with its m*n buffer, main.cpp:249-263.  open: 0 / -1; next: 0 / -2. */
typedef struct JbMatrixReader JbMatrixReader;
int jb_matrix_reader_open(const char* path, JbMatrixReader** out);
int jb_matrix_reader_next(JbMatrixReader* r, int64_t count, double* out);
void jb_matrix_reader_close(JbMatrixReader* r);
/* print_row main.cpp:284-295: "%.2f\t" per value, "\n" per row, into a caller buffer. */
int64_t jb_format_corner(const double* a, int64_t rows, int64_t cols, int64_t ld, char* buf,
                         int64_t buflen);

/* ------------------------------------------------------------------ CPU oracle (libjordan_host)
 * Intended-semantics block Gauss-Jordan (SURVEY D1 fixed), emulating p ranks in one process:
 * same pivot rule (min ||block^-1||inf, local ties -> lowest row, cross-rank ties -> highest
 * rank), same singular test.  a: n*n row-major, overwritten; x: n*n inverse.  piv_out (optional):
 * Nr chosen global block rows.  Returns 0 or JB_ERR_SINGULAR. */
int jb_cpu_block_jordan(double* a, double* x, int64_t n, int64_t m, int64_t p, int32_t* piv_out);
/* inverse_block main.cpp:746-820 restated: returns 1 if singular. */
int jb_cpu_inverse_block(double* a, double* b, double norm, int64_t n);
/* pivot_op main.cpp:729-735 applied as MPI's rank-ordered fold; returns winning index or -1. */
int jb_cpu_minloc_fold(const double* norms, const int32_t* nonsing, int64_t p);
double jb_cpu_residual(const double* a, const double* x, int64_t n);

/* Work list of the fused step kernel for one phase (CPU-side view of csrc/step_plan.h, used by the
 * scheduling-invariant tests).  out[5*i..] = {type, column tile, first local row, rows, item
 * belongs to the NEXT pivot row}.  phase: 0 ALL, 1 A, 2 B, 3 A2, 4 S, 5 B1, 6 B2; nb1 = number of
 * non-priority column tiles of phase B1, the first `nearly` of them flagged early; next_piv_* =
 * pivot of step t+1 (rank < 0: none). */
int64_t jb_plan_enumerate(int32_t m, int32_t p, int32_t rank, int32_t rloc, int64_t ldw, int32_t t,
                          int32_t piv_rank, int32_t piv_slot, int32_t do_swap, int32_t phase,
                          int32_t pct0, int32_t pct1, int32_t nb1, int32_t nearly,
                          int32_t next_piv_rank, int32_t next_piv_slot, int32_t* out, int64_t cap);

/* ------------------------------------------------------------------ GPU engine (libjordan_b200) */
typedef struct JbEngine JbEngine;

typedef struct {
  int64_t n;       /* matrix order */
  int32_t m;       /* block size */
  int32_t p;       /* ranks = GPUs */
  int32_t rank;
  int32_t device;  /* CUDA device ordinal of this rank */
  int32_t keep_copy; /* keep a pristine copy of A (needed for file input residual / refinement) */
  int32_t precision; /* 0 = fp64 DMMA path, 1 = tcgen05 split-precision update + refinement */
} JbConfig;

int jb_engine_create(const JbConfig* cfg, JbEngine** out);
void jb_engine_destroy(JbEngine* e);

/* Symmetric (peer-visible) region: pivot-row buffers, swap buffer, tile flags, min-loc mailbox.
 * Either let the engine allocate it (jb_engine_alloc_symm + IPC handle exchange) or bind memory
 * obtained elsewhere (e.g. torch symmetric memory, which also yields an NVLS multicast alias). */
size_t jb_engine_symm_bytes(const JbEngine* e);
int jb_engine_alloc_symm(JbEngine* e);
int jb_engine_bind_symm(JbEngine* e, void* const* peer_ptrs, void* multicast_ptr);
/* which: 0 = symmetric region, 1 = matrix slab W, 2 = refinement residual slab.
 * handle: 64 bytes (cudaIpcMemHandle_t). */
int jb_engine_ipc_handle(JbEngine* e, int which, void* handle64);
int jb_engine_open_peers(JbEngine* e, int which, const void* handles /* p * 64 bytes */);
/* Native symmetric heap (csrc/comm/symm_heap.cpp; no counterpart in the reference, which keeps
 * everything in host memory and exchanges it with MPI_Bcast / MPI_Send, main.cpp:1097, 1122-1129):
 * one cuMemCreate allocation per rank, shared as POSIX file descriptors that the CALLER moves between
 * the rank processes (SCM_RIGHTS over Unix sockets: jb_mpi_send_fd / jb_mpi_recv_fd of the MPI shim,
 * or parallel/comm.py under torchrun), mapped on every rank, plus an NVLS multicast alias
 * (cuMulticastCreate / AddDevice / BindMem) where the GPUs sit behind an NVSwitch.  Order of calls:
 *   create (all)  ->  exchange fds  ->  import (all)
 *   [multicast]   mc_create (rank 0) -> send fd to all -> mc_import (others) -> mc_add (all)
 *                 -> barrier -> mc_bind (all) -> barrier;  on any failure: mc_drop on ALL ranks
 * then hand the pointers to jb_engine_bind_symm.  The heap must outlive the engine. */
typedef struct JbSymmHeap JbSymmHeap;
int jb_symm_heap_create(size_t bytes, int device, int rank, int size, int want_multicast,
                        JbSymmHeap** out, int* export_fd);
int jb_symm_heap_import(JbSymmHeap* h, const int* fds /* [size], closed by the call */);
int jb_symm_heap_multicast_supported(const JbSymmHeap* h);
int jb_symm_heap_mc_create(JbSymmHeap* h, int* export_fd);
int jb_symm_heap_mc_import(JbSymmHeap* h, int fd);
int jb_symm_heap_mc_add(JbSymmHeap* h);
int jb_symm_heap_mc_bind(JbSymmHeap* h);
void jb_symm_heap_mc_drop(JbSymmHeap* h);
void* jb_symm_heap_ptr(const JbSymmHeap* h, int rank);   /* rank's memory as mapped here */
void* jb_symm_heap_mc_ptr(const JbSymmHeap* h);          /* multicast alias or NULL */
size_t jb_symm_heap_bytes(const JbSymmHeap* h);
void jb_symm_heap_destroy(JbSymmHeap* h);

/* Zero flags / mailboxes / step counter.  Call on all ranks, then barrier, e.g. after a singular
 * verdict left the step epochs of the ranks inconsistent. */
int jb_engine_reset_sync(JbEngine* e);

/* matrix set-up */
int jb_engine_fill_formula(JbEngine* e, int formula, void* stream);
int jb_engine_upload_rows(JbEngine* e, const double* host_rows, int64_t local_row0, int64_t nrows,
                          void* stream); /* rows are n wide; padding is handled */
int jb_engine_finish_upload(JbEngine* e, void* stream); /* identity padding, norm, optional copy */
int jb_engine_local_norm(JbEngine* e, double* out);     /* local ||A||inf (synchronises) */
int jb_engine_set_norm(JbEngine* e, double global_norm); /* after a MAX reduction over the ranks */

/* the timed region: enqueue all Nr elimination steps + final column un-permutation */
int jb_engine_invert(JbEngine* e, void* stream);
/* synchronise and fetch the verdict: JB_OK or JB_ERR_SINGULAR */
int jb_engine_status(JbEngine* e, void* stream);
int jb_engine_pivots(JbEngine* e, int32_t* piv_out /* Nr */);
int64_t jb_engine_launch_count(const JbEngine* e); /* kernels launched by this engine so far */
/* per-kernel CUDA-event timing of the last jb_engine_invert (adds a stream sync; off by default) */
int jb_engine_set_profiling(JbEngine* e, int on);
/* seg_ms[4], kernel durations summed over the steps (the kernels of the look-ahead schedule
 * overlap, so they do not add up to the inversion time).  Look-ahead: [0] phase A (swap + block
 * columns t, t+1), [1] phase B1 (K1(t+1) runs beside it), [2] phase B2 (next pivot row scaled +
 * broadcast, rest of the update), [3] K1(t+1) including its wait for free SMs.  Serial schedule:
 * [0] fused step kernel, [2] K1(t+1).  tail_ms: final column un-permutation. */
int jb_engine_phase_times(JbEngine* e, double* seg_ms, double* tail_ms);

/* results */
void* jb_engine_w_ptr(JbEngine* e);
int64_t jb_engine_ldw(const JbEngine* e);
int64_t jb_engine_local_rows(const JbEngine* e);
int jb_engine_download(JbEngine* e, int which /*0 = X (W), 1 = kept copy of A*/, int64_t local_row0,
                       int64_t nrows, int64_t col0, int64_t ncols, double* host, int64_t ldh);
/* local max_i sum_j |(A X - I)_ij| over this rank's rows; formula < 0 -> use the kept copy. */
int jb_engine_residual(JbEngine* e, int formula, double* out_local, void* stream);
/* Newton-Schulz X <- X + X (I - A X) in fp64 (K7): phase 0 = residual slab, barrier, phase 1 =
 * update, barrier. */
int jb_engine_refine_phase(JbEngine* e, int formula, int phase, void* stream);

/* stand-alone kernels for tests / tools */
int jb_block_inverse(const double* d_blocks, int64_t nblocks, int32_t m, double norm_a,
                     double* d_inv, double* d_norms, int32_t* d_ok, void* stream);
/* Same operation as jb_block_inverse, for callers that must not allocate (all device memory
 * already handed to a pool, or no room for cudaMalloc's device-wide synchronisation).  No
 * counterpart in the reference.  Staging lives in a fixed ~4 MiB device array inside the library
 * (resident on every GPU that loads it): accepted only if nblocks * Kp * Kp <= 524288 doubles
 * (Kp = m rounded up to 16) and m <= 256, otherwise JB_ERR_ARG.  Calls no cudaMalloc/cudaFree.
 * Outputs: d_inv = nblocks packed m x m inverses, d_norms and d_ok = nblocks entries; a singular
 * block (d_ok = 0) comes back as zeros.  Asynchronous on `stream` (synchronise before reading the
 * outputs).  All calls on a device share the array: calls that overlap on the GPU race.  The first
 * call on each device resolves the array's address; make it during initialisation, before any
 * stream capture begins, and do not reset the device afterwards. */
int jb_block_inverse_noalloc(const double* d_blocks, int64_t nblocks, int32_t m, double norm_a,
                             double* d_inv, double* d_norms, int32_t* d_ok, void* stream);
/* Message-passing litmus between two peer GPUs of this process (payload stores, fence,
 * st.release.sys flag / ld.acquire.sys spin): counts ordering violations over `iters` epochs. */
int jb_flag_litmus(int dev_producer, int dev_consumer, int iters, unsigned long long* errors_out);
/* Store / load bandwidth of the paths the broadcast uses (tools/nvlink_bw.py): mode 0 = 16-byte
 * stores to dst (local or peer mapping), 1 = multimem.st to an NVLS multicast alias, 2 = 16-byte
 * loads.  ms_out = average milliseconds per pass over `bytes`.  No counterpart in the reference. */
int jb_store_bandwidth(void* dst, size_t bytes, int mode, int iters, float* ms_out);
/* C (rows x cols) += L (rows x K) * Rn (K x cols), fp64 operands in device memory, computed on the
 * tcgen05 int8-split path (K3b) with S slices; rows must be a multiple of 128 (test / tool entry). */
int jb_tc05_gemm_update(double* C, int64_t ldc, int64_t rows, int64_t cols, const double* L,
                        int64_t ldl, const double* Rn, int64_t ldr, int32_t K, int32_t S, void* stream);
#ifdef __cplusplus
}
#endif
#endif

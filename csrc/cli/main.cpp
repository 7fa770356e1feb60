// bin/jordan — native command-line driver with the reference's contract (main.cpp:65-93, 343-519):
//
//     [jb_mpirun -np p] jordan n m [<file>]            exactly the reference's form
//     [jb_mpirun -np p] jordan n m r s [<file>]        extended: print size r, formula id s (0 = file)
//     flags --print r / --formula s are accepted anywhere
//
// stdout of rank 0, in this order: "A", the top-left min(n,r)^2 corner ("%.2f\t"), "glob_time: %.2f",
// "inverse matrix:", blank line, the corner of the inverse, "residual: %e".  Exit codes: 0 ok,
// 1 usage, 2 any failure ("cannot open %s", "cannot read %s", "singular matrix",
// "Not enough memory!").  One rank per GPU; ranks are bootstrapped through the MPI API (in-tree shim
// or a real MPI), which is also used for the few control-plane collectives the reference performs
// outside its hot loop (C2, C3, C4, C5, C6, C14 in SURVEY 2.3).  Everything per-step is done by the
// engine's kernels over peer memory.
#include <cuda_runtime.h>
#include <mpi.h>

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include <string>
#include <vector>

#include "jordan_b200.h"

static int g_rank = 0, g_size = 1;

static void print_corner(JbEngine* e, int which, int64_t n, int m, int nm) {
  // print_matrix (main.cpp:297-341): owners send their rows of the corner to rank 0 (C5)
  std::vector<double> row((size_t)nm);
  std::vector<char> text((size_t)nm * 32 + 16);
  for (int64_t gi = 0; gi < nm; ++gi) {
    const int owner = (int)jb_owner_of_row(gi, m, g_size);
    if (g_rank == owner) {
      if (jb_engine_download(e, which, jb_global_to_local_row(gi, m, g_size), 1, 0, nm, row.data(), nm) != JB_OK) {
        fprintf(stderr, "download failed: %s\n", jb_last_error());
        MPI_Abort(MPI_COMM_WORLD, 2);
      }
      if (owner != 0) MPI_Send(row.data(), nm, MPI_DOUBLE, 0, 0, MPI_COMM_WORLD);
    }
    if (g_rank == 0) {
      if (owner != 0) MPI_Recv(row.data(), nm, MPI_DOUBLE, owner, 0, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
      if (jb_format_corner(row.data(), 1, nm, nm, text.data(), (int64_t)text.size()) >= 0)
        fputs(text.data(), stdout);
    }
  }
  (void)n;
}

// Symmetric region of the engine.  Default: the native VMM heap (cuMemCreate + POSIX descriptors
// passed over the MPI shim's sockets + NVLS multicast object, csrc/comm/symm_heap.cpp).  JB_SYMM=ipc
// (or a build against an MPI without a descriptor channel) selects cudaMalloc + cudaIpc handles,
// unicast only.  Every decision is voted on so that all ranks take the same path.
static JbSymmHeap* g_heap = nullptr;
static const char* g_symm_mode = "local";

static int vote_min(int v) {
  int out = v;
  MPI_Allreduce(&v, &out, 1, MPI_INT, MPI_MIN, MPI_COMM_WORLD);
  return out;
}

static int setup_symm_ipc(JbEngine* e) {
  int rc = jb_engine_alloc_symm(e);
  if (vote_min(rc) != JB_OK) return rc ? rc : JB_ERR_NOMEM;
  char mine[64];
  std::vector<char> all((size_t)64 * g_size);
  if ((rc = jb_engine_ipc_handle(e, 0, mine)) != JB_OK) return rc;
  MPI_Allgather(mine, 64, MPI_BYTE, all.data(), 64, MPI_BYTE, MPI_COMM_WORLD);
  if ((rc = jb_engine_open_peers(e, 0, all.data())) != JB_OK) return rc;
  g_symm_mode = "cuda-ipc";
  return JB_OK;
}

static int setup_symm(JbEngine* e, int device) {
  if (g_size == 1) return jb_engine_alloc_symm(e);
  const char* mode = getenv("JB_SYMM");
#ifndef JB_MPI_SHIM_HAS_FD_PASSING
  mode = "ipc";
#endif
  if (mode && !strcmp(mode, "ipc")) return setup_symm_ipc(e);
#ifdef JB_MPI_SHIM_HAS_FD_PASSING
  int fd = -1;
  int rc = jb_symm_heap_create(jb_engine_symm_bytes(e), device, g_rank, g_size, getenv("JB_NO_MULTICAST") ? 0 : 1,
                               &g_heap, &fd);
  if (vote_min(rc) != JB_OK) {  // no VMM / descriptors on some rank: every rank falls back together
    if (g_heap) jb_symm_heap_destroy(g_heap);
    g_heap = nullptr;
    if (fd >= 0) close(fd);
    if (g_rank == 0 && getenv("JB_VERBOSE")) printf("symmetric heap: VMM path unavailable, using cudaIpc\n");
    return setup_symm_ipc(e);
  }
  std::vector<int> fds((size_t)g_size, -1);
  for (int r = 0; r < g_size; ++r)
    if (r != g_rank) jb_mpi_send_fd(fd, r);
  for (int r = 0; r < g_size; ++r)
    if (r != g_rank) fds[r] = jb_mpi_recv_fd(r);
  close(fd);
  rc = jb_symm_heap_import(g_heap, fds.data());
  if (vote_min(rc) != JB_OK) return rc ? rc : JB_ERR_CUDA;
  // NVLS multicast alias: all ranks must succeed at every stage, else everybody drops it
  int mc = vote_min(jb_symm_heap_multicast_supported(g_heap));
  if (mc) {
    int ok = 1;
    if (g_rank == 0) {
      int mfd = -1;
      ok = jb_symm_heap_mc_create(g_heap, &mfd) == JB_OK;
      for (int r = 1; r < g_size; ++r) {
        MPI_Send(&ok, 1, MPI_INT, r, 3, MPI_COMM_WORLD);
        if (ok) jb_mpi_send_fd(mfd, r);
      }
      if (mfd >= 0) close(mfd);
    } else {
      MPI_Recv(&ok, 1, MPI_INT, 0, 3, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
      if (ok) ok = jb_symm_heap_mc_import(g_heap, jb_mpi_recv_fd(0)) == JB_OK;
    }
    ok = vote_min(ok);
    if (ok) ok = vote_min(jb_symm_heap_mc_add(g_heap) == JB_OK);     // vote = barrier: all devices added
    if (ok) ok = vote_min(jb_symm_heap_mc_bind(g_heap) == JB_OK);    // vote = barrier: all bound
    if (!ok) {
      if (getenv("JB_VERBOSE")) fprintf(stderr, "[rank %d] multicast set-up failed: %s\n", g_rank, jb_last_error());
      jb_symm_heap_mc_drop(g_heap);
    }
    mc = ok;
  }
  std::vector<void*> ptrs((size_t)g_size);
  for (int r = 0; r < g_size; ++r) ptrs[r] = jb_symm_heap_ptr(g_heap, r);
  rc = jb_engine_bind_symm(e, ptrs.data(), mc ? jb_symm_heap_mc_ptr(g_heap) : nullptr);
  g_symm_mode = mc ? "vmm+nvls" : "vmm";
  return rc;
#else
  (void)device;
  return JB_ERR_ARG;
#endif
}

static int solve(int64_t n, int m, int print_size, int formula, const char* name) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    if (g_rank == 0) printf("no CUDA device\n");
    return -1;
  }
  // the reference accepts any m, including m > n (one block row of height n, main.cpp:124-127)
  if ((int64_t)m > n) m = (int)n;
  JbConfig cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.n = n; cfg.m = m; cfg.p = g_size; cfg.rank = g_rank; cfg.device = g_rank % ndev;
  cfg.keep_copy = name ? 1 : 0;
  if (const char* pr = getenv("JB_PRECISION")) cfg.precision = !strcmp(pr, "tc05") ? 1 : 0;
  JbEngine* e = nullptr;
  int rc_create = jb_engine_create(&cfg, &e);
  std::string create_msg = rc_create != JB_OK ? jb_last_error() : "";
  int worst = vote_min(rc_create);  // C2, main.cpp:371: every rank learns the most severe code
  if (worst == JB_OK) {
    rc_create = setup_symm(e, cfg.device);
    if (rc_create != JB_OK) create_msg = jb_last_error();
    worst = vote_min(rc_create);
  }
  if (worst != JB_OK) {
    if (g_rank == 0) {
      if (worst == JB_ERR_NOMEM) printf("Not enough memory!\n");  // main.cpp:375
      else if (worst == JB_ERR_ARG) printf("unsupported configuration: %s\n", create_msg.empty() ? "rejected by another rank" : create_msg.c_str());
      else printf("cannot set up the GPUs: %s\n", create_msg.empty() ? "failure on another rank" : create_msg.c_str());
    }
    if (rc_create != JB_OK && getenv("JB_VERBOSE")) fprintf(stderr, "[rank %d] %s\n", g_rank, create_msg.c_str());
    if (e) jb_engine_destroy(e);
    if (g_heap) jb_symm_heap_destroy(g_heap);
    return -1;
  }
  if (g_size > 1) {  // matrix slabs: peer-readable for the residual's direct X pulls (C13)
    char mine[64];
    std::vector<char> all((size_t)64 * g_size);
    if (jb_engine_ipc_handle(e, 1, mine) != JB_OK) { fprintf(stderr, "%s\n", jb_last_error()); MPI_Abort(MPI_COMM_WORLD, 2); }
    MPI_Allgather(mine, 64, MPI_BYTE, all.data(), 64, MPI_BYTE, MPI_COMM_WORLD);
    if (jb_engine_open_peers(e, 1, all.data()) != JB_OK) { fprintf(stderr, "%s\n", jb_last_error()); MPI_Abort(MPI_COMM_WORLD, 2); }
    MPI_Barrier(MPI_COMM_WORLD);
  }
  if (g_rank == 0 && getenv("JB_VERBOSE")) printf("symmetric region: %s\n", g_symm_mode);

  // ---- matrix: text file scattered block row by block row (read_matrix, C3/C4) or generator.
  //      Rank 0 scans the file once and holds ONE block row at a time (main.cpp:249-263).
  if (name) {
    int err = 0;
    JbMatrixReader* rd = nullptr;
    if (g_rank == 0) err = jb_matrix_reader_open(name, &rd);
    MPI_Bcast(&err, 1, MPI_INT, 0, MPI_COMM_WORLD);  // main.cpp:237
    const int64_t Nr = jb_num_block_rows(n, m);
    std::vector<double> buf((size_t)m * n);
    for (int64_t I = 0; I < Nr && !err; ++I) {
      const int owner = (int)(I % g_size);
      const int64_t h = (I == Nr - 1) ? n - I * m : m;
      if (g_rank == 0) err = jb_matrix_reader_next(rd, h * n, buf.data());
      MPI_Bcast(&err, 1, MPI_INT, 0, MPI_COMM_WORLD);  // fail fast, main.cpp:277
      if (err) break;
      if (g_rank == 0 && owner != 0) MPI_Send(buf.data(), (int)(h * n), MPI_DOUBLE, owner, 0, MPI_COMM_WORLD);
      if (g_rank == owner) {
        if (owner != 0) MPI_Recv(buf.data(), (int)(h * n), MPI_DOUBLE, 0, 0, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
        if (jb_engine_upload_rows(e, buf.data(), (I / g_size) * m, h, nullptr) != JB_OK) { fprintf(stderr, "%s\n", jb_last_error()); MPI_Abort(MPI_COMM_WORLD, 2); }
        cudaStreamSynchronize(nullptr);  // buf is reused for the next block row
      }
    }
    if (rd) jb_matrix_reader_close(rd);
    if (err) {
      if (g_rank == 0) {
        if (err == -1) printf("cannot open %s\n", name);
        else if (err == -2) printf("cannot read %s\n", name);
        else printf("unknown error %d in %s\n", err, name);
      }
      jb_engine_destroy(e);
      if (g_heap) jb_symm_heap_destroy(g_heap);
      return -2;
    }
    if (jb_engine_finish_upload(e, nullptr) != JB_OK) { fprintf(stderr, "%s\n", jb_last_error()); MPI_Abort(MPI_COMM_WORLD, 2); }
  } else {
    if (jb_engine_fill_formula(e, formula, nullptr) != JB_OK) { fprintf(stderr, "%s\n", jb_last_error()); MPI_Abort(MPI_COMM_WORLD, 2); }
  }
  double loc_norm = 0, glob_norm = 0;
  jb_engine_local_norm(e, &loc_norm);
  MPI_Allreduce(&loc_norm, &glob_norm, 1, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD);  // global ||A||inf (D3)
  jb_engine_set_norm(e, glob_norm);

  const int nm = (int)(n < print_size ? n : print_size);
  if (g_rank == 0) printf("A\n");
  print_corner(e, name ? 1 : 0, n, m, nm);

  // ---- the timed region: the elimination only (main.cpp:427-450), device-timed, max over ranks
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaDeviceSynchronize();
  MPI_Barrier(MPI_COMM_WORLD);
  cudaEventRecord(e0, nullptr);
  int rc = jb_engine_invert(e, nullptr);
  cudaEventRecord(e1, nullptr);
  if (rc == JB_OK) rc = jb_engine_status(e, nullptr);
  int rc_min = 0;
  MPI_Allreduce(&rc, &rc_min, 1, MPI_INT, MPI_MIN, MPI_COMM_WORLD);
  if (rc_min != JB_OK) {
    if (g_rank == 0) {
      if (rc_min == JB_ERR_NOMEM) printf("not enough memory for block\n");
      else if (rc_min == JB_ERR_SINGULAR) printf("singular matrix\n");
      else if (rc_min == JB_ERR_TIMEOUT) printf("a rank stopped responding (cross-GPU wait timed out)\n");
      else printf("unknown error: %d\n", rc_min);
    }
    jb_engine_destroy(e);
    if (g_heap) jb_symm_heap_destroy(g_heap);
    return -1;
  }
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  double loc_t = ms * 1e-3, glob_t = 0;
  MPI_Allreduce(&loc_t, &glob_t, 1, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD);  // C6
  if (g_rank == 0) {
    printf("glob_time: %.2f\n", glob_t);
    printf("inverse matrix:\n\n");
  }
  print_corner(e, 0, n, m, nm);

  // ---- residual with the original A (regenerated from the formula / the kept copy); always
  //      computed, also for p == 1
  MPI_Barrier(MPI_COMM_WORLD);
  double loc_res = 0, glob_res = 0;
  if (jb_engine_residual(e, name ? -1 : formula, &loc_res, nullptr) != JB_OK) {
    fprintf(stderr, "[rank %d] residual failed: %s\n", g_rank, jb_last_error());
    MPI_Abort(MPI_COMM_WORLD, 2);
  }
  MPI_Allreduce(&loc_res, &glob_res, 1, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD);  // C14
  if (g_rank == 0) {
    printf("residual: %e\n", glob_res);
    if (getenv("JB_VERBOSE"))
      printf("seconds: %.6f  gflops(2n^3): %.1f  launches: %lld\n", glob_t, 2.0 * n * n * n / glob_t / 1e9,
             (long long)jb_engine_launch_count(e));
  }
  MPI_Barrier(MPI_COMM_WORLD);
  jb_engine_destroy(e);
  if (g_heap) jb_symm_heap_destroy(g_heap);
  return 0;
}

int main(int argc, char* argv[]) {
  MPI_Init(&argc, &argv);
  MPI_Comm_size(MPI_COMM_WORLD, &g_size);
  MPI_Comm_rank(MPI_COMM_WORLD, &g_rank);
  int print_size = 10, formula = 1;
  std::vector<const char*> pos;
  bool bad = false;
  for (int i = 1; i < argc; ++i) {
    if (!strcmp(argv[i], "--print") && i + 1 < argc) print_size = atoi(argv[++i]);
    else if (!strcmp(argv[i], "--formula") && i + 1 < argc) formula = atoi(argv[++i]);
    else if (!strncmp(argv[i], "--", 2)) bad = true;
    else pos.push_back(argv[i]);
  }
  long long n = 0, m = 0;
  const char* name = nullptr;
  if (pos.size() >= 2) {
    n = atoll(pos[0]);
    m = atoll(pos[1]);
  }
  if (pos.size() == 3) {
    name = pos[2];
  } else if (pos.size() == 4 || pos.size() == 5) {  // n m r s [file]
    print_size = atoi(pos[2]);
    formula = atoi(pos[3]);
    if (pos.size() == 5) name = pos[4];
  }
  if (bad || pos.size() < 2 || pos.size() > 5 || n <= 0 || m <= 0 || print_size <= 0 ||
      (formula == 0 && !name) || formula < 0) {
    if (g_rank == 0) printf("usage:%s n m [<file>]\n", argv[0]);  // main.cpp:80
    MPI_Finalize();
    return 1;
  }
  int ret = 0;
  if (solve(n, (int)m, print_size, formula, name)) ret = 2;
  MPI_Finalize();
  return ret;
}

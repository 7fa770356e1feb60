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

static int solve(int64_t n, int m, int print_size, int formula, const char* name) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    if (g_rank == 0) printf("no CUDA device\n");
    return -1;
  }
  JbConfig cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.n = n; cfg.m = m; cfg.p = g_size; cfg.rank = g_rank; cfg.device = g_rank % ndev;
  cfg.keep_copy = name ? 1 : 0;
  JbEngine* e = nullptr;
  int loc_err = jb_engine_create(&cfg, &e) != JB_OK ? 1 : 0, glob_err = 0;
  if (!loc_err) loc_err = jb_engine_alloc_symm(e) != JB_OK ? 1 : 0;
  MPI_Allreduce(&loc_err, &glob_err, 1, MPI_INT, MPI_SUM, MPI_COMM_WORLD);  // C2, main.cpp:371
  if (glob_err) {
    if (g_rank == 0) printf("Not enough memory!\n");
    if (loc_err && getenv("JB_VERBOSE")) fprintf(stderr, "[rank %d] %s\n", g_rank, jb_last_error());
    if (e) jb_engine_destroy(e);
    return -1;
  }
  if (g_size > 1) {
    for (int which = 0; which < 2; ++which) {
      char mine[64];
      std::vector<char> all((size_t)64 * g_size);
      if (jb_engine_ipc_handle(e, which, mine) != JB_OK) { fprintf(stderr, "%s\n", jb_last_error()); MPI_Abort(MPI_COMM_WORLD, 2); }
      MPI_Allgather(mine, 64, MPI_BYTE, all.data(), 64, MPI_BYTE, MPI_COMM_WORLD);
      if (jb_engine_open_peers(e, which, all.data()) != JB_OK) { fprintf(stderr, "%s\n", jb_last_error()); MPI_Abort(MPI_COMM_WORLD, 2); }
    }
    MPI_Barrier(MPI_COMM_WORLD);
  }

  // ---- matrix: text file scattered block row by block row (read_matrix, C3/C4) or generator
  if (name) {
    int err = 0;
    std::vector<double> full;
    if (g_rank == 0) {
      full.resize((size_t)n * n);
      err = jb_read_matrix_full(name, n, full.data());
    }
    MPI_Bcast(&err, 1, MPI_INT, 0, MPI_COMM_WORLD);
    if (err) {
      if (g_rank == 0) {
        if (err == -1) printf("cannot open %s\n", name);
        else if (err == -2) printf("cannot read %s\n", name);
        else printf("unknown error %d in %s\n", err, name);
      }
      jb_engine_destroy(e);
      return -2;
    }
    const int64_t Nr = jb_num_block_rows(n, m);
    std::vector<double> buf((size_t)m * n);
    for (int64_t I = 0; I < Nr; ++I) {
      const int owner = (int)(I % g_size);
      const int64_t h = (I == Nr - 1) ? n - I * m : m;
      const double* src = nullptr;
      if (g_rank == 0) {
        src = full.data() + (size_t)I * m * n;
        if (owner != 0) MPI_Send((void*)src, (int)(h * n), MPI_DOUBLE, owner, 0, MPI_COMM_WORLD);
      }
      if (g_rank == owner) {
        if (owner != 0) {
          MPI_Recv(buf.data(), (int)(h * n), MPI_DOUBLE, 0, 0, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
          src = buf.data();
        }
        if (jb_engine_upload_rows(e, src, (I / g_size) * m, h, nullptr) != JB_OK) { fprintf(stderr, "%s\n", jb_last_error()); MPI_Abort(MPI_COMM_WORLD, 2); }
        cudaStreamSynchronize(nullptr);
      }
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
      else printf("unknown error: %d\n", rc_min);
    }
    jb_engine_destroy(e);
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

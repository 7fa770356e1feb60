"""The in-tree MPI shim: semantics the reference (and bin/jordan) rely on — rank-ordered fold of a
non-commutative user op, derived-type extent, simultaneous large Sendrecv_replace, Bcast/Allgather."""
import os
import subprocess
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

PROG = textwrap.dedent(r'''
    #include "mpi.h"
    #include <cstdlib>
    #include <vector>
    struct Piv { double min; int k; int row; int ok; };
    static void op(void* a, void* b, int* len, MPI_Datatype*) {   // same rule as pivot_op (main.cpp:729-735)
      Piv* x = (Piv*)a; Piv* y = (Piv*)b;
      for (int i = 0; i < *len; ++i) if (x[i].ok && (!y[i].ok || x[i].min < y[i].min)) y[i] = x[i];
    }
    int main(int argc, char** argv) {
      MPI_Init(&argc, &argv);
      int p, k; MPI_Comm_size(MPI_COMM_WORLD, &p); MPI_Comm_rank(MPI_COMM_WORLD, &k);
      int bad = 0;
      // derived type + non-commutative op: all ranks tie -> the HIGHEST rank must win
      Piv piv = {1.0, k, 10 + k, 1}, res = {0, 0, 0, 0};
      MPI_Aint base, a; MPI_Aint d[4]; int c[4] = {1, 1, 1, 1};
      MPI_Datatype t[4] = {MPI_DOUBLE, MPI_INT, MPI_INT, MPI_INT}, T; MPI_Op O;
      MPI_Address(&piv, &base);
      MPI_Address(&piv.min, &a); d[0] = a - base; MPI_Address(&piv.k, &a); d[1] = a - base;
      MPI_Address(&piv.row, &a); d[2] = a - base; MPI_Address(&piv.ok, &a); d[3] = a - base;
      MPI_Type_struct(4, c, d, t, &T); MPI_Type_commit(&T); MPI_Op_create(op, 0, &O);
      MPI_Allreduce(&piv, &res, 1, T, O, MPI_COMM_WORLD);
      if (res.k != p - 1 || res.row != 10 + p - 1) bad |= 1;
      piv.min = (k == 1 % p) ? 0.5 : 1.0;                       // a strictly better candidate wins
      MPI_Allreduce(&piv, &res, 1, T, O, MPI_COMM_WORLD);
      if (res.k != 1 % p) bad |= 2;
      piv.ok = (k == 0);                                          // non-singular beats singular
      piv.min = 5.0;
      MPI_Allreduce(&piv, &res, 1, T, O, MPI_COMM_WORLD);
      if (res.k != 0 || !res.ok) bad |= 4;
      MPI_Type_free(&T); MPI_Op_free(&O);
      // simultaneous large ring shift (main.cpp:639): everybody sends 8 MB at once
      const int n = 1 << 20; std::vector<double> buf(n, (double)k);
      MPI_Status st;
      MPI_Sendrecv_replace(buf.data(), n, MPI_DOUBLE, (k + p - 1) % p, 0, (k + 1) % p, 0, MPI_COMM_WORLD, &st);
      if (buf[0] != (double)((k + 1) % p) || buf[n - 1] != buf[0]) bad |= 8;
      // built-in reductions, broadcast, allgather
      int one = k + 1, sum = 0; MPI_Allreduce(&one, &sum, 1, MPI_INT, MPI_SUM, MPI_COMM_WORLD);
      if (sum != p * (p + 1) / 2) bad |= 16;
      double x = k, mx = -1; MPI_Allreduce(&x, &mx, 1, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD);
      if (mx != p - 1) bad |= 32;
      int root_val = (k == p - 1) ? 77 : 0; MPI_Bcast(&root_val, 1, MPI_INT, p - 1, MPI_COMM_WORLD);
      if (root_val != 77) bad |= 64;
      std::vector<int> all(p); MPI_Allgather(&k, 1, MPI_INT, all.data(), 1, MPI_INT, MPI_COMM_WORLD);
      for (int i = 0; i < p; ++i) if (all[i] != i) bad |= 128;
      int tot = 0; MPI_Allreduce(&bad, &tot, 1, MPI_INT, MPI_SUM, MPI_COMM_WORLD);
      if (k == 0) printf("shim-check p=%d bad=%d\n", p, tot);
      MPI_Finalize();
      return tot ? 3 : 0;
    }
''')


@pytest.mark.parametrize("p", [1, 2, 3, 8])
def test_shim_semantics(p, tmp_path):
    from mpi_jordan_crazy_acceleration_b200 import _build

    _build.build_shim()
    src = tmp_path / "t.cpp"
    src.write_text(PROG)
    exe = tmp_path / "t.out"
    lib = os.path.join(ROOT, "mpi_jordan_crazy_acceleration_b200", "lib", "libjbmpi.a")
    subprocess.run(["g++", "-O1", "-I", os.path.join(ROOT, "mpi_shim"), str(src), lib, "-o", str(exe)], check=True)
    r = subprocess.run([os.path.join(ROOT, "bin", "jb_mpirun"), "-np", str(p), str(exe)], capture_output=True,
                       text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert f"shim-check p={p} bad=0" in r.stdout


def test_launcher_propagates_exit_codes(tmp_path):
    src = tmp_path / "e.cpp"
    src.write_text('#include "mpi.h"\nint main(int c,char**v){MPI_Init(&c,&v);int k;MPI_Comm_rank(MPI_COMM_WORLD,&k);MPI_Finalize();return k==1?2:0;}\n')
    exe = tmp_path / "e.out"
    lib = os.path.join(ROOT, "mpi_jordan_crazy_acceleration_b200", "lib", "libjbmpi.a")
    subprocess.run(["g++", "-I", os.path.join(ROOT, "mpi_shim"), str(src), lib, "-o", str(exe)], check=True)
    r = subprocess.run([os.path.join(ROOT, "bin", "jb_mpirun"), "-np", "3", str(exe)])
    assert r.returncode == 2


FD_PROG = r'''
#include "mpi.h"
#include <stdlib.h>
#include <string.h>
#include <unistd.h>
#include <vector>
// every rank writes its number into an unlinked temp file and hands the DESCRIPTOR to every other
// rank (the channel csrc/comm/symm_heap.cpp uses for CUDA VMM handles); interleaved with ordinary
// messages to check that descriptors stay in order with the byte stream
int main(int argc, char** argv) {
  MPI_Init(&argc, &argv);
  int p, k; MPI_Comm_size(MPI_COMM_WORLD, &p); MPI_Comm_rank(MPI_COMM_WORLD, &k);
  char name[] = "/tmp/jbfdXXXXXX";
  int fd = mkstemp(name); unlink(name);
  int payload = 1000 + k;
  if (write(fd, &payload, sizeof(payload)) != (ssize_t)sizeof(payload)) return 9;
  int bad = 0;
  std::vector<double> big(1 << 18, (double)k);
  for (int r = 0; r < p; ++r) if (r != k) {
    MPI_Send(&k, 1, MPI_INT, r, 5, MPI_COMM_WORLD);
    if (jb_mpi_send_fd(fd, r) != 0) bad |= 1;
    MPI_Send(big.data(), (int)big.size(), MPI_DOUBLE, r, 6, MPI_COMM_WORLD);
  }
  for (int r = 0; r < p; ++r) if (r != k) {
    int who = -1; MPI_Recv(&who, 1, MPI_INT, r, 5, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
    int got = jb_mpi_recv_fd(r);
    std::vector<double> in(big.size());
    MPI_Recv(in.data(), (int)in.size(), MPI_DOUBLE, r, 6, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
    int val = -1;
    if (who != r || got < 0 || pread(got, &val, sizeof(val), 0) != (ssize_t)sizeof(val) || val != 1000 + r) bad |= 2;
    if (in[0] != (double)r || in.back() != (double)r) bad |= 4;
    if (got >= 0) close(got);
  }
  int tot = 0; MPI_Allreduce(&bad, &tot, 1, MPI_INT, MPI_SUM, MPI_COMM_WORLD);
  if (k == 0) printf("fd-check p=%d bad=%d\n", p, tot);
  MPI_Finalize();
  return tot ? 3 : 0;
}
'''


@pytest.mark.parametrize("p", [2, 4])
def test_shim_passes_file_descriptors(p, tmp_path):
    from mpi_jordan_crazy_acceleration_b200 import _build

    _build.build_shim()
    src = tmp_path / "fd.cpp"
    src.write_text(FD_PROG)
    exe = tmp_path / "fd.out"
    lib = os.path.join(ROOT, "mpi_jordan_crazy_acceleration_b200", "lib", "libjbmpi.a")
    subprocess.run(["g++", "-O1", "-I", os.path.join(ROOT, "mpi_shim"), str(src), lib, "-o", str(exe)], check=True)
    r = subprocess.run([os.path.join(ROOT, "bin", "jb_mpirun"), "-np", str(p), str(exe)], capture_output=True,
                       text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert f"fd-check p={p} bad=0" in r.stdout

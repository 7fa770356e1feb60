"""Invert the reference's default generator matrix |i-j| on however many GPUs torchrun gives us.

    python examples/invert_formula.py 8192 128
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 examples/invert_formula.py 32768 128
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from mpi_jordan_crazy_acceleration_b200 import Comm, JordanSolver  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    m = int(sys.argv[2]) if len(sys.argv) > 2 else 128
    precision = sys.argv[3] if len(sys.argv) > 3 else "fp64"      # or "tc05"
    comm = Comm.init_from_env()
    s = JordanSolver(n, m, comm=comm, precision=precision)
    s.fill_formula("absdiff")
    seconds = s.invert()
    residual = s.residual()
    if comm.rank == 0:
        print(f"n={n} m={m} ranks={comm.world} path={precision} symm={s.symm_mode}: {seconds:.4f} s, "
              f"{2 * n ** 3 / seconds / 1e12:.1f} TFLOP/s (2n^3), residual {residual:.3e}")
    s.close()


if __name__ == "__main__":
    main()

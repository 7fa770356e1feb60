"""mpi_jordan_crazy_acceleration_b200 — B200-native distributed dense matrix inversion
(block Gauss-Jordan with block-row pivoting), with the capabilities of the MPI/CPU program
``yusupov1alik/MPI-Jordan-crazy-acceleration``.

Layers:  ``models`` (JordanSolver = fused sm_100a engine; EagerJordan = NCCL+cuBLAS baseline / CPU),
``ops`` (single kernels + CPU oracle), ``parallel`` (block-cyclic layout, process-group plumbing),
``utils`` (text I/O, report, timing), ``cli`` (the reference's ``n m [file]`` command line).
"""
__version__ = "0.1.0"

from . import _native  # noqa: F401
from .parallel import Comm, Layout  # noqa: F401
from .models import FORMULAS, EagerJordan, JordanSolver, invert  # noqa: F401

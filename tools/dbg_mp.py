"""torchrun debug worker: residual of one inversion. usage: torchrun ... tools/dbg_mp.py n m fid [symm]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_jordan_crazy_acceleration_b200.models.jordan import JordanSolver
from mpi_jordan_crazy_acceleration_b200.parallel.comm import Comm
comm = Comm.init_from_env()
dev = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(dev)
n, m, fid = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
mode = sys.argv[4] if len(sys.argv) > 4 else "auto"
s = JordanSolver(n, m, comm=comm, device=dev, symm=mode)
out = []
for rep in range(3):
    s.fill_formula(fid)
    s.invert()
    out.append(s.residual())
if comm.rank == 0:
    print("DBG", n, m, fid, s.symm_mode, {k: os.environ.get(k) for k in ("JB_SCHED", "JB_NO_LOOKAHEAD", "JB_B2_SERIAL")}, out, flush=True)
s.close()

"""torchrun worker: per-phase times on every rank (JB_TIME_KERNELS semantics). usage: ... n m"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_jordan_crazy_acceleration_b200.models.jordan import JordanSolver
from mpi_jordan_crazy_acceleration_b200.parallel.comm import Comm
comm = Comm.init_from_env()
dev = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(dev)
n, m = int(sys.argv[1]), int(sys.argv[2])
s = JordanSolver(n, m, comm=comm, device=dev)
s.set_profiling(True)
for _ in range(2):
    s.fill_formula(1)
    sec = s.invert()
pt = s.phase_times()
Nr = s.layout.Nr
rec = {"rank": comm.rank, "sec": sec, **{k: (v * 1e3 / Nr if k.endswith("_ms") else v) for k, v in pt.items()}}
allr = comm.allgather_obj(rec)
if comm.rank == 0:
    print("per-step us (keys *_ms are us/step here):")
    for r in allr:
        print(json.dumps(r))
s.close()

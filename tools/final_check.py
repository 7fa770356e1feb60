"""torchrun worker: headline config (n=32768, m=128) and config 3 (n=65536, m=256) on the default
schedule, device-timed inversion + residual each, in one process group.  usage: torchrun ...
tools/final_check.py [n m ...]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_jordan_crazy_acceleration_b200.models.jordan import JordanSolver
from mpi_jordan_crazy_acceleration_b200.parallel.comm import Comm
comm = Comm.init_from_env()
dev = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(dev)
args = [int(x) for x in sys.argv[1:]] or [32768, 128, 65536, 256]
for n, m in zip(args[0::2], args[1::2]):
    s = JordanSolver(n, m, comm=comm, device=dev)
    secs = []
    for _ in range(2):
        s.fill_formula(1)
        secs.append(s.invert())
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    res = s.residual()
    e1.record()
    torch.cuda.synchronize()
    rsec = comm.allreduce_max(e0.elapsed_time(e1) * 1e-3)
    if comm.rank == 0:
        print("FINAL", json.dumps({"n": n, "m": m, "n_gpus": comm.world, "seconds": min(secs), "all_seconds": secs,
                                   "gflops": 2.0 * n ** 3 / min(secs) / 1e9, "residual": res,
                                   "residual_phase_seconds": rsec,
                                   "residual_phase_gflops": 2.0 * n ** 3 / rsec / 1e9, "symm": s.symm_mode}), flush=True)
    s.close()

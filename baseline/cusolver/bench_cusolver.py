"""Single-GPU library reference point: LU factorisation + solve against I (cuSOLVER getrf/getrs via
torch.linalg), i.e. what one would call without this engine.  cuSOLVER has no dense getri.
Reported only as a reference point (BASELINE.json north_star), never on a product path.
usage: python baseline/cusolver/bench_cusolver.py [n] [reps]"""
import json
import sys

import torch


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    dev = torch.device("cuda:0")
    i = torch.arange(n, dtype=torch.float64, device=dev)
    A = (i[:, None] - i[None, :]).abs()
    eye = torch.eye(n, dtype=torch.float64, device=dev)
    best = 1e30
    X = None
    for _ in range(reps + 1):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        LU, piv = torch.linalg.lu_factor(A)
        X = torch.linalg.lu_solve(LU, piv, eye)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e-3)
    res = float((A @ X - eye).abs().sum(dim=1).max())
    print(json.dumps({"impl": "cusolver getrf+getrs(I)", "n": n, "seconds": best,
                      "gflops_2n3": 2.0 * n ** 3 / best / 1e9, "residual": res}))


if __name__ == "__main__":
    main()

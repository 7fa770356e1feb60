"""Per-phase device times of one inversion: K1 (pivot search + min-loc) vs fused step kernels.
usage: python tools/phase_times.py n m [n m ...]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from mpi_jordan_crazy_acceleration_b200.models.jordan import JordanSolver, _SingleComm  # noqa: E402


def main():
    args = [int(x) for x in sys.argv[1:]] or [8192, 128]
    for n, m in zip(args[0::2], args[1::2]):
        s = JordanSolver(n, m, comm=_SingleComm())
        s.set_profiling(True)
        for _ in range(2):
            s.fill_formula(1)
            sec = s.invert()
        pt = s.phase_times()
        tf = 2.0 * n ** 3 / sec / 1e12
        Nr = s.layout.Nr
        per = {k.replace("_ms", "_us_per_step"): v * 1e3 / Nr for k, v in pt.items() if k.endswith("_ms")}
        print(json.dumps({"n": n, "m": m, "seconds": sec, "tflops": tf, **per, **pt}), flush=True)
        s.close()
        torch.cuda.synchronize()


if __name__ == "__main__":
    main()

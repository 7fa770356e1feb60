"""Measure the FP64 roofline denominators on this GPU: cuBLAS DGEMM (torch.matmul fp64) and a
device-to-device copy.  Writes gpurun_out/fp64_peaks.json."""
import json
import os
import sys

import torch


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    dev = torch.device("cuda:0")
    a = torch.randn(n, n, dtype=torch.float64, device=dev)
    b = torch.randn(n, n, dtype=torch.float64, device=dev)
    c = torch.empty_like(a)
    for _ in range(2):
        torch.matmul(a, b, out=c)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b, out=c)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    dgemm_tflops = 2.0 * n ** 3 / (best * 1e-3) / 1e12
    # rank-128 update shape (the engine's hot GEMM): C(n x n) -= L(n x 128) R(128 x n)
    L = torch.randn(n, 128, dtype=torch.float64, device=dev)
    R = torch.randn(128, n, dtype=torch.float64, device=dev)
    for _ in range(2):
        c.addmm_(L, R, alpha=-1.0)
    torch.cuda.synchronize()
    best2 = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        c.addmm_(L, R, alpha=-1.0)
        e1.record()
        torch.cuda.synchronize()
        best2 = min(best2, e0.elapsed_time(e1))
    rank_tflops = 2.0 * n * n * 128 / (best2 * 1e-3) / 1e12
    x = torch.empty(1 << 28, dtype=torch.float64, device=dev)
    y = torch.empty_like(x)
    y.copy_(x)
    torch.cuda.synchronize()
    bc = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        y.copy_(x)
        e1.record()
        torch.cuda.synchronize()
        bc = min(bc, e0.elapsed_time(e1))
    copy_gbs = 2 * x.numel() * 8 / (bc * 1e-3) / 1e9
    out = {"gpu": torch.cuda.get_device_name(0), "dgemm_n": n, "cublas_dgemm_tflops": dgemm_tflops,
           "cublas_rank128_update_tflops": rank_tflops, "copy_gbs": copy_gbs}
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/fp64_peaks.json", "w") as f:
        json.dump(out, f)
    print(json.dumps(out))


if __name__ == "__main__":
    main()

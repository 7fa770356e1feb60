"""The error-free int8 split behind the tcgen05 update path, pinned down on the CPU."""
import torch

from mpi_jordan_crazy_acceleration_b200.ops.ozaki_model import slice_rows, split_matmul


def test_digits_are_int8_and_reconstruct_exactly():
    torch.manual_seed(0)
    a = torch.randn(64, 128, dtype=torch.float64) * torch.exp(4 * torch.randn(64, 1, dtype=torch.float64))
    q, scale = slice_rows(a, 8)
    assert int(q.abs().max()) <= 64                       # fits signed int8 with head-room
    rec = sum(q[s].to(torch.float64) * 2.0 ** (-7 * s) for s in range(8)) * scale[:, None]
    # 8 slices keep 6 + 7*7 = 55 bits below each row maximum
    assert ((rec - a).abs() / a.abs().amax(dim=1, keepdim=True)).max() < 2.0 ** -54


def test_split_matmul_is_fp64_accurate_and_fits_int32():
    torch.manual_seed(1)
    L = torch.randn(96, 128, dtype=torch.float64) * torch.exp(3 * torch.randn(96, 1, dtype=torch.float64))
    R = torch.randn(128, 80, dtype=torch.float64)
    ref = L @ R
    out, acc_max = split_matmul(L, R, 8)
    assert acc_max < 2 ** 23                               # (d+1) * K * 64 * 64 <= 8 * 128 * 4096 = 2^22
    assert acc_max * 128 + acc_max < 2 ** 31               # adjacent diagonals can be merged in int32
    bound = (L.abs().amax(1, keepdim=True) * R.abs().amax(0, keepdim=True)) * 128
    assert ((out - ref).abs() / bound).max() < 2.0 ** -50
    assert ((out - ref).abs() / (L.abs() @ R.abs())).max() < 1e-12


def test_accuracy_degrades_with_fewer_slices_as_measured():
    torch.manual_seed(2)
    L = torch.randn(64, 128, dtype=torch.float64)
    R = torch.randn(128, 64, dtype=torch.float64)
    ref = L @ R
    errs = [float(((split_matmul(L, R, s)[0] - ref).abs() / (L.abs() @ R.abs())).max()) for s in (4, 6, 7, 8)]
    assert errs[0] > errs[1] > errs[2] >= errs[3]
    assert errs[0] < 1e-6 and errs[1] < 1e-10 and errs[3] < 1e-14

"""Allocation-free batched block inversion (``jb_block_inverse_noalloc``), driven through the C ABI.

Some users of this library cannot let it call cudaMalloc: their process has already handed the
whole device to a memory pool, or they sit in a loop where cudaMalloc/cudaFree's device-wide
synchronisation is unaffordable.  The entry point tested here does the same work as
``jb_block_inverse`` but stages through a fixed device array inside the library, so its answers must
match the allocating entry point - also on a second call, which finds the first call's leftovers in
that shared array.  Collected last on purpose (file name), so a ``-x`` run has exercised
everything else before it gets here.
"""
import ctypes as C

import pytest
import torch

from mpi_jordan_crazy_acceleration_b200 import _native

pytestmark = pytest.mark.gpu


def _lib():
    lib = _native.cuda()
    sig = [C.c_void_p, C.c_int64, C.c_int32, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.jb_block_inverse_noalloc.restype = C.c_int
    lib.jb_block_inverse_noalloc.argtypes = sig
    lib.jb_block_inverse.restype = C.c_int
    lib.jb_block_inverse.argtypes = sig
    return lib


def _call(fn, blocks, norm_a):
    k, m, _ = blocks.shape
    inv = torch.full_like(blocks, float("nan"))
    norms = torch.zeros(k, dtype=torch.float64, device=blocks.device)
    ok = torch.full((k,), -1, dtype=torch.int32, device=blocks.device)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = fn(C.c_void_p(blocks.data_ptr()), k, m, C.c_double(norm_a), C.c_void_p(inv.data_ptr()),
            C.c_void_p(norms.data_ptr()), C.c_void_p(ok.data_ptr()), st)
    torch.cuda.synchronize()
    return rc, inv, norms, ok


def _blocks(k, m, seed):
    g = torch.Generator().manual_seed(seed)
    b = torch.randn(k, m, m, dtype=torch.float64, generator=g)
    if k > 2:
        b[1, 0] = b[1, 1]      # exactly singular block: its output must come back as zeros
    return b.cuda()


def test_noalloc_matches_allocating_entry_point_twice_in_a_row():
    lib = _lib()
    # second call is smaller and uses another block size: it must not see the first call's leftovers
    for k, m, seed in [(12, 100, 1), (5, 37, 2)]:
        blocks = _blocks(k, m, seed)
        norm_a = float(blocks.abs().sum(dim=2).max())
        rc0, inv0, n0, ok0 = _call(lib.jb_block_inverse, blocks, norm_a)
        rc1, inv1, n1, ok1 = _call(lib.jb_block_inverse_noalloc, blocks, norm_a)
        assert rc0 == 0 and rc1 == 0, lib.jb_last_error()
        assert torch.equal(ok0, ok1) and int(ok1[0]) == 1 and int(ok1[1]) == 0
        good = ok0.bool()
        scale = inv0[good].abs().max()
        assert (inv1[good] - inv0[good]).abs().max() <= 1e-9 * scale
        assert (n1[good] - n0[good]).abs().max() <= 1e-9 * n0[good].abs().max()
        assert torch.equal(inv1[~good], torch.zeros_like(inv1[~good]))
        ref = torch.linalg.inv(blocks[0].cpu())
        assert (inv1[0].cpu() - ref).abs().max() <= 1e-8 * ref.abs().max()


def test_noalloc_rejects_oversized_requests():
    lib = _lib()
    blocks = torch.zeros(40, 128, 128, dtype=torch.float64, device="cuda")   # 40*128*128 > 524288
    rc, *_ = _call(lib.jb_block_inverse_noalloc, blocks, 1.0)
    assert rc == _native.JB_ERR_ARG
    assert b"static scratch" in lib.jb_last_error()
    big = torch.zeros(1, 272, 272, dtype=torch.float64, device="cuda")        # m > 256
    rc, *_ = _call(lib.jb_block_inverse_noalloc, big, 1.0)
    assert rc == _native.JB_ERR_ARG

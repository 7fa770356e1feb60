"""Index math vs the reference's formulas restated (main.cpp:95-127, 521-532)."""
import ctypes as C

from hypothesis import given, settings, strategies as st

from mpi_jordan_crazy_acceleration_b200 import _native
from mpi_jordan_crazy_acceleration_b200.parallel import layout as L


def ref_rows_p_process(Nr, p, k):  # main.cpp:101-116 restated
    sender = Nr % p - 1
    if sender < 0:
        sender += p
    if Nr % p == 0:
        return Nr // p
    return Nr // p + 1 if k <= sender else Nr // p


@settings(max_examples=300, deadline=None)
@given(st.integers(1, 5000), st.integers(1, 300), st.integers(1, 16))
def test_rows_per_process_matches_reference(n, m, p):
    Nr = L.num_block_rows(n, m)
    assert Nr == (n // m + 1 if n % m else n // m)  # main.cpp:126
    lib = _native.host()
    total = 0
    for k in range(p):
        r = L.rows_per_process(Nr, p, k)
        assert r == ref_rows_p_process(Nr, p, k)
        assert r == lib.jb_rows_per_process(Nr, p, k)
        total += r
    assert total == Nr
    assert L.find_sender(Nr, p) == (Nr - 1) % p == lib.jb_find_sender(Nr, p)


@settings(max_examples=200, deadline=None)
@given(st.integers(1, 2000), st.integers(1, 64), st.integers(1, 8))
def test_local_global_roundtrip(n, m, p):
    lib = _native.host()
    seen = set()
    for k in range(p):
        lay = L.Layout(n, m, p, k)
        rows = lay.global_rows()
        assert len(rows) == lay.rows_real
        for i, gi in enumerate(rows):
            assert gi == ((i // m) * p + k) * m + i % m  # main.cpp:121
            assert gi == lib.jb_local_to_global_row(i, m, p, k)
            assert L.owner_of_row(gi, m, p) == k == lib.jb_owner_of_row(gi, m, p)
            assert L.global_to_local(gi, m, p) == i == lib.jb_global_to_local_row(gi, m, p)
            assert gi < n
            seen.add(gi)
    assert seen == set(range(n))


def test_padded_ld_and_64bit():
    lib = _native.host()
    assert L.padded_ld(1200, 60) == 1200 and lib.jb_padded_ld(1200, 60) == 1200
    assert L.padded_ld(100, 7) == 112
    # n = 65536: the reference's int len = RpP*m*n overflows (main.cpp:366); we are 64-bit
    n, m = 65536, 256
    assert lib.jb_rows_per_process(lib.jb_num_block_rows(n, m), 1, 0) * m * n == 2 ** 32
    assert lib.jb_local_to_global_row(2 ** 33, 256, 3, 1) == ((2 ** 33 // 256) * 3 + 1) * 256


def test_p_greater_than_Nr():
    lay = [L.Layout(30, 8, 8, k) for k in range(8)]
    assert [x.rloc for x in lay] == [1, 1, 1, 1, 0, 0, 0, 0]
    assert lay[3].rows_real == 6 and lay[3].rows_padded == 8

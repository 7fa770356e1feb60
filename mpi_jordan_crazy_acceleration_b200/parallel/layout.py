"""1-D block-cyclic row distribution (the reference's only parallelism strategy).

Block row ``I`` lives on rank ``I % p`` in local slot ``I // p`` (reference ``main.cpp:118-123``);
``Nr = ceil(n/m)`` block rows (``main.cpp:124-127``); the last one may be ragged
(``l_h = n - (Nr-1) m``, ``main.cpp:958``).  All arithmetic is 64-bit.
The same closed forms exist natively in ``csrc/host/layout.cpp``; this module is the Python view
used by the I/O, CLI and test layers.
"""
from __future__ import annotations

from dataclasses import dataclass


def num_block_rows(n: int, m: int) -> int:
    return (n + m - 1) // m


def find_sender(Nr: int, p: int) -> int:
    """Rank owning the last block row (``main.cpp:526-532``)."""
    return (Nr % p - 1) % p


def rows_per_process(Nr: int, p: int, k: int) -> int:
    """Block rows owned by rank ``k`` (``main.cpp:101-116``)."""
    if k < 0 or k >= p:
        return 0
    return Nr // p + (1 if k < Nr % p else 0)


def local_to_global(i: int, m: int, p: int, k: int) -> int:
    return ((i // m) * p + k) * m + i % m


def global_to_local(gi: int, m: int, p: int) -> int:
    return ((gi // m) // p) * m + gi % m


def owner_of_row(gi: int, m: int, p: int) -> int:
    return (gi // m) % p


def padded_ld(n: int, m: int) -> int:
    w = num_block_rows(n, m) * m
    return (w + 15) // 16 * 16


@dataclass(frozen=True)
class Layout:
    n: int
    m: int
    p: int
    rank: int

    @property
    def Nr(self) -> int:
        return num_block_rows(self.n, self.m)

    @property
    def rloc(self) -> int:
        return rows_per_process(self.Nr, self.p, self.rank)

    @property
    def rows_padded(self) -> int:
        """Local rows including identity padding of a ragged last block."""
        return self.rloc * self.m

    @property
    def last_height(self) -> int:
        return self.n - (self.Nr - 1) * self.m

    @property
    def rows_real(self) -> int:
        """Local rows that belong to the real ``n x n`` matrix."""
        r = self.rows_padded
        if self.rloc and (self.Nr - 1) % self.p == self.rank:
            r -= self.m - self.last_height
        return r

    @property
    def ldw(self) -> int:
        return padded_ld(self.n, self.m)

    def global_rows(self) -> list[int]:
        return [local_to_global(i, self.m, self.p, self.rank) for i in range(self.rows_real)]

    def block_rows(self) -> list[int]:
        return [s * self.p + self.rank for s in range(self.rloc)]
